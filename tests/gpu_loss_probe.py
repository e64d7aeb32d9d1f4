"""GPU diagnostic (not a pytest): a few calls of the fused photometric loss (for ncu launch lists)."""
import torch
import helpers  # noqa: F401
from gsraster_b200 import losses
H, W = 1024, 1280
gt = torch.rand(3, H, W, device="cuda")
img = (gt + 0.1 * torch.randn(3, H, W, device="cuda")).clamp(0, 1)
pl = losses.PhotometricLoss(H, W, 0.2)
for _ in range(4):
    out3, grad = pl(img, gt)
torch.cuda.synchronize()
print(out3.tolist())
