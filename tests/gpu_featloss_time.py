"""GPU diagnostic (not a pytest): time of the fused feature loss at the headline image size (1280x1024, F = 16, C = 100),
tcgen05 kernel (default) vs the mma.sync kernel (GSR_FEATLOSS_MMA_SYNC=1). CUDA events, 20 calls after 5 warm-up."""
import os
import sys

import torch

import helpers  # noqa: F401
from gsraster_b200 import losses

F = int(sys.argv[1]) if len(sys.argv) > 1 else 16
H, W, C = 1024, 1280, 100
torch.manual_seed(3)
fl = losses.FeatureLoss(torch.nn.Conv2d(F, C, kernel_size=1).cuda())
obj = torch.randn(F, H, W, device="cuda")
gt = torch.randint(0, C, (H, W), device="cuda")


def timed(tag):
    for _ in range(5):
        out = fl(obj, gt, need_pred=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        out = fl(obj, gt, need_pred=True)
    e1.record()
    torch.cuda.synchronize()
    print("%s: %.4f ms per call, loss %.6f" % (tag, e0.elapsed_time(e1) / 20, float(out[0][0])), flush=True)
    return out


a = timed("tcgen05 (featloss_tc.cu)")
os.environ["GSR_FEATLOSS_MMA_SYNC"] = "1"
b = timed("mma.sync (featloss.cu)")
print("max |grad diff| / max |grad| = %.2e" % float((a[1] - b[1]).abs().max() / b[1].abs().max()))
