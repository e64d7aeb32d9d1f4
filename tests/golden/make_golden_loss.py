"""Generates tests/golden/loss_*.npz by importing the REFERENCE's own loss code
(/root/reference/src/gaussian_splatting/gaussian_splatting/utils/loss_utils.py, pure torch, CPU) -- run in the build
container, where /root/reference exists. loss_rgb = (1 - lambda) * Ll1 + lambda * (1 - L_ssim) as composed at
scripts/gaussian_splatting.py:805-807; the gradient comes from torch autograd through the reference functions."""
import importlib.util
import os
import sys

import numpy as np
import torch

REF = "/root/reference/src/gaussian_splatting/gaussian_splatting/utils/loss_utils.py"
spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)
here = os.path.dirname(os.path.abspath(__file__))
for name, (H, W, lam, seed, zero_frac) in dict(a=(40, 56, 0.2, 1, 0.1), b=(33, 47, 0.5, 2, 0.0), c=(16, 16, 0.2, 3, 0.5)).items():
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(3, H, W, generator=g)
    gt[torch.rand(3, H, W, generator=g) < zero_frac] = 0.0          # the masks of loss_utils.py:19,37
    img = (gt + 0.2 * torch.randn(3, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
    _, Ll1 = ref.l1_loss(img, gt)
    _, Ls = ref.ssim(img, gt)
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - Ls)
    loss.backward()
    np.savez_compressed(os.path.join(here, f"loss_{name}.npz"), img=img.detach().numpy(), gt=gt.numpy(), lam=np.float32(lam),
                        loss=np.float32(loss.item()), Ll1=np.float32(Ll1.item()), Lssim=np.float32(Ls.item()),
                        grad=img.grad.numpy())
    print(name, H, W, float(loss), float(Ll1), float(Ls))
