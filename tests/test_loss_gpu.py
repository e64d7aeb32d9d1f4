"""fused photometric loss (csrc/loss.cu) against (a) golden vectors produced by the reference's own
utils/loss_utils.py (tests/golden/make_golden_loss.py) and (b) a torch restatement of it at full size."""
import glob
import os

import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (puts the package on sys.path)
from gsraster_b200 import losses

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


from oracle.pyloss import photometric_loss as _torch_loss  # noqa: E402


def test_matches_golden_vectors_from_the_reference_loss_code():
    files = sorted(glob.glob(os.path.join(GOLD, "loss_*.npz")))
    assert len(files) >= 3
    for f in files:
        z = np.load(f)
        img, gt = torch.from_numpy(z["img"]).cuda(), torch.from_numpy(z["gt"]).cuda()
        H, W = img.shape[1:]
        out3, grad = losses.PhotometricLoss(H, W, float(z["lam"]))(img, gt)
        out3 = out3.cpu().numpy()
        assert abs(out3[0] - z["loss"]) <= 2e-6 and abs(out3[1] - z["Ll1"]) <= 2e-6 and abs(out3[2] - z["Lssim"]) <= 2e-6, f
        err = np.abs(grad.cpu().numpy() - z["grad"]).max() / np.abs(z["grad"]).max()
        assert err <= 2e-5, (f, err)


@pytest.mark.parametrize("H,W,lam", [(1024, 1280, 0.2), (1080, 1920, 0.2), (37, 501, 0.8), (1024, 1280, 0.0), (5, 7, 0.0)])
def test_matches_torch_restatement_at_full_size(H, W, lam):
    g = torch.Generator(device="cuda").manual_seed(H)
    gt = torch.rand(3, H, W, generator=g, device="cuda")
    gt[torch.rand(3, H, W, generator=g, device="cuda") < 0.05] = 0.0
    img = (gt + 0.1 * torch.randn(3, H, W, generator=g, device="cuda")).clamp(0, 1).requires_grad_(True)
    loss, l1, ss = _torch_loss(img, gt, lam)
    loss.backward()
    loss, l1, ss = loss.detach(), l1.detach(), ss.detach()
    out3, grad = losses.PhotometricLoss(H, W, lam)(img.detach(), gt)
    assert abs(float(out3[0]) - float(loss)) <= 5e-6 and abs(float(out3[1]) - float(l1)) <= 5e-6
    assert abs(float(out3[2]) - float(ss)) <= 5e-6 or lam == 0.0   # lambda = 0: the SSIM term is skipped, out3[2] = 0
    err = float((grad - img.grad).abs().max() / img.grad.abs().max())
    assert err <= 5e-5, err
    # value-only call, and run-to-run determinism of the value
    out3b, none = losses.PhotometricLoss(H, W, lam)(img.detach(), gt, need_grad=False)
    assert none is None and torch.equal(out3, out3b)
