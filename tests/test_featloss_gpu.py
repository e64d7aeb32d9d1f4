"""GPU parity of the fused segmentation-feature loss (csrc/featloss.cu through the C ABI): against the golden vectors of
the reference's call sequence, the numpy oracle, and torch's own Conv2d + CrossEntropyLoss at the headline image size.
Tolerances: loss 1e-5 absolute (values ~1), gradient 1e-4 of its max (3xTF32 products, fp32 accumulation)."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


class _Cls:
    def __init__(self, w, b):
        self.weight, self.bias = w, b


def _run(objects, gt, w, b, **kw):
    from gsraster_b200.losses import FeatureLoss
    fl = FeatureLoss(_Cls(w, b))
    return fl(objects, gt, need_pred=True, **kw)


def test_feature_loss_matches_golden_vectors():
    files = sorted(glob.glob(os.path.join(GOLD, "featloss_*.npz")))
    assert len(files) >= 3
    for f in files:
        z = np.load(f)
        dev = "cuda"
        out2, grad, pred = _run(torch.from_numpy(z["objects"]).to(dev), torch.from_numpy(z["gt"]).to(dev),
                                torch.from_numpy(z["weight"]).to(dev), torch.from_numpy(z["bias"]).to(dev))
        o = out2.cpu().numpy()
        assert abs(o[0] - z["loss"]) <= 1e-5 and abs(o[1] - z["mean_ce"]) <= 5e-5, (f, o, z["loss"])
        assert np.abs(grad.cpu().numpy() - z["grad"]).max() <= 1e-4 * np.abs(z["grad"]).max(), f
        assert (pred.cpu().numpy() == z["pred"]).mean() >= 0.998, f


@pytest.mark.parametrize("F,C,H,W", [(16, 100, 1024, 1280), (32, 100, 123, 77), (16, 256, 64, 50), (16, 128, 31, 33),
                                     (32, 200, 40, 40), (16, 2, 17, 19)])
def test_feature_loss_matches_torch_modules(F, C, H, W):
    torch.manual_seed(F + C)
    dev = "cuda"
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        classifier = torch.nn.Conv2d(F, C, kernel_size=1).to(dev)
        crit = torch.nn.CrossEntropyLoss(reduction='none')
        objects = (1.5 * torch.randn(F, H, W, device=dev)).requires_grad_(True)
        gt = torch.randint(0, C, (H, W), device=dev)
        gt[torch.rand(H, W, device=dev) < 0.05] = -100
        logits = classifier(objects)
        ref_pred = torch.argmax(logits, dim=0)
        ref = crit(logits.unsqueeze(0), gt.unsqueeze(0)).squeeze().mean() / torch.log(torch.tensor(float(C)))
        (0.7 * ref).backward()
    finally:
        torch.backends.cudnn.allow_tf32 = old
    out2, grad, pred = _run(objects.detach(), gt, classifier.weight.detach(), classifier.bias.detach(), grad_weight=0.7)
    assert abs(float(out2[0]) - float(ref)) <= 1e-5 * max(1.0, abs(float(ref)))
    gmax = float(objects.grad.abs().max())
    assert float((grad - objects.grad).abs().max()) <= 1e-4 * gmax
    assert float((pred.long() == ref_pred).float().mean()) >= 0.999
    # value-only call leaves the result unchanged and is deterministic
    out2b, g2, _ = _run(objects.detach(), gt, classifier.weight.detach(), classifier.bias.detach(), need_grad=False)
    assert g2 is None and float(out2b[0]) == float(out2[0])


def test_feature_loss_rejects_bad_arguments():
    dev = "cuda"
    w, b = torch.zeros(100, 16, device=dev), torch.zeros(100, device=dev)
    with pytest.raises(RuntimeError):
        _run(torch.zeros(16, 8, 8), torch.zeros(8, 8, dtype=torch.int64, device=dev), w, b)
    with pytest.raises(RuntimeError):
        _run(torch.zeros(16, 8, 8, device=dev), torch.zeros(8, 8, dtype=torch.int32, device=dev), w, b)
    with pytest.raises(RuntimeError):
        _run(torch.zeros(8, 8, 8, device=dev), torch.zeros(8, 8, dtype=torch.int64, device=dev),
             torch.zeros(100, 8, device=dev), b)


@pytest.mark.parametrize("F,C,H,W", [(16, 100, 480, 640), (32, 100, 123, 77), (16, 112, 64, 65), (32, 7, 33, 129),
                                     (16, 19, 1, 1), (16, 100, 1, 127)])
def test_tcgen05_kernel_matches_mma_sync_kernel(F, C, H, W):
    """featloss_tc.cu (tcgen05.mma kind::tf32, accumulators and the d operand in TMEM; default for C <= 112) against
    featloss.cu (mma.sync; GSR_FEATLOSS_MMA_SYNC=1) on the same inputs: same 3xTF32 split, so the two agree far inside
    the reference tolerance; ragged tiles (H*W not a multiple of 128), ignored labels, one-pixel images."""
    torch.manual_seed(7 * F + C)
    dev = "cuda"
    objects = 1.5 * torch.randn(F, H, W, device=dev)
    gt = torch.randint(0, C, (H, W), device=dev)
    gt[torch.rand(H, W, device=dev) < 0.1] = -100
    w, b = 0.3 * torch.randn(C, F, device=dev), 0.5 * torch.randn(C, device=dev)
    a = _run(objects, gt, w, b, grad_weight=1.3)
    os.environ["GSR_FEATLOSS_MMA_SYNC"] = "1"
    try:
        m = _run(objects, gt, w, b, grad_weight=1.3)
    finally:
        del os.environ["GSR_FEATLOSS_MMA_SYNC"]
    assert abs(float(a[0][0]) - float(m[0][0])) <= 2e-6 * max(1.0, abs(float(m[0][0])))
    assert abs(float(a[0][1]) - float(m[0][1])) <= 1e-5 * max(1.0, abs(float(m[0][1])))
    gmax = float(m[1].abs().max())
    assert float((a[1] - m[1]).abs().max()) <= 2e-5 * gmax + 1e-12
    assert float((a[2] == m[2]).float().mean()) >= 0.999
    # deterministic: a second call returns the same bits
    a2 = _run(objects, gt, w, b, grad_weight=1.3)
    assert torch.equal(a[1], a2[1]) and float(a[0][0]) == float(a2[0][0])
