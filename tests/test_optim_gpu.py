"""fused activation + Adam (csrc/optim.cu) against the reference's own chain --
torch activations (scene/gaussian_model.py:109-133), autograd, torch.optim.Adam(eps=1e-15) (:247-257)."""
import pytest
import torch

import helpers  # noqa: F401  (puts the package on sys.path)
from gsraster_b200 import optim

pytestmark = pytest.mark.gpu


def _raw_params(P, M, F, seed, device):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s, k=1.0: (torch.randn(*s, generator=g) * k).to(device)
    return dict(xyz=r(P, 3, k=3.0), f_dc=r(P, 1, 3), f_rest=r(P, M - 1, 3, k=0.2), opacity=r(P, 1, k=2.0),
                scaling=r(P, 3) - 3.0, rotation=r(P, 4), objects=r(P, 1, F))


def _rand_grads(P, M, F, seed, device, rows=None):
    g = torch.Generator().manual_seed(seed)
    n = P if rows is None else rows
    r = lambda *s: (torch.randn(*s, generator=g) * 1e-3).to(device)
    return dict(means3D=r(n, 3), sh=r(n, M, 3), objects=r(n, F), opacity=r(n, 1), scales=r(n, 3), rotations=r(n, 4))


def _torch_reference_steps(raw, grads_per_step, lr):
    """The reference chain: activations -> (stand-in for the rasterizer: <activated, upstream grad>) -> backward -> Adam."""
    p = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    opt = torch.optim.Adam([{"params": [p[k]], "lr": lr[k]} for k in optim.GROUPS], lr=0.0, eps=1e-15)
    for gr in grads_per_step:
        opt.zero_grad()
        shs = torch.cat((p["f_dc"], p["f_rest"]), dim=1)
        act = dict(means3D=p["xyz"], sh=shs, objects=p["objects"].reshape(gr["objects"].shape),
                   opacity=torch.sigmoid(p["opacity"]), scales=torch.exp(p["scaling"]),
                   rotations=torch.nn.functional.normalize(p["rotation"]))
        loss = sum((act[k] * gr[k]).sum() for k in gr)   # d loss / d activated == the given gradients
        loss.backward()
        opt.step()
    return {k: v.detach() for k, v in p.items()}


def _close(a, b, tol=2e-5):
    scale = float(b.abs().max()) + 1e-12
    return float((a - b).abs().max()) / scale <= tol


@pytest.mark.parametrize("M,F", [(16, 16), (1, 16), (4, 32)])
def test_dense_fused_adam_matches_torch_adam(M, F):
    P, dev = 3001, "cuda"
    raw = _raw_params(P, M, F, 1, dev)
    lr = dict(optim.DEFAULT_LR, xyz=1e-3)
    steps = [_rand_grads(P, M, F, 10 + i, dev) for i in range(4)]
    steps[2]["opacity"][::2] = 0.0   # zero rows still move with the momentum (dense Adam semantics)
    ref = _torch_reference_steps(raw, steps, lr)
    mine = {k: v.clone() for k, v in raw.items()}
    fa = optim.FusedGaussianAdam(mine, lr=lr)
    # activation of the initial state
    assert _close(fa.scales, torch.exp(raw["scaling"])) and _close(fa.opacities, torch.sigmoid(raw["opacity"]))
    assert _close(fa.rotations, torch.nn.functional.normalize(raw["rotation"]))
    assert torch.equal(fa.shs, torch.cat((raw["f_dc"], raw["f_rest"]), dim=1))
    for gr in steps:
        fa.step(gr)
    for k in optim.GROUPS:
        if mine[k].numel():
            assert _close(mine[k], ref[k]), k
    # the re-activated tensors belong to the UPDATED raw parameters
    assert _close(fa.scales, torch.exp(ref["scaling"])) and _close(fa.opacities, torch.sigmoid(ref["opacity"]))
    assert _close(fa.rotations, torch.nn.functional.normalize(ref["rotation"]))
    assert _close(fa.shs, torch.cat((ref["f_dc"], ref["f_rest"]), dim=1))


def test_sparse_fused_adam_steps_only_the_listed_rows():
    P, M, F, dev = 5000, 16, 16, "cuda"
    raw = _raw_params(P, M, F, 2, dev)
    lr = dict(optim.DEFAULT_LR, xyz=1e-3)
    idx = torch.sort(torch.randperm(P, generator=torch.Generator().manual_seed(3))[:700])[0].to(dev)
    cap, count = 1024, torch.tensor([700], dtype=torch.int32, device=dev)
    idx_buf = torch.zeros(cap, dtype=torch.int32, device=dev)
    idx_buf[:700] = idx.to(torch.int32)
    steps = [_rand_grads(P, M, F, 20 + i, dev, rows=cap) for i in range(3)]
    # reference: torch Adam on the sub-model made of the listed rows
    sub = {k: v[idx].clone() for k, v in raw.items()}
    ref = _torch_reference_steps(sub, [{k: v[:700] for k, v in gr.items()} for gr in steps], lr)
    mine = {k: v.clone() for k, v in raw.items()}
    fa = optim.FusedGaussianAdam(mine, lr=lr)
    for gr in steps:
        fa.step(gr, idx=idx_buf, count=count)
    rest = torch.ones(P, dtype=torch.bool, device=dev)
    rest[idx] = False
    for k in optim.GROUPS:
        assert _close(mine[k][idx], ref[k]), k
        assert torch.equal(mine[k][rest], raw[k][rest]), k       # untouched rows: bit-identical
        assert float(fa.m[k][rest].abs().max()) == 0.0
    assert _close(fa.scales[idx], torch.exp(ref["scaling"]))


def test_cpu_tensors_are_rejected():
    raw = _raw_params(8, 4, 16, 4, "cpu")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        optim.FusedGaussianAdam(raw)
