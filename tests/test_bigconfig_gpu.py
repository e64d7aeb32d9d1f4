"""GPU parity tests at the full sizes of BASELINE.json's configs 3 and 4 and on the bench's named workloads, against the
unmodified reference CUDA build run live (oracle/_ref). Everything is compared ON THE DEVICE (no multi-GB host
copies); a one-line summary per case is printed (kept under profiles/ from the round's GPU runs)."""
import pytest

import helpers as h
from helpers import scenes

pytestmark = pytest.mark.gpu


def _compare_on_device(name, sc, cam, grads):
    import torch
    P, W, H = sc.P, cam.W, cam.H
    mine = h.run_mine(sc, cam, grads)
    ref = h.run_ref(sc, cam, grads)
    R = mine["R"]
    assert R == ref["R"] and R > 0
    for k in ("radii", "is_used", "tiles_touched"):
        assert torch.equal(mine[k].view(-1), ref[k].view(-1)), k
    for k in ("n_contrib", "ranges", "point_list", "sorted_keys"):   # run_ref decodes the reference's scratch buffers
        assert torch.equal(mine[k].view(-1), ref[k].view(-1)), k
    assert torch.equal(mine["final_T"].view(torch.int32), ref["final_T"].view(torch.int32)), "final_T bits"
    img_err = {k: float((mine[k] - ref[k]).abs().max()) for k in ("color", "depth", "objects")}
    assert max(img_err.values()) <= 1e-4, img_err
    ref2 = h.run_ref(sc, cam, grads)
    worst = {}
    for k in h.GRAD_KEYS:
        a, b1, b2 = mine[k].double().reshape(-1), ref[k].double().reshape(-1), ref2[k].double().reshape(-1)
        if a.numel() == 0:
            continue
        mx = float(b1.abs().max())
        if mx == 0:
            assert float(a.abs().max()) == 0
            continue
        spread_max = float((b1 - b2).abs().max()) / mx
        spread_l2 = float((b1 - b2).norm() / b1.norm())
        e_max = float((a - b1).abs().max()) / mx
        e_l2 = float((a - b1).norm() / b1.norm())
        assert e_max <= 1e-3 + 3 * spread_max, (k, "max-norm", e_max, spread_max)
        assert e_l2 <= 1e-3 + 3 * spread_l2, (k, "rel-L2", e_l2, spread_l2)
        refm = 0.5 * (b1 + b2)
        sel = refm.abs() >= 1e-4 * mx
        noise = (b1 - b2).abs()   # the reference's own run-to-run spread (helpers.grad_elementwise_violations)
        viol = int((((a - refm).abs() > 1e-2 * refm.abs() + 2.0 * noise + noise.max() + 1e-6 * mx) & sel).sum())
        assert viol == 0, (k, "per-element", viol, int(sel.sum()))
        nz_a = a.view(P, -1).abs().sum(1) > 0
        nz_b = (b1.view(P, -1).abs().sum(1) > 0) | (b2.view(P, -1).abs().sum(1) > 0)
        assert torch.equal(nz_a, nz_b), (k, "non-zero rows")
        worst[k] = (e_max, e_l2)
    wk = max(worst, key=lambda k: worst[k][0])
    print("bigconfig %s: P=%d %dx%d F=%d R=%d: integer state bit-exact, images max-abs %.1e, gradients worst max-norm "
          "%.1e / rel-L2 %.1e (%s), per-element and non-zero-row checks clean"
          % (name, P, W, H, sc.sh_objs.shape[-1], R, max(img_err.values()), worst[wk][0], worst[wk][1], wk))


@pytest.mark.skipif(not h.have_ref(32), reason="oracle/_ref F=32 build not shipped")
def test_config4_6M_1920x1080_f32():
    """BASELINE config 4: 6M Gaussians, 1920x1080, 32-d feature (R ~ 4.5e8 tile instances)."""
    W, H, P, F = 1920, 1080, 6_000_000, 32
    sc = scenes.make_scene(P, sh_degree=3, F=F, seed=1238, W=W, H=H)
    _compare_on_device("c4", sc, scenes.make_camera(W, H, 0), scenes.make_upstream_grads(W, H, F))


@pytest.mark.skipif(not h.have_ref(), reason="oracle/_ref not shipped")
def test_config3_2M_keyframe():
    """BASELINE config 3: one keyframe of the 8-view batch over 2M Gaussians (trajectory scene, distinct pose)."""
    W, H, P, F = 1280, 1024, 2_000_000, 16
    sc = scenes.make_trajectory_scene(P, 8, sh_degree=3, F=F, seed=1234, W=W, H=H)
    _compare_on_device("c3", sc, scenes.make_camera(W, H, 5), scenes.make_upstream_grads(W, H, F))


@pytest.mark.skipif(not h.have_ref(), reason="oracle/_ref not shipped")
@pytest.mark.parametrize("name", ["trained_like_x3", "non_saturating", "survey8d_view0"])
def test_bench_workloads_vs_reference(name):
    """The named single-view workloads of bench.py's `workloads` block, at their full size."""
    W, H, F = 1280, 1024, 16
    sc = scenes.make_workload_scene(name, 3, F, W, H)
    _compare_on_device(name, sc, scenes.make_camera(W, H, 0), scenes.make_upstream_grads(W, H, F))


@pytest.mark.skipif(not h.have_ref(), reason="oracle/_ref not shipped")
def test_headline_trajectory_keyframe_vs_reference():
    """A keyframe of the bench's headline trajectory scene (16 poses, 1M Gaussians)."""
    W, H, F = 1280, 1024, 16
    sc = scenes.make_trajectory_scene(1_000_000, 16, sh_degree=3, F=F, seed=1234, W=W, H=H)
    _compare_on_device("headline_traj_pose9", sc, scenes.make_camera(W, H, 9), scenes.make_upstream_grads(W, H, F))
