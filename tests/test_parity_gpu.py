"""GPU parity tests: the product (C ABI, include/gsraster_b200.h) against
 (a) golden fixtures produced by the unmodified reference CUDA build (tests/golden/make_golden.py),
 (b) the reference CUDA build run live (oracle/_ref, when shipped to the box),
 (c) the CPU oracle (oracle/oracle.cpp) on small scenes,
with the bars BASELINE.json states: bit-exact radii / tiles_touched / sorted keys / tile ranges /
n_contrib; max-abs 1e-4 on colour, depth, features; relative 1e-3 on gradients."""
import glob
import os

import numpy as np
import pytest

import helpers as h
from helpers import scenes

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXACT_KEYS = ("radii", "is_used", "tiles_touched", "point_list", "sorted_keys", "ranges", "n_contrib")
IMG_TOL = 1e-4   # max-abs, colour / depth / features  (north_star)
GRAD_TOL = 1e-3  # relative (tensor max-norm), gradients (north_star)


def _check_exact(mine, ref, keys=EXACT_KEYS):
    for k in keys:
        a, b = h.npy(mine[k]), h.npy(ref[k])
        assert a.shape == b.shape or a.size == b.size, (k, a.shape, b.shape)
        assert (a.reshape(-1).astype(np.int64) == b.reshape(-1).astype(np.int64)).all(), f"{k} not bit-exact"


def _check_images(mine, ref):
    for k in ("color", "depth", "objects"):
        err = float(np.abs(h.npy(mine[k]) - h.npy(ref[k])).max())
        assert err <= IMG_TOL, (k, err)


def _check_grads(mine, ref, tol=GRAD_TOL, ref_again=None):
    """ref_again: a second run of the reference on the same inputs. Its float atomics make the reference's
    own gradients differ run to run (measured on B200: up to 3.3e-4 on dL_dcov3D for screen-filling splats,
    tests/gpu_gradnoise.py; the product's spread is ~5e-6), so the bar is 1e-3 plus 3x that spread.
    Three bars per tensor: (i) max-norm max|a-b| / max|b| <= tol, (ii) relative L2 <= tol, both widened by 3x the
    reference's own spread, and -- when a second reference run is given -- (iii) per element on the entries with
    |ref| >= 1e-4 max|ref|: relative error <= 1e-2 against the mean of the two reference runs (noise-aware, see
    helpers.grad_elementwise_violations); plus (iv) the set of Gaussians with a non-zero gradient row is the
    reference's."""
    for k in h.GRAD_KEYS:
        a, b = h.npy(mine[k]), h.npy(ref[k])
        if a.size == 0 and b.size == 0:
            continue
        a = a.reshape(b.shape)
        err = h.grad_rel_err(a, b)
        t = tol + (3.0 * h.grad_rel_err(ref_again[k], ref[k]) if ref_again is not None else 0.0)
        assert err <= t, (k, "max-norm", err, t)
        l2 = h.grad_rel_l2(a, b)
        t2 = tol + (3.0 * h.grad_rel_l2(ref_again[k], ref[k]) if ref_again is not None else 0.0)
        assert l2 <= t2, (k, "rel-L2", l2, t2)
        if ref_again is not None:
            viol, n, worst = h.grad_elementwise_violations(a, b, ref_again[k])
            assert viol == 0, (k, "per-element", viol, n, worst)
            P = b.shape[0]
            assert (h.nonzero_rows(a, P) == (h.nonzero_rows(b, P) | h.nonzero_rows(ref_again[k], P))).all(), (k, "non-zero rows")


def _golden_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                  if not os.path.basename(p).startswith(("knn", "loss_", "featloss_", "cls3d_")))


@pytest.mark.parametrize("name", _golden_cases())
def test_product_vs_golden(name):
    from golden.make_golden import load_case
    sc, cam, grads, colors, cov, ref = load_case(os.path.join(GOLDEN, name + ".npz"))
    mine = h.run_mine(sc, cam, grads, colors_precomp=colors, cov3D_precomp=cov)
    assert mine["R"] == int(ref["R"][0])
    _check_exact(mine, ref)
    vis = ref["radii"] > 0
    for k in ("depths", "means2D", "conic_opacity", "rgb", "clamped"):  # per-Gaussian state, bit-exact too
        a, b = h.npy(mine[k])[vis], ref[k][vis]
        assert (a.view(np.uint8) == b.view(np.uint8)).all(), k
    assert (h.npy(mine["final_T"]).view(np.uint32) == ref["final_T"].view(np.uint32)).all()
    _check_images(mine, ref)
    _check_grads(mine, ref)


def test_knn_vs_golden():
    import torch
    from gsraster_b200 import capi
    for f in ("knn.npz", "knn_few.npz"):
        z = np.load(os.path.join(GOLDEN, f))
        out = h.npy(capi.distCUDA2(torch.from_numpy(z["in_points"]).cuda()))
        assert (out.view(np.uint32) == z["ref_mean_dists"].view(np.uint32)).all(), f


# ---- live reference (oracle/_ref) ----------------------------------------------------------------
needs_ref = pytest.mark.skipif(not h.have_ref(), reason="oracle/_ref not shipped")


@needs_ref
@pytest.mark.parametrize("P,W,H,deg,view,mul", [
    (10_000, 640, 512, 0, 0, 1.0),      # config C1
    (10_000, 640, 512, 3, 3, 3.0),
    (500_000, 1280, 1024, 3, 0, 1.0),   # config C2
    (200_000, 1280, 1024, 3, 2, 3.0),   # "trained-like": larger splats, much larger R
    (30_000, 1000, 700, 2, 1, 2.0),     # image size not a multiple of 16
])
def test_product_vs_reference_live(P, W, H, deg, view, mul):
    sc = scenes.make_scene(P, sh_degree=deg, F=16, seed=1234 + P % 97 + deg, W=W, H=H, scale_mul=mul)
    cam = scenes.make_camera(W, H, view)
    grads = scenes.make_upstream_grads(W, H, 16)
    mine = h.run_mine(sc, cam, grads)
    ref = h.run_ref(sc, cam, grads)
    assert mine["R"] == ref["R"] and ref["R"] > 0
    _check_exact(mine, ref)
    _check_images(mine, ref)
    _check_grads(mine, ref, ref_again=h.run_ref(sc, cam, grads))


@pytest.mark.skipif(not h.have_ref(32), reason="oracle/_ref F=32 build not shipped")
def test_product_vs_reference_f32_features():
    """Config C4's 32-d feature (reference rebuilt with NUM_OBJECTS=32)."""
    W, H, P = 960, 544, 150_000
    sc = scenes.make_scene(P, sh_degree=3, F=32, seed=4, W=W, H=H, scale_mul=2.0)
    cam = scenes.make_camera(W, H, 1)
    grads = scenes.make_upstream_grads(W, H, 32)
    mine = h.run_mine(sc, cam, grads)
    ref = h.run_ref(sc, cam, grads)
    assert mine["R"] == ref["R"]
    _check_exact(mine, ref)
    _check_images(mine, ref)
    _check_grads(mine, ref, ref_again=h.run_ref(sc, cam, grads))


@needs_ref
def test_full_size_headline_properties():
    """1M Gaussians, 1280x1024, F=16 (the headline shape): bit-exact integer state vs the live
    reference plus size-independent properties: sortedness of the keys, ranges partition [0,R),
    n_contrib <= range length, sum(tiles_touched) == R."""
    W, H, P = 1280, 1024, 1_000_000
    sc = scenes.make_scene(P, sh_degree=3, F=16, seed=1234, W=W, H=H)
    cam = scenes.make_camera(W, H, 0)
    grads = scenes.make_upstream_grads(W, H, 16)
    mine = h.run_mine(sc, cam, grads)
    keys = h.npy(mine["sorted_keys"]).astype(np.uint64)
    assert (np.diff(keys.view(np.int64)) >= 0).all()  # keys < 2^45, signed view is safe
    assert int(h.npy(mine["tiles_touched"]).astype(np.int64).sum()) == mine["R"]
    rng = h.npy(mine["ranges"]).astype(np.int64)
    ne = rng[rng[:, 1] > rng[:, 0]]
    assert ne[0, 0] == 0 and ne[-1, 1] == mine["R"] and (ne[1:, 0] == ne[:-1, 1]).all()
    T_x = (W + 15) // 16
    nc = h.npy(mine["n_contrib"]).astype(np.int64).reshape(H, W)
    tile_of = (np.arange(H)[:, None] // 16) * T_x + (np.arange(W)[None, :] // 16)
    assert (nc <= (rng[:, 1] - rng[:, 0])[tile_of]).all()
    ref = h.run_ref(sc, cam, grads)
    _check_exact(mine, ref)
    _check_images(mine, ref)
    _check_grads(mine, ref, ref_again=h.run_ref(sc, cam, grads))


@needs_ref
def test_mark_visible_vs_reference():
    import ctypes as C
    import torch
    from gsraster_b200 import capi
    sc = scenes.make_scene(100_003, sh_degree=0, seed=3)
    cam = scenes.make_camera(1280, 1024, 4)
    d = h.to_torch(sc, cam)
    mine = capi.mark_visible(d["means3D"], d["viewmatrix"], d["projmatrix"])
    ref = torch.zeros(sc.P, dtype=torch.bool, device="cuda")
    L = h.ref_lib()
    L.ref_mark_visible.argtypes = [C.c_int] + [C.c_void_p] * 4
    L.ref_mark_visible(sc.P, d["means3D"].data_ptr(), d["viewmatrix"].data_ptr(), d["projmatrix"].data_ptr(),
                       ref.data_ptr())
    torch.cuda.synchronize()
    assert (mine == ref).all()
    from oracle import pyoracle as po
    assert (h.npy(mine) == po.mark_visible(sc.means3D, cam.viewmatrix)).all()


@pytest.mark.skipif(not os.path.exists(h.REF_KNN_LIB), reason="oracle/_ref knn not shipped")
@pytest.mark.parametrize("P", [1, 2, 4, 255, 257, 100_000, 1_000_000])
def test_knn_vs_reference_live(P):
    import torch
    from gsraster_b200 import capi
    pts = scenes.make_points(P, seed=4321 + P)
    if P == 257:
        pts[10:20] = pts[0]  # duplicates count as distance 0 (simple_knn.cu:131-145)
    mine = h.npy(capi.distCUDA2(torch.from_numpy(pts).cuda()))
    ref = h.npy(h.run_ref_knn(pts))
    assert (mine.view(np.uint32) == ref.view(np.uint32)).all()


# ---- CPU oracle ----------------------------------------------------------------------------------
@pytest.mark.parametrize("P,W,H,deg,F", [(10_000, 640, 512, 0, 16), (6_000, 200, 150, 3, 16), (5_000, 256, 192, 2, 32)])
def test_product_vs_oracle(P, W, H, deg, F):
    """CPU restatement vs product: integer state may differ on isolated Gaussians where the host
    compiler and nvcc round differently (FMA contraction), so compare what is well conditioned."""
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=21 + F, W=W, H=H, scale_mul=2.0)
    cam = scenes.make_camera(W, H, 1)
    grads = scenes.make_upstream_grads(W, H, F)
    mine = h.run_mine(sc, cam, grads)
    orc = h.run_oracle(sc, cam, grads)
    mism = int((h.npy(mine["radii"]) != orc["radii"]).sum())
    assert mism <= max(1, P // 2000), mism
    if mism == 0 and mine["R"] == orc["R"]:
        # host libm exp vs MUFU.EX2 can flip the alpha >= 1/255 test on isolated pixels (a 0.004 * colour
        # jump), so the CPU comparison is on the 99.99 % quantile; the strict max-abs bar is held against
        # the reference CUDA build (golden + live tests above)
        for k in ("color", "depth", "objects"):
            dd = np.abs(h.npy(mine[k]) - orc[k])
            assert np.quantile(dd, 0.9999) <= IMG_TOL and dd.max() <= 2e-2, (k, dd.max())
        _check_grads(mine, orc, tol=5e-3)
    else:
        for k in ("color", "depth", "objects"):
            assert float(np.median(np.abs(h.npy(mine[k]) - orc[k]))) < 1e-6


# ---- edge cases ----------------------------------------------------------------------------------
def test_all_culled_and_tiny():
    import torch
    W, H = 64, 48
    sc = scenes.make_scene(50, sh_degree=1, seed=5, W=W, H=H, behind_frac=1.0)  # everything behind the camera
    sc.bg = np.array([0.2, 0.4, 0.6], np.float32)
    cam = scenes.make_camera(W, H, 0)
    grads = scenes.make_upstream_grads(W, H, 16)
    mine = h.run_mine(sc, cam, grads)
    assert mine["R"] == 0
    assert (h.npy(mine["radii"]) == 0).all() and not h.npy(mine["is_used"]).any()
    assert torch.allclose(mine["color"], torch.tensor(sc.bg).cuda()[:, None, None].expand(3, H, W))
    assert (mine["depth"] == 15.0).all() and (mine["objects"] == 0).all()
    for k in h.GRAD_KEYS:
        assert (h.npy(mine[k]) == 0).all(), k
    one = scenes.make_scene(1, sh_degree=0, seed=6, W=W, H=H, behind_frac=0.0, scale_mul=5.0)
    m1 = h.run_mine(one, cam, grads)
    o1 = h.run_oracle(one, cam, grads)
    assert m1["R"] == o1["R"]
    _check_images(m1, o1)


def test_argument_errors():
    import torch
    from gsraster_b200 import capi
    sc = scenes.make_scene(10, sh_degree=0, seed=5, W=64, H=48)
    cam = scenes.make_camera(64, 48, 0)
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    with pytest.raises(RuntimeError):  # neither shs nor colors_precomp
        capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, 48, 64, e, d["sh_objs"],
                                 0, d["campos"], False, False)
    with pytest.raises(RuntimeError):  # means3D wrong shape (rasterize_points.cu:58-60)
        capi.rasterize_gaussians(d["bg"], d["means3D"].view(-1), e, d["opacities"], d["scales"], d["rotations"], 1.0,
                                 e, d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, 48, 64, d["shs"],
                                 d["sh_objs"], 0, d["campos"], False, False)
    with pytest.raises(RuntimeError):  # CPU tensor: there is no CPU fallback
        capi.rasterize_gaussians(d["bg"].cpu(), d["means3D"].cpu(), e, d["opacities"].cpu(), d["scales"].cpu(),
                                 d["rotations"].cpu(), 1.0, e, d["viewmatrix"].cpu(), d["projmatrix"].cpu(),
                                 cam.tanfovx, cam.tanfovy, 48, 64, d["shs"].cpu(), d["sh_objs"].cpu(), 0,
                                 d["campos"].cpu(), False, False)


def test_forward_is_deterministic_and_backward_stable():
    sc = scenes.make_scene(50_000, sh_degree=3, seed=9, W=640, H=512, scale_mul=2.0)
    cam = scenes.make_camera(640, 512, 0)
    grads = scenes.make_upstream_grads(640, 512, 16)
    a = h.run_mine(sc, cam, grads)
    b = h.run_mine(sc, cam, grads)
    for k in ("color", "depth", "objects", "n_contrib", "point_list"):
        assert (h.npy(a[k]) == h.npy(b[k])).all(), k
    for k in h.GRAD_KEYS:
        assert h.grad_rel_err(a[k], b[k]) < 5e-5, k  # float atomics: run-to-run spread, bar is 1e-3


# ---- drop-in Python surface (reference diff_gaussian_rasterization/__init__.py, simple_knn/_C) -----
def test_dropin_module_autograd_matches_capi():
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from simple_knn._C import distCUDA2
    W, H, P = 320, 256, 20_000
    sc = scenes.make_scene(P, sh_degree=2, F=16, seed=31, W=W, H=H, scale_mul=2.0)
    cam = scenes.make_camera(W, H, 1)
    grads = scenes.make_upstream_grads(W, H, 16)
    d = h.to_torch(sc, cam)
    prm = {k: d[k].clone().requires_grad_(True) for k in ("means3D", "shs", "sh_objs", "opacities", "scales", "rotations")}
    means2D = torch.zeros_like(prm["means3D"], requires_grad=True)  # like render_4 (:356)
    settings = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, d["viewmatrix"],
                                             d["projmatrix"], sc.sh_degree, d["campos"], False, False)
    rast = GaussianRasterizer(settings)
    depth, color, radii, objects, is_used = rast(means3D=prm["means3D"], means2D=means2D, opacities=prm["opacities"],
                                                 shs=prm["shs"], sh_objs=prm["sh_objs"], scales=prm["scales"],
                                                 rotations=prm["rotations"])
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]
    ((depth * gd).sum() + (color * gc).sum() + (objects * go).sum()).backward()
    mine = h.run_mine(sc, cam, grads)
    assert (radii == mine["radii"]).all() and (is_used == mine["is_used"]).all()
    assert torch.equal(color, mine["color"]) and torch.equal(depth, mine["depth"]) and torch.equal(objects, mine["objects"])
    pairs = (("means3D", "dL_dmeans3D"), ("shs", "dL_dsh"), ("sh_objs", "dL_dobjects"), ("opacities", "dL_dopacity"),
             ("scales", "dL_dscales"), ("rotations", "dL_drotations"))
    for k, gk in pairs:
        assert prm[k].grad.shape == prm[k].shape
        assert h.grad_rel_err(prm[k].grad.reshape(mine[gk].shape), mine[gk]) < 1e-5, k
    assert h.grad_rel_err(means2D.grad, mine["dL_dmeans2D"]) < 1e-5
    vis = rast.markVisible(d["means3D"])
    assert vis.dtype == torch.bool and int(vis.sum()) >= int((radii > 0).sum())
    assert distCUDA2(d["means3D"]).shape == (P,)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        rast(means3D=prm["means3D"], means2D=means2D, opacities=prm["opacities"], sh_objs=prm["sh_objs"],
             scales=prm["scales"], rotations=prm["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair"):
        rast(means3D=prm["means3D"], means2D=means2D, opacities=prm["opacities"], shs=prm["shs"],
             sh_objs=prm["sh_objs"], scales=prm["scales"])


def test_libtorch_binding_matches_ctypes_binding():
    """csrc/torch_binding.cpp (C++/LibTorch, the reference's signatures) and the ctypes binding drive the same
    C ABI: forward outputs identical, gradients equal up to atomic ordering."""
    import torch
    from gsraster_b200 import capi, torch_ext
    m = torch_ext.load()
    W, H, P = 400, 300, 30_000
    sc = scenes.make_scene(P, sh_degree=3, F=16, seed=77, W=W, H=H, scale_mul=2.0)
    cam = scenes.make_camera(W, H, 2)
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    gd, gc, go = [torch.from_numpy(g).cuda() for g in scenes.make_upstream_grads(W, H, 16)]
    outs = []
    for mod in (m, capi):
        r = mod.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                    d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"],
                                    d["sh_objs"], 3, d["campos"], False, False)
        g = mod.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e,
                                             d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, gd, gc,
                                             d["shs"], go, d["sh_objs"], 3, d["campos"], r[6], r[0], r[7], r[8], False)
        outs.append((r, g))
    (ra, ga), (rb, gb) = outs
    assert ra[0] == rb[0] > 0
    for i in range(1, 6):
        assert torch.equal(ra[i], rb[i]), i
    for a, b in zip(ga, gb):
        assert a.shape == b.shape and h.grad_rel_err(a, b) < 1e-5
    assert torch.equal(m.mark_visible(d["means3D"], d["viewmatrix"], d["projmatrix"]),
                       capi.mark_visible(d["means3D"], d["viewmatrix"], d["projmatrix"]))
    assert torch.equal(m.distCUDA2(d["means3D"]), capi.distCUDA2(d["means3D"]))


# ---- camera-pose gradient (extension; parity unpinned by the reference) -------------------
def _camera_grad_run(sc, cam, grads):
    import torch
    from gsraster_b200 import capi
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, cam.H, cam.W, d["shs"],
                                 d["sh_objs"], sc.sh_degree, d["campos"], False, False)
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]
    g = capi.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e,
                                          d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, gd, gc, d["shs"], go,
                                          d["sh_objs"], sc.sh_degree, d["campos"], r[6], r[0], r[7], r[8], False,
                                          camera_grads=True)
    dV, dPV, dC = [t.double().cpu().numpy() for t in capi.rasterize_gaussians_backward.last_camera_grads]
    return r, g, dV, dPV, dC


@pytest.mark.parametrize("P,W,H,mul", [(20_000, 320, 256, 2.0), (1_000_000, 1280, 1024, 1.0)])
def test_camera_gradient_rigid_motion_identity(P, W, H, mul):
    """SH degree 0: L(V G, PV G) == L with every Gaussian moved by G (p -> G p, Sigma -> R Sigma R^T), so for any
    twist (omega, t) the directional derivative through the camera gradients must equal the one through
    dL_dmeans3D / dL_dcov3D. Second case = BASELINE config 5 size (1M Gaussians)."""
    sc = scenes.make_scene(P, sh_degree=0, F=16, seed=91, W=W, H=H, scale_mul=mul)
    cam = scenes.make_camera(W, H, 2)
    grads = scenes.make_upstream_grads(W, H, 16, seed=5)
    grads = tuple(g * (W * H) for g in grads)  # O(1) upstream gradients
    # No depth upstream gradient here: the camera gradient is the TRUE derivative of the forward pass (pinned by finite
    # differences, tests/test_pose_fd_gpu.py), while dL_dmeans3D / dL_dcov3D reproduce three shortcuts of the
    # reference's depth backward (backward.cu:263-268, :629, :631 vs :671), so the identity holds where those vanish.
    grads = (np.zeros_like(grads[0]), grads[1], grads[2])
    r, g, dV, dPV, dC = _camera_grad_run(sc, cam, grads)
    assert np.abs(dC).max() == 0  # colour does not depend on the view direction at degree 0
    dmean = g[4].double().cpu().numpy()
    dcov = g[5].double().cpu().numpy()
    # reconstruct Sigma (6) per Gaussian in double from scale / rotation (forward.cu:122-168 convention)
    q = sc.rotations.astype(np.float64)
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    Rm = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                   2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                   2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
    S2 = sc.scales.astype(np.float64) ** 2
    Sigma = np.einsum("nij,nj,nkj->nik", Rm, S2, Rm)
    dS = np.zeros_like(Sigma)  # dL/dSigma as a full symmetric-matrix gradient: off-diagonals split in two
    dS[:, 0, 0], dS[:, 1, 1], dS[:, 2, 2] = dcov[:, 0], dcov[:, 3], dcov[:, 5]
    dS[:, 0, 1] = dS[:, 1, 0] = 0.5 * dcov[:, 1]
    dS[:, 0, 2] = dS[:, 2, 0] = 0.5 * dcov[:, 2]
    dS[:, 1, 2] = dS[:, 2, 1] = 0.5 * dcov[:, 4]
    V = cam.viewmatrix.T.astype(np.float64)    # math matrices: flat index 4c + r
    PV = cam.projmatrix.T.astype(np.float64)
    gV, gPV = dV.T, dPV.T                       # dL/dV[r][c]
    p = sc.means3D.astype(np.float64)
    rng = np.random.default_rng(0)
    for _ in range(4):
        om, t = rng.normal(size=3), rng.normal(size=3)
        K = np.array([[0, -om[2], om[1]], [om[2], 0, -om[0]], [-om[1], om[0], 0]])
        Gd = np.zeros((4, 4)); Gd[:3, :3] = K; Gd[:3, 3] = t  # d/d eps of G = exp(eps * twist) at 0
        cam_side = (gV * (V @ Gd)).sum() + (gPV * (PV @ Gd)).sum()
        dp = p @ K.T + t
        dSig = K @ Sigma + Sigma @ K.T
        gauss_side = (dmean * dp).sum() + (dS * dSig).sum()
        assert abs(cam_side - gauss_side) <= 2e-3 * max(abs(cam_side), abs(gauss_side), 1e-6), (cam_side, gauss_side)


def test_camera_centre_gradient_finite_differences():
    """campos only enters through the SH view direction (forward.cu:25-27): colours are smooth in it and the
    geometry does not move, so central differences of the forward pass are clean."""
    import torch
    from gsraster_b200 import capi
    W, H, P = 320, 256, 20_000
    sc = scenes.make_scene(P, sh_degree=3, F=16, seed=92, W=W, H=H, scale_mul=2.0)
    sc.shs[:, 1:, :] *= 6.0  # make the view-dependent part matter
    cam = scenes.make_camera(W, H, 1)
    grads = tuple(g * (W * H) for g in scenes.make_upstream_grads(W, H, 16, seed=6))
    r, g, dV, dPV, dC = _camera_grad_run(sc, cam, grads)
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    gc = torch.from_numpy(grads[1]).cuda().double()

    def loss(c):
        rr = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                      d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"],
                                      d["sh_objs"], 3, torch.tensor(c, dtype=torch.float32, device="cuda"), False, False)
        return float((rr[2].double() * gc).sum())

    c0 = cam.campos.astype(np.float64)
    for k in range(3):
        eps = 2e-2
        cp, cm = c0.copy(), c0.copy()
        cp[k] += eps; cm[k] -= eps
        fd = (loss(cp) - loss(cm)) / (2 * eps)
        assert abs(fd - dC[k]) <= 2e-2 * max(abs(fd), np.abs(dC).max()), (k, fd, dC[k])


def test_tracking_autograd_wrapper():
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    from gsraster_b200 import tracking
    W, H, P = 160, 128, 5_000
    sc = scenes.make_scene(P, sh_degree=1, F=16, seed=93, W=W, H=H, scale_mul=3.0)
    cam = scenes.make_camera(W, H, 0)
    d = h.to_torch(sc, cam)
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, d["viewmatrix"], d["projmatrix"],
                                       1, d["campos"], False, False)
    V = d["viewmatrix"].clone().requires_grad_(True)
    PV = d["projmatrix"].clone().requires_grad_(True)
    c = d["campos"].clone().requires_grad_(True)
    depth, color, radii, objects, used = tracking.rasterize_with_camera_grads(
        V, PV, c, d["means3D"], d["shs"], d["sh_objs"], d["opacities"], d["scales"], d["rotations"], rs)
    (color.sum() + 0.1 * depth.sum()).backward()
    assert V.grad.shape == (4, 4) and PV.grad.shape == (4, 4) and c.grad.shape == (3,)
    assert float(V.grad.abs().sum()) > 0 and float(PV.grad.abs().sum()) > 0 and float(c.grad.abs().sum()) > 0
    assert float(V.grad[:, 3].abs().sum()) == 0  # flat index 4c + 3 == row 3 of V: unused by the rasterizer


@pytest.mark.skipif(not os.path.exists(h.REF_KNN_LIB), reason="oracle/_ref knn not shipped")
def test_knn_config5_5M_points():
    import torch
    from gsraster_b200 import capi
    pts = scenes.make_points(5_000_000, seed=55)
    mine = capi.distCUDA2(torch.from_numpy(pts).cuda())
    ref = h.run_ref_knn(pts)
    assert torch.equal(mine.view(torch.int32), ref.view(torch.int32))


@pytest.mark.gpu
def test_touched_mask_marks_exactly_the_nonzero_gradient_rows():
    """gsr_ext::touched_mask: Gaussians the backward blend did not reach have exactly-zero parameter gradients
    (that is what lets the multi-GPU exchange skip them); touched ones are a small part of the scene."""
    import torch
    from gsraster_b200 import capi
    W, H, P = 640, 512, 60000
    sc = scenes.make_scene(P, sh_degree=3, F=16, seed=21, W=W, H=H)
    cam = scenes.make_camera(W, H, 1)
    grads = scenes.make_upstream_grads(W, H, 16)
    mask = torch.zeros(P, dtype=torch.uint8, device="cuda")
    mine = h.run_mine(sc, cam, grads, ext=capi.Ext(touched_mask=mask))
    m = mask.bool()
    assert 0 < int(m.sum()) < P
    for k in h.GRAD_KEYS:
        gk = mine[k].reshape(P, -1)
        assert float(gk[~m].abs().max()) == 0.0, k
    # and every row with a non-zero gradient is flagged
    nz = torch.zeros(P, dtype=torch.bool, device="cuda")
    for k in h.GRAD_KEYS:
        nz |= mine[k].reshape(P, -1).abs().sum(1) > 0
    assert bool((nz & ~m).sum() == 0)


@pytest.mark.gpu
def test_exchange_kernels_match_torch_reference():
    """csrc/exchange.cu: union + compaction of gathered flags and the row gather / scatter, against torch ops."""
    import torch
    from gsraster_b200 import capi
    torch.manual_seed(5)
    for P, nranks, dens in ((1, 1, 1.0), (4099, 3, 0.02), (250001, 8, 0.3), (100000, 2, 0.0)):
        Ppad = (P + 3) // 4 * 4
        flags = torch.zeros(nranks, Ppad, dtype=torch.uint8, device="cuda")
        flags[:, :P] = (torch.rand(nranks, P, device="cuda") < dens).to(torch.uint8)
        ref_idx = torch.nonzero(flags[:, :P].amax(0)).squeeze(1)
        U = int(ref_idx.numel())
        for cap in (P, max(U // 2, 1)):
            idx = torch.full((P + 1,), -1, dtype=torch.int32, device="cuda")
            count = torch.zeros(1, dtype=torch.int32, device="cuda")
            scratch = torch.empty(capi.exchange_scratch_bytes(P) + 256, dtype=torch.uint8, device="cuda")
            capi.exchange_compact(flags.view(-1), nranks, P, Ppad, cap, idx, count, scratch)
            assert int(count) == U
            k = min(U, cap)
            assert (idx[:k].long() == ref_idx[:k]).all()
            assert (idx[k:] == -1).all()            # nothing written past min(count, cap)
            sizes = (3, 48, 16, 1)
            dense = [torch.randn(P, n, device="cuda") for n in sizes]
            comp = [torch.full((cap, n), 7.0, device="cuda") for n in sizes]
            capi.exchange_rows(False, dense, comp, sizes, idx, count, cap)
            for d_, c_ in zip(dense, comp):
                assert (c_[:k] == d_[ref_idx[:k]]).all() and (c_[k:] == 0).all()
            back = [torch.zeros(P, n, device="cuda") for n in sizes]
            capi.exchange_rows(True, back, comp, sizes, idx, count, cap)
            for d_, b_ in zip(dense, back):
                exp = torch.zeros_like(d_)
                exp[ref_idx[:k]] = d_[ref_idx[:k]]
                assert (b_ == exp).all()


def test_sparse_exchange_with_overlapped_flag_exchange_single_rank_nccl():
    """The overlapped exchange (blend-done event -> side stream all-gather + compaction beside the per-Gaussian
    backward) on a 1-rank NCCL group: stream plumbing + result against the dense bucket. The N-rank version is
    tests/gpu_exchange_check.py (torchrun)."""
    import socket
    import torch
    import torch.distributed as dist
    import gpu_exchange_check as gx
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1,
                                device_id=torch.device("cuda", 0))
    try:
        worst, rows, fb = gx.run_check(P=50000, W=320, H=256, steps=3)
        assert worst <= 1e-5 and rows > 0 and fb <= 1, (worst, rows, fb)   # (two backward runs: atomic order differs)
        # two views per rank accumulated into the sparse rows of one bucket (gsr_ext::accumulate)
        worst, rows, fb = gx.run_check(P=50000, W=320, H=256, steps=3, V=2)
        assert worst <= 1e-5 and rows > 0 and fb <= 1, (worst, rows, fb)
    finally:
        if created:
            dist.destroy_process_group()
