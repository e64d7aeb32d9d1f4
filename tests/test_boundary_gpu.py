"""GPU tests of the drop-in boundary: the C ABI has no mode switches and no per-process guards, so it
can be called from several host threads and on several devices of one process; everything a call does differently
from the reference travels with that call (gsr_ext). Also the reference-visible debug behaviour of the Python module
(snapshot dumps, diff_gaussian_rasterization/__init__.py:87-94, 140-147)."""
import os
import threading

import numpy as np
import pytest

import helpers as h
from helpers import scenes

pytestmark = pytest.mark.gpu


def _scene(P=20_000, W=320, H=256, deg=3, F=16, seed=41, view=1, mul=2.0):
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=seed, W=W, H=H, scale_mul=mul)
    return sc, scenes.make_camera(W, H, view), scenes.make_upstream_grads(W, H, F)


def _same(a, b, tol=5e-5):
    for k in ("color", "depth", "objects", "n_contrib", "point_list", "radii"):
        assert (h.npy(a[k]) == h.npy(b[k])).all(), k
    for k in h.GRAD_KEYS:
        assert h.grad_rel_err(a[k], b[k]) < tol, k


def test_two_host_threads_concurrently():
    """One process, two host threads, each with its own stream on cuda:0: results equal the single-threaded run."""
    import torch
    sc, cam, grads = _scene()
    base = h.run_mine(sc, cam, grads)
    out, err = [None, None], []

    def work(i):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                for _ in range(3):
                    out[i] = h.run_mine(sc, cam, grads)
        except Exception as ex:  # pragma: no cover
            err.append(ex)

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not err, err
    for o in out:
        _same(o, base)


def test_two_devices_from_one_process():
    """cudaFuncSetAttribute is per device and the helper streams / events belong to a device: forward + backward on
    cuda:0 and cuda:1 from the same process (the backward blend needs ~100 KB of opted-in dynamic shared memory)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sc, cam, grads = _scene()
    a = h.run_mine(sc, cam, grads, device="cuda:0")
    b = h.run_mine(sc, cam, grads, device="cuda:1")
    c = h.run_mine(sc, cam, grads, device="cuda:0")
    _same(a, {k: (v.to("cuda:0") if hasattr(v, "to") else v) for k, v in b.items()})
    _same(a, c)


@pytest.mark.parametrize("mul,opacity_cap,min_frac", [(1.5, None, 0.0), (0.5, 0.08, 0.3)])
def test_sparse_rows_accumulate_workspace_match_dense(mul, opacity_cap, min_frac):
    """gsr_ext: sparse_rows + accumulate + row_workspace + touched_idx. Two views accumulated into one set of gradient
    tensors equal the sum of the two dense backward passes on the flagged rows; unflagged rows are not written; the
    persistent workspace is all-zero again after every call. Second case: faint small splats, so that pixels do not
    saturate and most Gaussians are reached (the coalesced, row-masked store path of the per-Gaussian backward)."""
    import torch
    from gsraster_b200 import capi
    P, W, H, F, deg = 30_000, 400, 300, 16, 3
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=43, W=W, H=H, scale_mul=mul)
    if opacity_cap is not None:
        sc.opacities = np.minimum(sc.opacities, np.float32(opacity_cap))
    grads = scenes.make_upstream_grads(W, H, F)
    dense = [h.run_mine(sc, scenes.make_camera(W, H, v), grads) for v in (0, 2)]
    mask = torch.zeros(P, dtype=torch.uint8, device="cuda")
    ws = capi.new_row_workspace(P, F, "cuda")
    idx = torch.full((P,), -1, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    ext = capi.Ext(touched_mask=mask, sparse_rows=True, accumulate=True, row_workspace=ws, touched_idx=idx,
                   touched_count=cnt)
    M = (deg + 1) ** 2
    o = dict(device="cuda", dtype=torch.float32)
    SENT = 12345.0
    out = dict(means3D=torch.full((P, 3), SENT, **o), sh=torch.full((P, M, 3), SENT, **o), objects=torch.full((P, F), SENT, **o),
               opacity=torch.full((P, 1), SENT, **o), scales=torch.full((P, 3), SENT, **o), rotations=torch.full((P, 4), SENT, **o))
    e = torch.empty(0, device="cuda")
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]
    for v in (0, 2):
        cam = scenes.make_camera(W, H, v)
        d = h.to_torch(sc, cam)
        r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                     d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"], d["sh_objs"],
                                     deg, d["campos"], False, False)
        capi.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e,
                                          d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, gd, gc, d["shs"], go,
                                          d["sh_objs"], deg, d["campos"], r[6], r[0], r[7], r[8], False, out=out, ext=ext)
        assert int(ws.count_nonzero()) == 0, "row workspace must be clean after the call"
    m = mask.bool()
    U = int(cnt)
    assert U == int(m.sum()) and min_frac * P < U < P
    assert torch.equal(torch.sort(idx[:U].long())[0], torch.nonzero(m).squeeze(1))
    pairs = (("means3D", "dL_dmeans3D"), ("sh", "dL_dsh"), ("objects", "dL_dobjects"), ("opacity", "dL_dopacity"),
             ("scales", "dL_dscales"), ("rotations", "dL_drotations"))
    for k, gk in pairs:
        want = (dense[0][gk] + dense[1][gk]).reshape(P, -1)
        got = out[k].reshape(P, -1)
        assert bool((got[~m] == SENT).all()), (k, "unflagged rows must be left untouched")
        assert float(want[~m].abs().max()) == 0.0, (k, "a row with a gradient is not flagged")
        assert h.grad_rel_err(got[m], want[m]) < 2e-5, k


def test_tracking_mode_gives_the_same_camera_gradients():
    """gsr_ext::tracking: the backward blend skips the feature-channel reductions and the per-Gaussian
    backward writes nothing per Gaussian; the camera gradients equal those of the full backward."""
    import torch
    from gsraster_b200 import capi
    sc, cam, grads = _scene(P=40_000, W=480, H=320, deg=3, view=2)
    grads = tuple(g * (480 * 320) for g in grads)
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, cam.H, cam.W, d["shs"],
                                 d["sh_objs"], sc.sh_degree, d["campos"], False, False)
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]
    args = (d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e, d["viewmatrix"], d["projmatrix"],
            cam.tanfovx, cam.tanfovy, gd, gc, d["shs"], go, d["sh_objs"], sc.sh_degree, d["campos"], r[6], r[0], r[7], r[8],
            False)
    full = capi.Ext()
    capi.rasterize_gaussians_backward(*args, camera_grads=True, ext=full)
    trk = capi.Ext(tracking=True)
    res = capi.rasterize_gaussians_backward(*args, camera_grads=True, ext=trk)
    assert all(t.numel() == 0 for t in res)
    for a, b in zip(full.camera_grads, trk.camera_grads):
        assert float(a.abs().max()) > 0
        assert h.grad_rel_err(b, a) < 2e-4
    with pytest.raises(RuntimeError):  # tracking without the camera outputs has nothing to compute
        capi.rasterize_gaussians_backward(*args, camera_grads=False, ext=capi.Ext(tracking=True))


def test_ext_argument_checks():
    import torch
    from gsraster_b200 import capi
    sc, cam, grads = _scene(P=500, W=64, H=48, deg=0)
    with pytest.raises(RuntimeError, match="stale mask"):   # a mask of another map size is refused, not overrun
        h.run_mine(sc, cam, grads, ext=capi.Ext(touched_mask=torch.zeros(400, dtype=torch.uint8, device="cuda")))
    with pytest.raises(RuntimeError, match="sparse_rows needs touched_mask"):
        h.run_mine(sc, cam, grads, ext=capi.Ext(sparse_rows=True))
    with pytest.raises(RuntimeError, match="accumulate"):
        h.run_mine(sc, cam, grads, ext=capi.Ext(accumulate=True))
    with pytest.raises(RuntimeError, match="smaller than"):
        h.run_mine(sc, cam, grads, ext=capi.Ext(row_workspace=torch.zeros(1024, dtype=torch.uint8, device="cuda")))


def test_stage_timer_and_stats_are_per_call():
    from gsraster_b200 import capi
    sc, cam, grads = _scene()
    t = capi.Timer()
    ext = capi.Ext(timer=t)
    import torch
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, cam.H, cam.W, d["shs"],
                                 d["sh_objs"], sc.sh_degree, d["campos"], False, False, ext=ext)
    assert ext.stats.num_rendered == r[0] > 0 and ext.stats.launches >= 5 and ext.stats.l1_total > 0
    import ctypes as C
    L = capi.lib()
    L.gsr_binning_bytes_bound.restype = C.c_size_t
    L.gsr_binning_bytes_bound.argtypes = [C.c_int64, C.c_int, C.c_int, C.c_int]
    # the pre-sizing bound covers point list + blend masks + the supertile lists this call really built
    assert L.gsr_binning_bytes_bound(r[0], sc.P, cam.W, cam.H) >= L.gsr_binning_bytes(r[0]) + 8 * ext.stats.l1_total
    ms, gaps = t.read()
    assert all(ms[k] > 0 for k in ("preprocess", "depth_sort_offsets", "binning", "render_fwd"))
    assert ms["render_bwd"] == 0 and ms["preprocess_bwd"] == 0
    h.run_mine(sc, cam, grads)            # a call without the timer records nothing into it
    ms2, _ = t.read()
    assert sum(ms2.values()) == 0


def test_num_rendered_beyond_int32_is_refused():
    """R = sum of tiles touched must fit an int (the reference reads it into one, rasterizer_impl.cu:289-290): 600k
    screen-filling splats on 1920x1080 give R = 4.9e9; the call fails cleanly before any R-sized allocation."""
    import torch
    from gsraster_b200 import capi
    W, H, P = 1920, 1080, 600_000
    sc = scenes.make_scene(P, sh_degree=0, F=16, seed=3, W=W, H=H, behind_frac=0.0)
    sc.means3D[:, :2] *= 0.05
    sc.means3D[:, 2] = 1.0
    sc.scales[:] = 3.0
    cam = scenes.make_camera(W, H, 0)
    asked = []
    with pytest.raises(RuntimeError, match="exceeds int32"):
        h.run_mine(sc, cam)
    torch.cuda.synchronize()
    small = h.run_mine(*_scene(P=2000, W=128, H=96)[:2])   # the library is usable afterwards
    assert small["R"] > 0


def test_debug_flag_writes_the_reference_snapshots(tmp_path, monkeypatch):
    """debug=True: CPU copies of the arguments are dumped to snapshot_fw.dump / snapshot_bw.dump when the extension
    throws (reference :87-94, :140-147); the per-stage checks of the C ABI run (gsr_params::debug)."""
    import torch
    import diff_gaussian_rasterization as dgr
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    monkeypatch.chdir(tmp_path)
    W, H, P = 160, 128, 3000
    sc = scenes.make_scene(P, sh_degree=1, F=16, seed=17, W=W, H=H, scale_mul=2.0)
    cam = scenes.make_camera(W, H, 0)
    d = h.to_torch(sc, cam)
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, d["viewmatrix"], d["projmatrix"], 1,
                                       d["campos"], False, True)
    rast = GaussianRasterizer(rs)
    prm = {k: d[k].clone().requires_grad_(True) for k in ("means3D", "shs", "sh_objs", "opacities", "scales", "rotations")}
    m2 = torch.zeros_like(prm["means3D"], requires_grad=True)
    depth, color, radii, objects, used = rast(means3D=prm["means3D"], means2D=m2, opacities=prm["opacities"], shs=prm["shs"],
                                              sh_objs=prm["sh_objs"], scales=prm["scales"], rotations=prm["rotations"])
    color.sum().backward()      # debug path, no error: no dump
    assert not os.path.exists("snapshot_fw.dump") and not os.path.exists("snapshot_bw.dump")
    # forward failure: a bad means3D shape makes the extension throw (rasterize_points.cu:58-60)
    with pytest.raises(Exception):
        dgr.rasterize_gaussians(prm["means3D"].detach().view(-1), m2, prm["shs"], prm["sh_objs"], torch.Tensor([]),
                                prm["opacities"], prm["scales"], prm["rotations"], torch.Tensor([]), rs)
    assert os.path.exists("snapshot_fw.dump")
    snap = torch.load("snapshot_fw.dump")
    assert len(snap) == 20 and snap[1].device.type == "cpu" and snap[1].numel() == 3 * P
    # backward failure: upstream colour gradient of the wrong size
    depth, color, radii, objects, used = rast(means3D=prm["means3D"], means2D=m2, opacities=prm["opacities"], shs=prm["shs"],
                                              sh_objs=prm["sh_objs"], scales=prm["scales"], rotations=prm["rotations"])
    orig = dgr._C.rasterize_gaussians_backward

    def boom(*a, **k):
        raise RuntimeError("injected backward failure")

    monkeypatch.setattr(dgr._C, "rasterize_gaussians_backward", boom)
    with pytest.raises(RuntimeError, match="injected"):
        color.sum().backward()
    monkeypatch.setattr(dgr._C, "rasterize_gaussians_backward", orig)
    assert os.path.exists("snapshot_bw.dump")
    assert len(torch.load("snapshot_bw.dump")) == 24


def test_view_pipeline_two_streams_matches_one_stream():
    """mapping.ViewPipeline: the views of a step alternate between two streams (forward of view j + 1 beside backward of
    view j, backward passes chained by an event). Four views accumulated through the pipeline equal the same four views
    accumulated on one stream (same flags, rows equal up to the order of the float atomics), repeatedly, and the
    scratch pool never hands a buffer of one stream to a call on another."""
    import torch
    from gsraster_b200 import capi, mapping
    P, W, H, F, deg = 60_000, 640, 512, 16, 3
    sc = scenes.make_trajectory_scene(P, 4, sh_degree=deg, F=F, seed=47, W=W, H=H)
    grads = scenes.make_upstream_grads(W, H, F)
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]
    e = torch.empty(0, device="cuda")
    cams = [scenes.make_camera(W, H, v) for v in range(4)]
    ds = [h.to_torch(sc, c) for c in cams]
    M = (deg + 1) ** 2

    def run(pipe):
        mask = torch.zeros(P, dtype=torch.uint8, device="cuda")
        ws = capi.new_row_workspace(P, F, "cuda")
        ext = capi.Ext(touched_mask=mask, sparse_rows=True, accumulate=True, row_workspace=ws)
        o = dict(device="cuda", dtype=torch.float32)
        out = dict(means3D=torch.zeros((P, 3), **o), sh=torch.zeros((P, M, 3), **o), objects=torch.zeros((P, F), **o),
                   opacity=torch.zeros((P, 1), **o), scales=torch.zeros((P, 3), **o), rotations=torch.zeros((P, 4), **o))
        if pipe:
            pipe.begin()
        for cam, d in zip(cams, ds):
            with (pipe.view() if pipe else torch.cuda.stream(torch.cuda.current_stream())):
                r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                             d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"],
                                             d["sh_objs"], deg, d["campos"], False, False)
                if pipe:
                    pipe.before_backward()
                capi.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e,
                                                  d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, gd, gc, d["shs"],
                                                  go, d["sh_objs"], deg, d["campos"], r[6], r[0], r[7], r[8], False, out=out,
                                                  ext=ext)
                if pipe:
                    pipe.after_backward()
        if pipe:
            pipe.end()
        torch.cuda.synchronize()
        assert int(ws.count_nonzero()) == 0
        return mask.bool(), out

    m0, ref = run(None)
    vp = mapping.ViewPipeline(torch.device("cuda"))
    for _ in range(3):
        m1, got = run(vp)
        assert torch.equal(m0, m1) and int(m0.sum()) > 0
        for k in ref:
            assert h.grad_rel_err(got[k].reshape(P, -1)[m0], ref[k].reshape(P, -1)[m0]) < 2e-5, k
