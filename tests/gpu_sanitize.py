"""GPU diagnostic (not a pytest): one small pass over every product kernel on small, odd-sized scenes. Written for
`compute-sanitizer --tool memcheck`; that tool is closed on this GPU pool (profiles/r02_sanitizer_memcheck.log), so the
script brings its own bounds checks: every scratch buffer the C ABI asks for gets a canary zone behind the requested
size that is verified when the call returns (capi._Resizer.debug_canary), the backward outputs live inside canary-
padded buffers, and results are compared with the CPU oracle."""
import torch

import helpers as h
from helpers import scenes
from gsraster_b200 import capi, optim

capi._Resizer.debug_canary = True


def padded(n, dtype=torch.float32):
    """n elements in the middle of a buffer whose 1 KB margins hold a pattern; check() verifies the margins."""
    buf = torch.full((n + 512,), 7.25 if dtype.is_floating_point else 77, dtype=dtype, device="cuda")
    return buf, buf[256:256 + n]


def check(buf, n, what):
    ref = 7.25 if buf.dtype.is_floating_point else 77
    assert bool((buf[:256] == ref).all()) and bool((buf[256 + n:] == ref).all()), "out-of-bounds write around " + what


for (P, W, H, F, deg) in ((3000, 200, 136, 16, 3), (1500, 131, 77, 32, 1), (5, 64, 64, 16, 0)):
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=3, W=W, H=H)
    cam = scenes.make_camera(W, H, 1)
    grads = scenes.make_upstream_grads(W, H, F)
    mask = torch.zeros(P, dtype=torch.uint8, device="cuda")
    out = h.run_mine(sc, cam, grads, ext=capi.Ext(touched_mask=mask))   # forward, export (full pull), backward
    # the per-call extensions: sparse rows accumulated over two calls into a persistent row workspace + index list,
    # the camera gradient in full and in tracking mode
    ws = capi.new_row_workspace(P, F, "cuda")
    m2 = torch.zeros(P, dtype=torch.uint8, device="cuda")
    tidx = torch.zeros(P, dtype=torch.int32, device="cuda")
    tcnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    ext = capi.Ext(touched_mask=m2, sparse_rows=True, accumulate=True, row_workspace=ws, touched_idx=tidx, touched_count=tcnt)
    for _ in range(2):
        h.run_mine(sc, cam, grads, ext=ext)
    assert int(ws.count_nonzero()) == 0 and int(tcnt) == int(m2.sum())
    # dense backward into canary-padded output tensors, checked against the CPU oracle
    Mm = (deg + 1) ** 2
    sizes_o = dict(means3D=3, sh=3 * Mm, objects=F, opacity=1, scales=3, rotations=4, means2D=3, conic=6, depths=1,
                   colors=3, cov3D=6)
    bufs = {k: padded(P * n) for k, n in sizes_o.items()}
    res = h.run_mine(sc, cam, grads)
    dd = h.to_torch(sc, cam)
    ee = torch.empty(0, device="cuda")
    rr = capi.rasterize_gaussians(dd["bg"], dd["means3D"], ee, dd["opacities"], dd["scales"], dd["rotations"], 1.0, ee,
                                  dd["viewmatrix"], dd["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, dd["shs"],
                                  dd["sh_objs"], deg, dd["campos"], False, False)
    gg = [torch.from_numpy(x).cuda() for x in grads]
    capi.rasterize_gaussians_backward(dd["bg"], dd["means3D"], rr[4], ee, dd["scales"], dd["rotations"], 1.0, ee,
                                      dd["viewmatrix"], dd["projmatrix"], cam.tanfovx, cam.tanfovy, gg[0], gg[1], dd["shs"],
                                      gg[2], dd["sh_objs"], deg, dd["campos"], rr[6], rr[0], rr[7], rr[8], False,
                                      out={k: v[1] for k, v in bufs.items()})
    torch.cuda.synchronize()
    for k, n in sizes_o.items():
        check(bufs[k][0], P * n, "dL_d" + k)
    orc = h.run_oracle(sc, cam, grads)
    if res["R"] == orc["R"]:
        for k in ("dL_dmeans3D", "dL_dsh", "dL_dobjects"):
            assert h.grad_rel_err(res[k], orc[k]) < 5e-3, k
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"], d["sh_objs"],
                                 deg, d["campos"], False, False)
    gt = [torch.from_numpy(x).cuda() for x in grads]
    for trk in (False, True):
        capi.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e, d["viewmatrix"],
                                          d["projmatrix"], cam.tanfovx, cam.tanfovy, gt[0], gt[1], d["shs"], gt[2],
                                          d["sh_objs"], deg, d["campos"], r[6], r[0], r[7], r[8], False, camera_grads=True,
                                          ext=capi.Ext(tracking=trk))
    print("scene", P, W, H, F, "R", out["R"], "touched", int(mask.sum()))
    Ppad = (P + 3) // 4 * 4
    flags = torch.zeros(2, Ppad, dtype=torch.uint8, device="cuda")
    flags[0, :P] = mask
    flags[1, :P] = (torch.rand(P, device="cuda") < 0.1).to(torch.uint8)
    idx = torch.zeros(P + 1, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    scratch = torch.empty(capi.exchange_scratch_bytes(P) + 256, dtype=torch.uint8, device="cuda")
    cap = max(P // 2, 1)
    capi.exchange_compact(flags.view(-1), 2, P, Ppad, cap, idx, cnt, scratch)
    sizes = (3, 3 * (deg + 1) ** 2, F, 1, 3, 4)
    dense = [torch.randn(P, n, device="cuda") for n in sizes]
    comp = [torch.empty(cap, n, device="cuda") for n in sizes]
    capi.exchange_rows(False, dense, comp, sizes, idx, cnt, cap)
    capi.exchange_rows(True, dense, comp, sizes, idx, cnt, cap)
    M = (deg + 1) ** 2
    raw = dict(xyz=torch.randn(P, 3, device="cuda"), f_dc=torch.randn(P, 1, 3, device="cuda"),
               f_rest=torch.randn(P, M - 1, 3, device="cuda"), opacity=torch.randn(P, 1, device="cuda"),
               scaling=torch.randn(P, 3, device="cuda"), rotation=torch.randn(P, 4, device="cuda"),
               objects=torch.randn(P, 1, F, device="cuda"))
    fa = optim.FusedGaussianAdam(raw)
    g = dict(means3D=dense[0], sh=dense[1].view(P, M, 3), objects=dense[2], opacity=dense[3], scales=dense[4],
             rotations=dense[5])
    fa.step(g)
    gc = dict(means3D=comp[0], sh=comp[1], objects=comp[2], opacity=comp[3], scales=comp[4], rotations=comp[5])
    fa.step(gc, idx=idx[:cap], count=cnt)
    torch.cuda.synchronize()
pts = scenes.make_points(2000, seed=5)
capi.distCUDA2(torch.from_numpy(pts).cuda())
torch.cuda.synchronize()
print("sanitize pass done")
