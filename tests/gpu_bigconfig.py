"""GPU diagnostic (not a pytest): large configurations (BASELINE config 4: 6M Gaussians, 1920x1080, F=32) --
product vs the unmodified reference build, compared ON THE DEVICE (no multi-GB host copies), plus timings.
Usage: python tests/gpu_bigconfig.py P W H F [deg]"""
import sys
import time

import numpy as np
import torch

import helpers as h
from helpers import scenes
from gsraster_b200 import capi
from oracle import refcuda

P, W, H, F = [int(x) for x in sys.argv[1:5]]
deg = int(sys.argv[5]) if len(sys.argv) > 5 else 3
sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=1238, W=W, H=H)
cam = scenes.make_camera(W, H, 0)
d = h.to_torch(sc, cam)
g = [torch.from_numpy(x).cuda() for x in scenes.make_upstream_grads(W, H, F)]
e = torch.empty(0, device="cuda")
M = (deg + 1) ** 2


def mine():
    r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"], d["sh_objs"],
                                 deg, d["campos"], False, False)
    gr = capi.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e,
                                           d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, g[0], g[1], d["shs"],
                                           g[2], d["sh_objs"], deg, d["campos"], r[6], r[0], r[7], r[8], False)
    return r, gr


def timeit(fn, n=5):
    for _ in range(2):
        out = fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        out = fn()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / n, out


ms_mine, (r, gr) = timeit(mine)
R = r[0]
print(f"P={P} {W}x{H} F={F} deg={deg}: R={R} mine fwd+bwd {ms_mine:.3f} ms  peak mem {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
st = capi.export_state(P, deg, M, F, W, H, R, r[6], r[7], r[8])
if refcuda.available(F):
    ref = refcuda.RefRasterizer(F)
    dd = dict(d, sh_objs=d["sh_objs"])

    def reff():
        Rr, depth, color, objects, radii, is_used, bufs = ref.forward(dd, W, H, cam.tanfovx, cam.tanfovy, deg)
        gg = ref.backward(dd, W, H, cam.tanfovx, cam.tanfovy, deg, Rr, radii, bufs, g[0], g[1], g[2])
        return (Rr, depth, color, objects, radii, is_used, bufs), gg

    ms_ref, (rr, gg) = timeit(reff, n=3)
    print(f"  reference fwd+bwd {ms_ref:.3f} ms  -> x{ms_ref / ms_mine:.2f}")
    assert rr[0] == R, (rr[0], R)
    print("  radii equal", bool(torch.equal(rr[4], r[4])), " is_used equal", bool(torch.equal(rr[5], r[5])))
    import ctypes as C
    L = ref.L
    L.ref_decode_img.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    L.ref_decode_binning.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    ptrs = (C.c_void_p * 6)()
    ga, ba, ia = rr[6]
    L.ref_decode_img(ia.t.data_ptr(), W * H, ptrs)
    T = ((W + 15) // 16) * ((H + 15) // 16)

    def view(base, addr, count, dtype):
        off = addr - base.data_ptr()
        return base[off:off + count * torch.empty(0, dtype=dtype).element_size()].view(dtype)

    print("  n_contrib equal", bool(torch.equal(view(ia.t, ptrs[3], W * H, torch.int32), st["n_contrib"])),
          " ranges equal", bool(torch.equal(view(ia.t, ptrs[5], 2 * T, torch.int32).view(T, 2), st["ranges"])))
    ptrs = (C.c_void_p * 4)()
    L.ref_decode_binning(ba.t.data_ptr(), R, ptrs)
    print("  point_list equal", bool(torch.equal(view(ba.t, ptrs[0], R, torch.int32), st["point_list"])),
          " sorted keys equal", bool(torch.equal(view(ba.t, ptrs[2], R, torch.int64), st["sorted_keys"])))
    for name, a, b in (("color", r[2], rr[2]), ("depth", r[1], rr[1]), ("objects", r[3], rr[3])):
        print(f"  {name} max abs diff {float((a - b).abs().max()):.2e}")
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dobjects", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")
    for n_, a in zip(names, gr):
        b = gg[n_]
        print(f"  {n_:14s} rel err {float((a - b.view(a.shape)).abs().max() / b.abs().max()):.2e}")
