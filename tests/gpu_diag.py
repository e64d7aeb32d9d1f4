"""GPU diagnostic (not a pytest): product vs unmodified reference vs CPU oracle on seeded scenes.
Usage: python tests/gpu_diag.py [P] [W] [H] [deg]"""
import sys
import time

import numpy as np
import torch

from helpers import GRAD_KEYS, grad_rel_err, have_ref, npy, run_mine, run_oracle, run_ref, scenes


def cmp_exact(name, a, b, mask=None):
    a, b = npy(a), npy(b)
    if mask is not None:
        a, b = a[mask], b[mask]
    if a.dtype.kind == "f":
        a, b = a.view(np.uint32), b.view(np.uint32)
    bad = int((a != b).sum())
    print(f"  [exact] {name:14s} mismatches {bad} / {a.size}")
    return bad


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 640
    H = int(sys.argv[3]) if len(sys.argv) > 3 else 512
    deg = int(sys.argv[4]) if len(sys.argv) > 4 else 3
    do_oracle = P <= 50000
    sc = scenes.make_scene(P, sh_degree=deg, F=16, seed=1235, W=W, H=H)
    cam = scenes.make_camera(W, H, 0)
    grads = scenes.make_upstream_grads(W, H, 16)
    print(f"== scene P={P} {W}x{H} deg={deg}")
    mine = run_mine(sc, cam, grads, debug=True)
    print("  mine R", mine["R"], "launches fwd/bwd", mine["launches_fwd"], mine["launches_bwd"])
    if have_ref():
        ref = run_ref(sc, cam, grads)
        print("  ref  R", ref["R"])
        vis = npy(ref["radii"]) > 0
        cmp_exact("radii", mine["radii"], ref["radii"])
        cmp_exact("tiles_touched", mine["tiles_touched"], ref["tiles_touched"])
        cmp_exact("is_used", mine["is_used"], ref["is_used"])
        cmp_exact("depths", mine["depths"], ref["depths"], vis)
        cmp_exact("means2D", mine["means2D"], ref["means2D"], vis)
        cmp_exact("conic_opacity", mine["conic_opacity"], ref["conic_opacity"], vis)
        cmp_exact("rgb", mine["rgb"], ref["rgb"], vis)
        cmp_exact("clamped", mine["clamped"], ref["clamped"], vis)
        if mine["R"] == ref["R"] and ref["R"] > 0:
            cmp_exact("point_list", mine["point_list"], ref["point_list"])
            cmp_exact("sorted_keys", mine["sorted_keys"], ref["sorted_keys"])
        cmp_exact("ranges", mine["ranges"], ref["ranges"])
        cmp_exact("n_contrib", mine["n_contrib"], ref["n_contrib"])
        cmp_exact("final_T", mine["final_T"], ref["final_T"])
        for k in ("color", "depth", "objects"):
            print(f"  [1e-4 ] {k:14s} max abs diff {float((mine[k] - ref[k]).abs().max()):.3e}")
        for k in GRAD_KEYS:
            print(f"  [1e-3 ] {k:14s} rel err vs ref {grad_rel_err(mine[k], ref[k]):.3e}   max|ref| {float(ref[k].abs().max()):.3e}")
    if do_oracle:
        orc = run_oracle(sc, cam, grads)
        tgt = ref if have_ref() else mine
        print("  oracle R", orc["R"], "vs", "ref" if have_ref() else "mine")
        for k in ("radii", "tiles_touched", "n_contrib"):
            print(f"  [orc  ] {k:14s} mismatches {(npy(tgt[k]).astype(np.int64) != orc[k].astype(np.int64)).sum()}")
        for k in ("color", "depth", "objects"):
            print(f"  [orc  ] {k:14s} max abs diff {np.abs(npy(tgt[k]) - orc[k]).max():.3e}")
        for k in GRAD_KEYS:
            print(f"  [orc  ] {k:14s} rel err {grad_rel_err(orc[k], tgt[k]):.3e}")
    # quick timing (events, warm)
    from gsraster_b200 import capi
    from helpers import to_torch
    d = to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]

    def step():
        R, depth, color, objects, radii, is_used, gb, bb, ib = capi.rasterize_gaussians(
            d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, d["viewmatrix"],
            d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"], d["sh_objs"], deg, d["campos"], False, False)
        capi.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                          d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, gd, gc, d["shs"],
                                          go, d["sh_objs"], deg, d["campos"], gb, R, bb, ib, False)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    print(f"  mine fwd+bwd {1e3 * (time.time() - t0) / 10:.3f} ms/iter")
    if have_ref():
        for _ in range(2):
            run_ref(sc, cam, grads)
        torch.cuda.synchronize()


if __name__ == "__main__":
    main()
