"""GPU diagnostic (not a pytest): gradient error statistics of the product against the live reference build, per tensor:
max-norm, relative L2, and the per-element relative error over the entries with |ref| >= 1e-4 max|ref| -- next to the
reference's own run-to-run spread (its float atomics make it non-deterministic). Used to set the bars of
tests/test_parity_gpu.py::_check_grads.  Usage: python tests/gpu_gradstats.py [case ...]"""
import sys

import numpy as np

import helpers as h
from helpers import scenes

CASES = {
    "c1_10k": (10_000, 640, 512, 0, 0, 1.0, 16),
    "c2_500k": (500_000, 1280, 1024, 3, 0, 1.0, 16),
    "trained_200k": (200_000, 1280, 1024, 3, 2, 3.0, 16),
    "odd_30k": (30_000, 1000, 700, 2, 1, 2.0, 16),
    "headline_1M": (1_000_000, 1280, 1024, 3, 0, 1.0, 16),
}


def stats(a, b1, b2):
    a, b1, b2 = [h.npy(x).astype(np.float64).reshape(-1) for x in (a, b1, b2)]
    ref = 0.5 * (b1 + b2)
    mx = np.abs(ref).max()
    if mx == 0:
        return None
    sel = np.abs(ref) >= 1e-4 * mx
    err = np.abs(a - ref)
    spread = np.abs(b1 - b2)
    rel = err[sel] / np.abs(ref[sel])
    rel_ref = spread[sel] / np.abs(ref[sel])
    excess = err[sel] - (1e-2 * np.abs(ref[sel]) + 2.0 * spread[sel] + spread.max())
    return dict(maxnorm=err.max() / mx, ref_maxnorm=spread.max() / mx, l2=np.linalg.norm(a - ref) / np.linalg.norm(ref),
                ref_l2=np.linalg.norm(b1 - b2) / np.linalg.norm(ref), n=int(sel.sum()), elem_max=rel.max(),
                elem_p999=np.quantile(rel, 0.999), ref_elem_max=rel_ref.max(), viol=int((excess > 0).sum()),
                nz_mismatch=int(((a != 0) != (ref != 0)).sum()))


for name in (sys.argv[1:] or CASES):
    P, W, H, deg, view, mul, F = CASES[name]
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=1234 + P % 97 + deg, W=W, H=H, scale_mul=mul)
    cam = scenes.make_camera(W, H, view)
    grads = scenes.make_upstream_grads(W, H, F)
    mine = h.run_mine(sc, cam, grads)
    r1, r2 = h.run_ref(sc, cam, grads), h.run_ref(sc, cam, grads)
    print("== %s R=%d" % (name, mine["R"]))
    for k in h.GRAD_KEYS:
        s = stats(mine[k], r1[k], r2[k])
        if s is None:
            continue
        print("  %-14s max %.2e (ref %.2e)  L2 %.2e (ref %.2e)  elem[n=%d] max %.2e p99.9 %.2e (ref max %.2e) viol %d  nz-mismatch %d"
              % (k, s["maxnorm"], s["ref_maxnorm"], s["l2"], s["ref_l2"], s["n"], s["elem_max"], s["elem_p999"],
                 s["ref_elem_max"], s["viol"], s["nz_mismatch"]))
