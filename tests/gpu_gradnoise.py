"""GPU diagnostic (not a pytest): gradient agreement mine vs reference vs run-to-run noise of each
(float atomics make both non-deterministic). Usage: python tests/gpu_gradnoise.py P W H deg view mul"""
import sys

import numpy as np

import helpers as h
from helpers import scenes

P, W, H, deg, view = [int(x) for x in sys.argv[1:6]]
mul = float(sys.argv[6])
sc = scenes.make_scene(P, sh_degree=deg, F=16, seed=1234 + P % 97 + deg, W=W, H=H, scale_mul=mul)
cam = scenes.make_camera(W, H, view)
grads = scenes.make_upstream_grads(W, H, 16)
r1 = h.run_ref(sc, cam, grads)
r2 = h.run_ref(sc, cam, grads)
m1 = h.run_mine(sc, cam, grads)
m2 = h.run_mine(sc, cam, grads)
print(f"{'tensor':14s} {'ref-ref':>10s} {'mine-mine':>10s} {'mine-ref':>10s}   worst idx (mine-ref): |ref| there / max|ref|")
for k in h.GRAD_KEYS:
    a, b = h.npy(m1[k]).astype(np.float64), h.npy(r1[k]).astype(np.float64)
    d = np.abs(a - b).reshape(P, -1).max(1)
    i = int(d.argmax())
    print(f"{k:14s} {h.grad_rel_err(r1[k], r2[k]):10.2e} {h.grad_rel_err(m1[k], m2[k]):10.2e} {h.grad_rel_err(m1[k], r1[k]):10.2e}"
          f"   idx {i} {np.abs(b.reshape(P, -1)[i]).max() / np.abs(b).max():.3f} radius {int(h.npy(r1['radii'])[i])}"
          f" depth {float(h.npy(r1['depths'])[i]):.3f} tiles {int(h.npy(r1['tiles_touched'])[i])}")
