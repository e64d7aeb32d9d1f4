"""GPU diagnostic (not a pytest): camera-pose gradient vs central differences, per upstream-gradient component."""
import numpy as np
import helpers as h
from helpers import scenes
from pose_fd import PoseFD, drop_clamped
from test_pose_fd_gpu import _camera_grads, TWISTS

for deg, view, seed in ((3, 2, 101), (0, 2, 101)):
    W, H, P, F = 128, 96, 600, 16
    cam = scenes.make_camera(W, H, view)
    sc = drop_clamped(scenes.make_scene(P, sh_degree=deg, F=F, seed=seed, W=W, H=H, scale_mul=4.0, behind_frac=0.0), cam)
    sc.shs[:, 1:, :] *= 6.0
    for bgv in ((0, 0, 0), (0.3, 0.1, 0.5)):
        sc.bg = np.array(bgv, np.float32)
        g0 = tuple(g * (W * H) for g in scenes.make_upstream_grads(W, H, F, seed=seed))
        for name, mask in (("color", (0, 1, 0)), ("depth", (1, 0, 0)), ("objects", (0, 0, 1))):
            grads = tuple(g * m for g, m in zip(g0, mask))
            fd = PoseFD(sc, cam, grads)
            dV, dPV, dC, _ = _camera_grads(sc, cam, grads, tracking=False)
            out = []
            for xi in TWISTS[:6]:
                d_fd, rich = fd.derivative(xi)
                d_an = fd.project(xi, dV, dPV, dC)
                out.append("%.1f/%.1f(%.1e)" % (d_an, d_fd, abs(d_an - d_fd) / max(abs(d_fd), 1e-9)))
            print("deg", deg, "bg", bgv, name, " ".join(out))
