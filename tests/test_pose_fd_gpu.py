"""Camera-pose gradient (BASELINE config 5; an extension, the reference never differentiates w.r.t. the camera:
PARITY UNPINNED BY THE REFERENCE) pinned the way SURVEY.md 8c / 11 asks: central differences, in double precision, of
the reference's forward pass (CPU oracle, -DORC_REAL=double) over a 6-dof twist applied consistently to viewmatrix,
projmatrix and campos, at SH degree 3, with colour, depth AND feature upstream gradients, Richardson-checked steps --
against gsr_backward_pose projected on the twist, in full and in tracking mode. Reference lines the derivation rests on:
forward.cu:74-117 (EWA Jacobian), :216-218 (projection), :268 (depth), :25-27 (SH direction); backward.cu:251-293,
:404-417."""
import numpy as np
import pytest

import helpers as h
from helpers import scenes
from pose_fd import PoseFD, drop_clamped

pytestmark = pytest.mark.gpu

TWISTS = [(1, 0, 0, 0, 0, 0), (0, 1, 0, 0, 0, 0), (0, 0, 1, 0, 0, 0), (0, 0, 0, 1, 0, 0), (0, 0, 0, 0, 1, 0),
          (0, 0, 0, 0, 0, 1), (0.3, -0.5, 0.2, 0.4, 0.1, -0.6)]


def _camera_grads(sc, cam, grads, tracking):
    import torch
    from gsraster_b200 import capi
    d = h.to_torch(sc, cam)
    e = torch.empty(0, device="cuda")
    r = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                 d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, cam.H, cam.W, d["shs"],
                                 d["sh_objs"], sc.sh_degree, d["campos"], False, False)
    gd, gc, go = [torch.from_numpy(g).cuda() for g in grads]
    ext = capi.Ext(tracking=tracking)
    capi.rasterize_gaussians_backward(d["bg"], d["means3D"], r[4], e, d["scales"], d["rotations"], 1.0, e, d["viewmatrix"],
                                      d["projmatrix"], cam.tanfovx, cam.tanfovy, gd, gc, d["shs"], go, d["sh_objs"],
                                      sc.sh_degree, d["campos"], r[6], r[0], r[7], r[8], False, camera_grads=True, ext=ext)
    dV, dPV, dC = [t.double().cpu().numpy() for t in ext.camera_grads]
    return dV.T, dPV.T, dC, r     # math layout dL/dM[r][c] (the tensors are flat 4c + r like the inputs)


@pytest.mark.parametrize("deg,view,seed", [(3, 2, 101), (1, 0, 102)])
def test_pose_gradient_matches_central_differences_over_a_twist(deg, view, seed):
    W, H, P, F = 128, 96, 600, 16
    cam = scenes.make_camera(W, H, view)
    # (Gaussians outside the reference's +-1.3 tan(fov) clamp are left out: see pose_fd.drop_clamped)
    sc = drop_clamped(scenes.make_scene(P, sh_degree=deg, F=F, seed=seed, W=W, H=H, scale_mul=4.0, behind_frac=0.0), cam)
    sc.shs[:, 1:, :] *= 6.0           # make the view-dependent colour (the campos path) matter
    sc.bg = np.array([0.3, 0.1, 0.5], np.float32)
    grads = tuple(g * (W * H) for g in scenes.make_upstream_grads(W, H, F, seed=seed))   # O(1) upstream gradients
    fd = PoseFD(sc, cam, grads)
    full = _camera_grads(sc, cam, grads, tracking=False)
    track = _camera_grads(sc, cam, grads, tracking=True)
    # the product's forward pass is the oracle's (so both differentiate the same function)
    st = h.run_oracle(sc, cam)
    assert float(np.abs(h.npy(full[3][2]) - st["color"]).max()) < 1e-4
    ders = [fd.derivative(xi) for xi in TWISTS]
    scale = max(abs(d[0]) for d in ders)
    for xi, (d_fd, rich) in zip(TWISTS, ders):
        assert rich <= 1e-4 * scale, ("finite differences not converged (a discrete event inside the stencil?)", xi, rich)
        for name, (dV, dPV, dC, _) in (("full", full), ("tracking", track)):
            d_an = fd.project(xi, dV, dPV, dC)
            assert abs(d_an - d_fd) <= 2e-3 * max(abs(d_fd), abs(d_an)) + 2e-4 * scale, (name, xi, d_an, d_fd)
