"""Finite-difference machinery for the camera-pose gradient (SURVEY.md 8c last row, 11): a 6-dof twist applied
consistently to viewmatrix, projmatrix and campos, the loss evaluated by the DOUBLE-precision build of the CPU oracle
(oracle/oracle.cpp with -DORC_REAL=double: the reference's forward pass restated, smooth enough for steps of 1e-7),
central differences with a Richardson check of the step. Test infrastructure only."""
import numpy as np

from gsraster_b200 import scenes
from oracle import pyoracle as po


def twist_matrix(xi):
    """xi = (omega, t) -> the 4x4 generator G^ with exp(eps G^) the rigid motion applied on the camera side."""
    om, t = np.asarray(xi[:3], np.float64), np.asarray(xi[3:], np.float64)
    G = np.zeros((4, 4))
    G[:3, :3] = np.array([[0, -om[2], om[1]], [om[2], 0, -om[0]], [-om[1], om[0], 0]])
    G[:3, 3] = t
    return G


def expm_se3(G):
    """Matrix exponential of a 4x4 twist generator (series: the steps are tiny, 12 terms are exact to double)."""
    out, term = np.eye(4), np.eye(4)
    for k in range(1, 13):
        term = term @ G / k
        out = out + term
    return out


class PoseFD:
    """L(eps) = <color, gc> + <depth, gd> + <objects, go> of the double-precision oracle under V(eps) = exp(eps G^) V0."""

    def __init__(self, scene, cam, grads):
        self.scene, self.cam0 = scene, cam
        self.gd, self.gc, self.go = [np.asarray(g, np.float64) for g in grads]
        # the fp32 camera arrays, taken exactly, are the base point: V0, PV0, c0 (math matrices: flat index 4c + r)
        self.V0 = cam.viewmatrix.T.astype(np.float64)
        self.PV0 = cam.projmatrix.T.astype(np.float64)
        self.c0 = cam.campos.astype(np.float64)
        self.V0inv = np.linalg.inv(self.V0)
        self.Proj = self.PV0 @ self.V0inv          # so that Proj @ V0 == PV0 exactly as given

    def camera(self, G, eps):
        V = expm_se3(eps * G) @ self.V0
        c = self.c0 + (np.linalg.inv(V) - self.V0inv)[:3, 3]
        return scenes.Camera(self.cam0.W, self.cam0.H, self.cam0.tanfovx, self.cam0.tanfovy,
                             np.ascontiguousarray(V.T), np.ascontiguousarray((self.Proj @ V).T), c)

    def loss(self, G, eps):
        st = po.forward(po.Problem(self.scene, self.camera(G, eps), f64=True))
        return float((st["out_color"] * self.gc).sum() + (st["out_depth"] * self.gd).sum() +
                     (st["out_objects"] * self.go).sum())

    def derivative(self, xi, steps=(2e-7, 5e-8, 1.3e-8, 3e-9), rtol=2e-5):
        """-> (Richardson-extrapolated dL/d eps at 0, |D(h) - D(h/2)|) for the first step size h whose two central
        differences agree to rtol: a discrete event of the rasterizer (a pixel crossing the alpha >= 1/255 or
        T < 1e-4 threshold, a radius or tile rect changing, two depths swapping) inside the stencil shows up as a large
        disagreement of the two step sizes, and a smaller stencil steps over it."""
        G = twist_matrix(xi)
        D = lambda s: (self.loss(G, s) - self.loss(G, -s)) / (2 * s)
        best = None
        for h in steps:
            d1, d2 = D(h), D(h / 2)
            cand = ((4 * d2 - d1) / 3.0, abs(d1 - d2))
            if best is None or cand[1] < best[1]:
                best = cand
            if cand[1] <= rtol * max(abs(d1), abs(d2), 1e-9):
                return cand
        return best

    def project(self, xi, dV, dPV, dC):
        """Analytic directional derivative from the three camera gradients (math-matrix layout dL/dM[r][c])."""
        G = twist_matrix(xi)
        dVde = G @ self.V0
        dPVde = self.Proj @ dVde
        dcde = (-self.V0inv @ G)[:3, 3]
        return float((dV * dVde).sum() + (dPV * dPVde).sum() + (dC * dcde).sum())


def drop_clamped(scene, cam, margin=0.97):
    """Keeps the Gaussians whose camera-space direction lies inside the reference's +-1.3 tan(fov) clamp of the EWA
    Jacobian (forward.cu:82-87). Outside it the reference's backward pass treats the clamped t.x = lim * t.z as
    independent of t.z (backward.cu:170-178, 282-284: a 0/1 mask on dL/dt.x only) -- a modelling shortcut of the
    reference that the product reproduces, so analytic and numerical derivatives differ there by construction."""
    V = cam.viewmatrix.T.astype(np.float64)
    t = scene.means3D.astype(np.float64) @ V[:3, :3].T + V[:3, 3]
    keep = (t[:, 2] > 0.2) & (np.abs(t[:, 0] / t[:, 2]) < margin * 1.3 * cam.tanfovx) & \
           (np.abs(t[:, 1] / t[:, 2]) < margin * 1.3 * cam.tanfovy)
    return scenes.Scene(scene.means3D[keep], scene.scales[keep], scene.rotations[keep], scene.opacities[keep],
                        scene.shs[keep], scene.sh_objs[keep], scene.sh_degree, scene.bg, scene.scale_modifier)
