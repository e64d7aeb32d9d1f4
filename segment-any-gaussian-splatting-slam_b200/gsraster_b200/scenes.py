"""Seeded synthetic scenes for the rasterizer hot path.

Camera: FAST-LIVO2 intrinsics used by the reference driver
(src/gaussian_splatting/scripts/gaussian_splatting.py:174-177), znear/zfar and matrix conventions of
SharedCam (src/gaussian_splatting/gaussian_splatting/scene/shared_objs.py:232-243) and
getProjectionMatrix (.../utils/graphics_utils.py:51-71). Everything is numpy/float32 on the host;
callers move it to the device. No torch import here so the oracle-side tests stay light.
"""
import math
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

FX_FULL, FY_FULL = 1293.56944, 1293.3155  # at 1280x1024


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: np.ndarray   # (4,4) float32 = V^T  (flat index 4*c+r == V[r][c])
    projmatrix: np.ndarray   # (4,4) float32 = (Proj @ V)^T
    campos: np.ndarray       # (3,)


@dataclass
class Scene:
    means3D: np.ndarray
    scales: np.ndarray
    rotations: np.ndarray    # (P,4) xyzw, unit
    opacities: np.ndarray    # (P,1) activated
    shs: np.ndarray          # (P,M,3)
    sh_objs: np.ndarray      # (P,1,F)
    sh_degree: int
    bg: np.ndarray = field(default_factory=lambda: np.zeros(3, np.float32))
    scale_modifier: float = 1.0

    @property
    def P(self):
        return self.means3D.shape[0]


def projection_matrix(znear, zfar, tanx, tany):
    P = np.zeros((4, 4), np.float64)
    top, right = tany * znear, tanx * znear
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W, H, view_index=0, znear=0.01, zfar=100.0):
    """View k: yaw k*5 degrees about +y and a 0.25 m lateral step per view."""
    sx = W / 1280.0 if W != 1920 else 1.0
    sy = H / 1024.0 if W != 1920 else 1.0
    fx, fy = FX_FULL * sx, FY_FULL * sy
    tanx, tany = W / (2 * fx), H / (2 * fy)
    a = math.radians(5.0 * view_index)
    R = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])
    t = np.array([-0.25 * view_index, 0.0, 0.0])
    V = np.eye(4)
    V[:3, :3] = R
    V[:3, 3] = t
    Pm = projection_matrix(znear, zfar, tanx, tany)
    full = Pm @ V
    campos = np.linalg.inv(V)[:3, 3]
    return Camera(W, H, float(np.float32(tanx)), float(np.float32(tany)),
                  np.ascontiguousarray(V.T.astype(np.float32)),
                  np.ascontiguousarray(full.T.astype(np.float32)),
                  campos.astype(np.float32))


def make_scene(P, sh_degree=3, F=16, seed=1234, W=1280, H=1024, scale_mul=1.0, behind_frac=0.05):
    """'lidar-init' Gaussians in a frustum 1.3x wider than the FoV of view 0."""
    rng = np.random.default_rng(seed)
    cam = make_camera(W, H, 0)
    u = rng.uniform(-1.3, 1.3, P) * cam.tanfovx
    v = rng.uniform(-1.3, 1.3, P) * cam.tanfovy
    z = np.exp(rng.uniform(math.log(0.5), math.log(40.0), P))
    behind = rng.uniform(size=P) < behind_frac
    z = np.where(behind, -z, z)
    means = np.stack([u * np.abs(z), v * np.abs(z), z], 1).astype(np.float32)
    scales = (np.exp(rng.normal(math.log(0.02), 0.5, (P, 3))) * scale_mul).astype(np.float32)
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(-1.0, 1.5, (P, 1))))).astype(np.float32)
    M = (sh_degree + 1) ** 2
    shs = np.zeros((P, M, 3), np.float32)
    shs[:, 0, :] = (rng.uniform(size=(P, 3)) - 0.5) / 0.28209
    if M > 1:
        shs[:, 1:, :] = rng.normal(0, 0.05, (P, M - 1, 3))
    objs = ((rng.uniform(size=(P, 1, F)) - 0.5) / 0.28209).astype(np.float32)
    return Scene(means, scales, q.astype(np.float32), opac, shs, objs, sh_degree)


def make_trajectory_scene(P, num_views, sh_degree=3, F=16, seed=1234, W=1280, H=1024, scale_mul=1.0, behind_frac=0.05,
                          opacity_mean=-1.0):
    """'lidar-init' Gaussians over the UNION of the frusta of a keyframe trajectory (views 0 .. num_views-1 of
    make_camera: 5 degrees of yaw and 0.25 m of lateral motion per keyframe): every Gaussian is drawn with the
    SURVEY 8d sampler in the camera frame of a uniformly chosen keyframe and moved to the world frame, so every
    keyframe looks at the same kind of population (a SLAM map covers what the sensor has swept, not one frustum).
    num_views = 1 is make_scene. Same attribute distributions as make_scene."""
    rng = np.random.default_rng(seed)
    cam0 = make_camera(W, H, 0)
    u = rng.uniform(-1.3, 1.3, P) * cam0.tanfovx
    v = rng.uniform(-1.3, 1.3, P) * cam0.tanfovy
    z = np.exp(rng.uniform(math.log(0.5), math.log(40.0), P))
    behind = rng.uniform(size=P) < behind_frac
    z = np.where(behind, -z, z)
    local = np.stack([u * np.abs(z), v * np.abs(z), z], 1)
    owner = rng.integers(0, num_views, P)
    means = np.empty((P, 3), np.float64)
    for k in range(num_views):
        sel = owner == k
        if not sel.any():
            continue
        Vk = make_camera(W, H, k).viewmatrix.T.astype(np.float64)       # world -> camera k
        Rk, tk = Vk[:3, :3], Vk[:3, 3]
        means[sel] = (local[sel] - tk) @ Rk                              # p_world = R^T (p_cam - t)
    scales = (np.exp(rng.normal(math.log(0.02), 0.5, (P, 3))) * scale_mul).astype(np.float32)
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(opacity_mean, 1.5, (P, 1))))).astype(np.float32)
    M = (sh_degree + 1) ** 2
    shs = np.zeros((P, M, 3), np.float32)
    shs[:, 0, :] = (rng.uniform(size=(P, 3)) - 0.5) / 0.28209
    if M > 1:
        shs[:, 1:, :] = rng.normal(0, 0.05, (P, M - 1, 3))
    objs = ((rng.uniform(size=(P, 1, F)) - 0.5) / 0.28209).astype(np.float32)
    return Scene(means.astype(np.float32), scales, q.astype(np.float32), opac, shs, objs, sh_degree)


# Named single-view workloads of bench.py's `workloads` block (both arms build them from here, so the scenes are
# identical): the SURVEY 8d scene of round 1, BASELINE config 2, the "trained-like" x3 scene, and a scene whose pixels
# do NOT saturate early (small, faint splats: most visible Gaussians receive a gradient).
WORKLOADS = {
    "survey8d_view0": dict(P=1_000_000, scale_mul=1.0, opacity_cap=None,
                           note="SURVEY 8d lidar-init scene in the frustum of view 0 (the round-1 headline)"),
    "c2_500k": dict(P=500_000, scale_mul=1.0, opacity_cap=None, note="BASELINE config 2: 500k Gaussians"),
    "trained_like_x3": dict(P=1_000_000, scale_mul=3.0, opacity_cap=None, note="SURVEY 8d 'trained-like': scales x3"),
    "non_saturating": dict(P=1_000_000, scale_mul=0.3, opacity_cap=0.1,
                           note="scales x0.3 and opacity <= 0.1: pixels do not saturate, most visible Gaussians are touched"),
}


def make_workload_scene(name, sh_degree=3, F=16, W=1280, H=1024, seed=1234):
    w = WORKLOADS[name]
    sc = make_scene(w["P"], sh_degree=sh_degree, F=F, seed=seed, W=W, H=H, scale_mul=w["scale_mul"])
    if w["opacity_cap"] is not None:
        sc.opacities = np.minimum(sc.opacities, np.float32(w["opacity_cap"])).astype(np.float32)
    return sc


def make_upstream_grads(W, H, F, seed=99):
    rng = np.random.default_rng(seed)
    n = float(H * W)
    return ((rng.normal(size=(1, H, W)) / n).astype(np.float32),
            (rng.normal(size=(3, H, W)) / n).astype(np.float32),
            (rng.normal(size=(F, H, W)) / n).astype(np.float32))


def make_points(P, seed=4321):
    """kNN input: same position sampler as make_scene."""
    return make_scene(P, 0, 1, seed).means3D
