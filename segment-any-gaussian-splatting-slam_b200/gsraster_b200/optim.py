"""Fused activation + Adam around the rasterizer (csrc/optim.cu through the C ABI).

Mirrors the part of the reference's GaussianModel that sits on either side of the rasterizer call:
parameter groups and learning rates of `training_setup` (scene/gaussian_model.py:240-257, arguments/__init__.py:74-81),
the activations of `get_scaling / get_rotation / get_opacity / get_features` (:109-133) and
`torch.optim.Adam(l, lr=0.0, eps=1e-15)`. There is no CPU fallback: tensors must be CUDA tensors.
"""
import ctypes as C
from typing import Dict, Optional

import torch

from . import capi

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation", "objects")
# reference defaults: arguments/__init__.py:74-81 (position_lr_init * spatial_lr_scale for xyz, feature_lr / 20 for f_rest)
DEFAULT_LR = dict(xyz=0.0000016, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05, scaling=0.005, rotation=0.001,
                  objects=0.0025)


class AdamTensors(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                [g for g in GROUPS] + ["m_" + g for g in GROUPS] + ["v_" + g for g in GROUPS] +
                ["g_means3D", "g_sh", "g_objects", "g_opacity", "g_scales", "g_rotations",
                 "a_shs", "a_opacities", "a_scales", "a_rotations"]]


def _bind():
    L = capi.lib()
    if not getattr(L, "_optim_bound", False):
        L.gsr_activate.argtypes = [C.POINTER(AdamTensors), C.c_int, C.c_int, C.c_void_p]
        L.gsr_adam_step.argtypes = [C.POINTER(AdamTensors), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_float,
                                    C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.gsr_adam_step_ex.argtypes = L.gsr_adam_step.argtypes[:-1] + [C.c_int, C.c_void_p]
        L._optim_bound = True
    return L


class FusedGaussianAdam:
    """Raw parameters + Adam moments + the activated tensors the rasterizer reads.

    params: dict of CUDA fp32 tensors xyz (P,3), f_dc (P,1,3), f_rest (P,M-1,3), opacity (P,1), scaling (P,3),
    rotation (P,4), objects (P,1,F) or (P,F) -- the reference's `_xyz, _features_dc, ...` (they are updated in place).
    """

    def __init__(self, params: Dict[str, torch.Tensor], lr: Optional[Dict[str, float]] = None, betas=(0.9, 0.999),
                 eps: float = 1e-15):
        for k in GROUPS:
            t = params[k]
            if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                raise RuntimeError(f"gsraster_b200.optim: {k} must be a contiguous CUDA float32 tensor (no CPU fallback)")
        self.p = {k: params[k] for k in GROUPS}
        self.P = self.p["xyz"].shape[0]
        self.M = 1 + self.p["f_rest"].shape[1]
        self.F = self.p["objects"].numel() // max(self.P, 1)
        self.m = {k: torch.zeros_like(v) for k, v in self.p.items()}
        self.v = {k: torch.zeros_like(v) for k, v in self.p.items()}
        self.lr = dict(DEFAULT_LR)
        if lr:
            self.lr.update(lr)
        self.betas, self.eps, self.step_count = betas, eps, 0
        dev = self.p["xyz"].device
        o = dict(device=dev, dtype=torch.float32)
        self.shs = torch.empty((self.P, self.M, 3), **o)
        self.opacities = torch.empty((self.P, 1), **o)
        self.scales = torch.empty((self.P, 3), **o)
        self.rotations = torch.empty((self.P, 4), **o)
        self.activate()

    def _tensors(self, grads=None):
        t = AdamTensors()
        ptr = lambda x: x.data_ptr() if x is not None and x.numel() else None
        for k in GROUPS:
            setattr(t, k, ptr(self.p[k]))
            setattr(t, "m_" + k, ptr(self.m[k]))
            setattr(t, "v_" + k, ptr(self.v[k]))
        if grads is not None:
            for f, k in (("g_means3D", "means3D"), ("g_sh", "sh"), ("g_objects", "objects"), ("g_opacity", "opacity"),
                         ("g_scales", "scales"), ("g_rotations", "rotations")):
                g = grads[k]
                if not g.is_cuda or g.dtype != torch.float32 or not g.is_contiguous():
                    raise RuntimeError(f"gsraster_b200.optim: gradient {k} must be a contiguous CUDA float32 tensor")
                setattr(t, f, g.data_ptr())
        t.a_shs, t.a_opacities = self.shs.data_ptr(), self.opacities.data_ptr()
        t.a_scales, t.a_rotations = self.scales.data_ptr(), self.rotations.data_ptr()
        return t

    def activate(self):
        """shs / opacities / scales / rotations of all Gaussians from the raw parameters."""
        t = self._tensors()
        with torch.cuda.device(self.shs.device):
            capi._check(_bind().gsr_activate(C.byref(t), self.P, self.M, capi._stream_ptr()))

    def rasterizer_inputs(self):
        """-> dict(means3D, shs, sh_objs, opacities, scales, rotations) for rasterize_gaussians."""
        return dict(means3D=self.p["xyz"], shs=self.shs, sh_objs=self.p["objects"].view(self.P, self.F),
                    opacities=self.opacities, scales=self.scales, rotations=self.rotations)

    def step(self, grads: Dict[str, torch.Tensor], idx: Optional[torch.Tensor] = None,
             count: Optional[torch.Tensor] = None, dense_rows: bool = False):
        """One Adam step. grads: means3D, sh, objects, opacity, scales, rotations (the rasterizer's gradients w.r.t.
        the activated values; mapping.GradBucket.views or the compact views of a SparseGradExchange).
        idx/count None: dense (= torch.optim.Adam on all P rows). idx (int32, sorted) + count (int32 device scalar, or
        None = all of idx): sparse, only those Gaussians are stepped. dense_rows (sparse only): grads are the P-row
        outputs of a sparse-row backward (capi.Ext(sparse_rows=True)): Gaussian idx[r] takes its own row idx[r]."""
        self.step_count += 1
        t = self._tensors(grads)
        lr = (C.c_float * 7)(*[self.lr[k] for k in GROUPS])
        cap = 0
        if idx is not None:
            if idx.dtype != torch.int32 or not idx.is_cuda:
                raise RuntimeError("gsraster_b200.optim: idx must be a CUDA int32 tensor")
            cap = int(idx.numel())
            if count is None:
                count = torch.full((1,), cap, dtype=torch.int32, device=idx.device)
        with torch.cuda.device(self.shs.device):
            capi._check(_bind().gsr_adam_step_ex(C.byref(t), self.P, self.M, self.F, lr, self.betas[0], self.betas[1],
                                                 self.eps, self.step_count, None if idx is None else idx.data_ptr(),
                                                 None if idx is None else count.data_ptr(), cap,
                                                 1 if (dense_rows and idx is not None) else 0, capi._stream_ptr()))
