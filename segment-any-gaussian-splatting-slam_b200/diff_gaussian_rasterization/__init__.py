"""Drop-in for the reference's `diff_gaussian_rasterization` Python module
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py): same names, same
argument order, same outputs, so `render_4` (gaussian_renderer/__init__.py:325-480) runs unchanged.

    GaussianRasterizationSettings   12-field NamedTuple            (reference :166-178)
    GaussianRasterizer(settings)    .forward(...) -> (depth, color, radii, objects, is_used), .markVisible(positions)
    rasterize_gaussians(...)        autograd entry point            (reference :21-44)

Underneath sits the LibTorch binding lib/gsraster_b200_torch.so (or gsraster_b200.capi) -> lib/libgsraster_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/gsraster_b200.h). There is no CPU fallback: CPU tensors or a missing library raise.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from gsraster_b200 import torch_ext as _torch_ext

# `_C` is the C++/LibTorch binding (csrc/torch_binding.cpp), like the reference's pybind `_C`; the ctypes binding
# gsraster_b200.capi has the same functions and is used when the LibTorch shim has not been built.
if _torch_ext.available():
    _C = _torch_ext.load()
else:
    from gsraster_b200 import capi as _C


def cpu_deep_copy_tuple(input_tuple):
    """Reference :17-19: CPU copies of the tensor arguments, taken BEFORE the call so that a crash dump shows the inputs
    as they were."""
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def rasterize_gaussians(means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, backward_ext=None):
    """`backward_ext` (extension, not in the reference): a gsraster_b200.capi.Ext handed to the BACKWARD call of this
    rasterization -- touched flags / sparse rows for a SparseGradExchange, timers -- wherever autograd runs it (torch
    executes backward nodes on its own worker thread, so the options travel with the call, not with the thread)."""
    return _RasterizeGaussians.apply(means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, backward_ext)


class _RasterizeGaussians(torch.autograd.Function):
    """Reference :46-164. Gradient slots: (means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales,
    rotations, cov3Ds_precomp, None)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, sh_objs, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, backward_ext=None):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh, sh_objs,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        if rs.debug:  # reference :87-94: keep CPU copies of the arguments, dump them if the extension throws
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = _C.rasterize_gaussians(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            out = _C.rasterize_gaussians(*args)
        num_rendered, depth, color, objects, radii, is_used, geomBuffer, binningBuffer, imgBuffer = out
        ctx.backward_ext = backward_ext
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, sh_objs,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii, is_used)
        return depth, color, radii, objects, is_used

    @staticmethod
    def backward(ctx, grad_depth, grad_out_color, grad_radii, grad_out_objects, grad_is_used):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, sh_objs, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        H, W = rs.image_height, rs.image_width
        z = lambda g, c: g if g is not None else torch.zeros((c, H, W), dtype=torch.float32, device=means3D.device)
        F = sh_objs.numel() // max(means3D.size(0), 1)
        grad_depth, grad_out_color, grad_out_objects = z(grad_depth, 1), z(grad_out_color, 3), z(grad_out_objects, F)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_depth, grad_out_color, sh, grad_out_objects, sh_objs,
                rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        if ctx.backward_ext is not None:  # per-call extensions exist on the ctypes binding only
            from gsraster_b200 import capi as _capi
            call = lambda: _capi.rasterize_gaussians_backward(*args, ext=ctx.backward_ext)
        else:
            call = lambda: _C.rasterize_gaussians_backward(*args)
        if rs.debug:  # reference :140-147
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                grads = call()
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            grads = call()
        (grad_means2D, grad_colors_precomp, grad_objects, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = grads
        return (grad_means3D, grad_means2D, grad_sh, grad_objects.unsqueeze(1), grad_colors_precomp, grad_opacities,
                grad_scales, grad_rotations, grad_cov3Ds_precomp, None, None)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings, backward_ext=None):
        super().__init__()
        self.raster_settings = raster_settings
        self.backward_ext = backward_ext   # extension: see rasterize_gaussians

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, sh_objs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        # same two checks and messages as the reference (:200-204)
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        e = torch.Tensor([])  # "absent" == 0-element tensor -> NULL pointer (reference :206-219)
        shs = e if shs is None else shs
        colors_precomp = e if colors_precomp is None else colors_precomp
        scales = e if scales is None else scales
        rotations = e if rotations is None else rotations
        cov3D_precomp = e if cov3D_precomp is None else cov3D_precomp
        sh_objs = e if sh_objs is None else sh_objs
        return rasterize_gaussians(means3D, means2D, shs, sh_objs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings, self.backward_ext)
