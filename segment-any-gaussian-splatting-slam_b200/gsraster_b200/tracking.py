"""Camera-pose gradient (tracking path, BASELINE config 5) -- an EXTENSION with no reference equivalent: the reference
takes poses from FAST-LIVO2 odometry and its rasterizer returns no camera gradients.

`rasterize_with_camera_grads` is an autograd function whose differentiable inputs are the three camera tensors of
GaussianRasterizationSettings (viewmatrix, projmatrix, campos: same layouts). Build them from a pose parametrisation
in PyTorch (e.g. a 6-dof twist) and autograd chains through. Gaussian parameters are treated as constants here."""
import torch

from . import capi


class _RasterizeCamera(torch.autograd.Function):
    @staticmethod
    def forward(ctx, viewmatrix, projmatrix, campos, means3D, sh, sh_objs, opacities, scales, rotations, rs, tracking):
        e = torch.empty(0, device=means3D.device)
        (R, depth, color, objects, radii, is_used, gb, bb, ib) = capi.rasterize_gaussians(
            rs.bg, means3D, e, opacities, scales, rotations, rs.scale_modifier, e, viewmatrix, projmatrix, rs.tanfovx,
            rs.tanfovy, rs.image_height, rs.image_width, sh, sh_objs, rs.sh_degree, campos, rs.prefiltered, rs.debug)
        ctx.rs, ctx.R, ctx.tracking = rs, R, bool(tracking)
        ctx.save_for_backward(viewmatrix, projmatrix, campos, means3D, sh, sh_objs, scales, rotations, radii, gb, bb, ib)
        ctx.mark_non_differentiable(radii, is_used)
        return depth, color, radii, objects, is_used

    @staticmethod
    def backward(ctx, g_depth, g_color, _g_radii, g_objects, _g_used):
        rs = ctx.rs
        viewmatrix, projmatrix, campos, means3D, sh, sh_objs, scales, rotations, radii, gb, bb, ib = ctx.saved_tensors
        H, W = rs.image_height, rs.image_width
        F = sh_objs.numel() // max(means3D.size(0), 1)
        z = lambda g, c: g if g is not None else torch.zeros((c, H, W), dtype=torch.float32, device=means3D.device)
        e = torch.empty(0, device=means3D.device)
        # tracking mode (default): the Gaussians are constants here, so the backward blend skips the feature-channel
        # reductions and the per-Gaussian backward writes none of its eleven tensors (gsr_ext::tracking)
        ext = capi.Ext(tracking=ctx.tracking)
        capi.rasterize_gaussians_backward(rs.bg, means3D, radii, e, scales, rotations, rs.scale_modifier, e, viewmatrix,
                                          projmatrix, rs.tanfovx, rs.tanfovy, z(g_depth, 1), z(g_color, 3), sh,
                                          z(g_objects, F), sh_objs, rs.sh_degree, campos, gb, ctx.R, bb, ib, rs.debug,
                                          camera_grads=True, ext=ext)
        dV, dPV, dC = ext.camera_grads
        return dV, dPV, dC, None, None, None, None, None, None, None, None


def rasterize_with_camera_grads(viewmatrix, projmatrix, campos, means3D, sh, sh_objs, opacities, scales, rotations,
                                raster_settings, tracking=True):
    """-> (depth, color, radii, objects, is_used); gradients flow to viewmatrix (4,4), projmatrix (4,4), campos (3,).
    The camera fields of `raster_settings` are ignored in favour of the tensor arguments. tracking=False runs the full
    mapping backward underneath (same camera gradients, for cross-checks)."""
    return _RasterizeCamera.apply(viewmatrix, projmatrix, campos, means3D, sh, sh_objs, opacities, scales, rotations,
                                  raster_settings, tracking)
