"""Fused photometric loss (csrc/loss.cu through the C ABI).

Same value as the reference's composition (utils/loss_utils.py:17-20 `l1_loss`, :36-69 `ssim`, both with the `gt != 0`
mask; scripts/gaussian_splatting.py:805-807): loss_rgb = (1 - lambda_dssim) * Ll1 + lambda_dssim * (1 - L_ssim),
plus d loss_rgb / d image in the same pass. No CPU fallback.
"""
import ctypes as C

import torch

from . import capi


def _bind():
    L = capi.lib()
    if not getattr(L, "_loss_bound", False):
        L.gsr_photometric_loss_scratch_bytes.restype = C.c_size_t
        L.gsr_photometric_loss_scratch_bytes.argtypes = [C.c_int, C.c_int]
        L.gsr_photometric_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p]
        L._loss_bound = True
    return L


class PhotometricLoss:
    """Holds the scratch for one image size. __call__(image, gt) -> (out3, dL_dimage); out3 = device tensor
    [loss_rgb, Ll1, L_ssim] (no host synchronisation)."""

    def __init__(self, H, W, lambda_dssim=0.2, device="cuda"):
        self.H, self.W, self.lam = int(H), int(W), float(lambda_dssim)
        n = _bind().gsr_photometric_loss_scratch_bytes(self.H, self.W)
        self.scratch = torch.empty(int(n) + 256, dtype=torch.uint8, device=device)

    def __call__(self, image, gt, need_grad=True, out=None, grad_out=None):
        for t in (image, gt):
            if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or tuple(t.shape) != (3, self.H, self.W):
                raise RuntimeError("gsraster_b200.losses: expected contiguous CUDA float32 (3, H, W) tensors (no CPU fallback)")
        out3 = out if out is not None else torch.empty(3, dtype=torch.float32, device=image.device)
        grad = None
        if need_grad:
            grad = grad_out if grad_out is not None else torch.empty_like(image)
        with torch.cuda.device(image.device):
            capi._check(_bind().gsr_photometric_loss(image.data_ptr(), gt.data_ptr(), self.H, self.W, self.lam,
                                                     None if grad is None else grad.data_ptr(), out3.data_ptr(),
                                                     self.scratch.data_ptr(), capi._stream_ptr()))
        return out3, grad
