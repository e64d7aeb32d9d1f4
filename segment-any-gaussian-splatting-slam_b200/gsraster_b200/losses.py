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


def _bind_feat():
    L = capi.lib()
    if not getattr(L, "_featloss_bound", False):
        L.gsr_feature_loss_scratch_bytes.restype = C.c_size_t
        L.gsr_feature_loss_scratch_bytes.argtypes = []
        L.gsr_feature_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L._featloss_bound = True
    return L


class FeatureLoss:
    """Segmentation-feature loss of the mapping step (csrc/featloss.cu): the reference's
    `logits = classifier(objects); loss_obj = cls_criterion(logits[None], gt[None]).squeeze().mean() / log(num_classes)`
    and `pred_obj = argmax(logits, 0)` (scripts/gaussian_splatting.py:215-216, 798-803) in one kernel that never writes
    the logits. `classifier` is a torch.nn.Conv2d(F, C, kernel_size=1) (or any object with .weight (C,F,1,1) / .bias).

    __call__(objects (F,H,W), gt (H,W) int64) -> (out2, dL_dobjects, pred); out2 = device tensor [loss_obj, mean CE]."""

    def __init__(self, classifier, device="cuda"):
        self.classifier = classifier
        n = _bind_feat().gsr_feature_loss_scratch_bytes()
        self.scratch = torch.empty(int(n) + 256, dtype=torch.uint8, device=device)

    def __call__(self, objects, gt, grad_weight=1.0, need_grad=True, need_pred=False, out=None, grad_out=None):
        w, b = self.classifier.weight.detach(), self.classifier.bias.detach()
        Cn, F = int(w.shape[0]), int(w.shape[1])
        w = w.reshape(Cn, F)
        if not objects.is_cuda or objects.dtype != torch.float32 or not objects.is_contiguous() or objects.dim() != 3 \
                or objects.shape[0] != F:
            raise RuntimeError("gsraster_b200.losses: objects must be a contiguous CUDA float32 (F, H, W) tensor (no CPU fallback)")
        H, W = int(objects.shape[1]), int(objects.shape[2])
        if not gt.is_cuda or gt.dtype != torch.int64 or not gt.is_contiguous() or gt.numel() != H * W:
            raise RuntimeError("gsraster_b200.losses: gt must be a contiguous CUDA int64 (H, W) tensor")
        if not (w.is_cuda and w.is_contiguous() and w.dtype == torch.float32 and b.is_cuda and b.dtype == torch.float32):
            raise RuntimeError("gsraster_b200.losses: classifier weight / bias must be CUDA float32")
        out2 = out if out is not None else torch.empty(2, dtype=torch.float32, device=objects.device)
        grad = (grad_out if grad_out is not None else torch.empty_like(objects)) if need_grad else None
        pred = torch.empty((H, W), dtype=torch.int32, device=objects.device) if need_pred else None
        with torch.cuda.device(objects.device):
            capi._check(_bind_feat().gsr_feature_loss(
                objects.data_ptr(), gt.data_ptr(), H, W, F, Cn, w.data_ptr(), b.data_ptr(), float(grad_weight),
                None if grad is None else grad.data_ptr(), None if pred is None else pred.data_ptr(), out2.data_ptr(),
                self.scratch.data_ptr(), capi._stream_ptr()))
        return out2, grad, pred


def _bind_cls3d():
    L = capi.lib()
    if not getattr(L, "_cls3d_bound", False):
        L.gsr_cls3d_loss_scratch_bytes.restype = C.c_size_t
        L.gsr_cls3d_loss_scratch_bytes.argtypes = [C.c_int, C.c_int]
        L.gsr_cls3d_loss.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p]
        L._cls3d_bound = True
    return L


class Cls3DLoss:
    """3-D feature-consistency regulariser (csrc/cls3d.cu): the reference's
    `prob = softmax(classifier(_objects_dc.permute(2,0,1)), 0)...; loss_cls_3d(_xyz, prob, 5, 2, 300000, 1000)`
    (scripts/gaussian_splatting.py:812-814, utils/loss_utils.py:71-112) without the (C, P) probability tensor and without
    the sample x subset distance matrix. The two index sets the reference draws with torch.randperm are arguments:
    `subset` (int32 indices into the P Gaussians, or None = all) and `sample` (int32 indices into the subset);
    `draw_indices` draws them the way the reference does.

    __call__(xyz (P,3), objects (P,F) or (P,1,F), subset, sample) -> (out1, neighbors); the gradient (times grad_weight)
    is ADDED to `grad_out` (P,F) when given."""

    def __init__(self, classifier, device="cuda"):
        self.classifier = classifier
        self.device = device
        self.scratch = None

    @staticmethod
    def draw_indices(P, max_points=300000, sample_size=1000, device="cuda", generator=None):
        """The reference's two torch.randperm draws (loss_utils.py:86-95) -> (subset or None, sample) as int32."""
        subset = None
        n = P
        if P > max_points:
            subset = torch.randperm(P, generator=generator)[:max_points].to(device=device, dtype=torch.int32)
            n = max_points
        sample = torch.randperm(n, generator=generator)[:sample_size].to(device=device, dtype=torch.int32)
        return subset, sample

    def __call__(self, xyz, objects, subset, sample, k=5, lambda_val=2.0, grad_weight=1.0, grad_out=None, out=None):
        w, b = self.classifier.weight.detach(), self.classifier.bias.detach()
        Cn, F = int(w.shape[0]), int(w.shape[1])
        w = w.reshape(Cn, F)
        P = int(xyz.shape[0])
        for t, shp in ((xyz, P * 3), (objects, P * F)):
            if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != shp:
                raise RuntimeError("gsraster_b200.losses: xyz (P,3) / objects (P,F) must be contiguous CUDA float32 (no CPU fallback)")
        for t in (subset, sample):
            if t is not None and (not t.is_cuda or t.dtype != torch.int32 or not t.is_contiguous()):
                raise RuntimeError("gsraster_b200.losses: subset / sample must be contiguous CUDA int32 tensors")
        if grad_out is not None and (not grad_out.is_cuda or grad_out.dtype != torch.float32 or not grad_out.is_contiguous()
                                     or grad_out.numel() != P * F):
            raise RuntimeError("gsraster_b200.losses: grad_out must be a contiguous CUDA float32 tensor of P*F elements")
        S = int(sample.numel())
        n = P if subset is None else int(subset.numel())
        L = _bind_cls3d()
        need = int(L.gsr_cls3d_loss_scratch_bytes(S, n)) + 256
        if self.scratch is None or self.scratch.numel() < need:
            self.scratch = torch.empty(need, dtype=torch.uint8, device=xyz.device)
        out1 = out if out is not None else torch.empty(1, dtype=torch.float32, device=xyz.device)
        neighbors = torch.empty((S, int(k)), dtype=torch.int32, device=xyz.device)
        with torch.cuda.device(xyz.device):
            capi._check(L.gsr_cls3d_loss(xyz.data_ptr(), objects.data_ptr(), P, F, Cn, w.data_ptr(), b.data_ptr(),
                                         None if subset is None else subset.data_ptr(), n, sample.data_ptr(), S, int(k),
                                         float(lambda_val), float(grad_weight),
                                         None if grad_out is None else grad_out.data_ptr(), out1.data_ptr(),
                                         neighbors.data_ptr(), self.scratch.data_ptr(), capi._stream_ptr()))
        return out1, neighbors
