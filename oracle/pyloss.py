"""TEST INFRASTRUCTURE (oracle): torch restatement of the reference's photometric loss
(src/gaussian_splatting/gaussian_splatting/utils/loss_utils.py:17-20 l1_loss, :27-69 gaussian / create_window / ssim /
_ssim; composed as at scripts/gaussian_splatting.py:805-807). Pinned against golden vectors produced by importing the
reference module itself (tests/golden/make_golden_loss.py). Only tests/ may import this; the product path is
csrc/loss.cu."""
import numpy as np
import torch
import torch.nn.functional as F


def photometric_loss(img, gt, lam):
    """-> (loss_rgb, Ll1, L_ssim) as 0-d tensors; differentiable w.r.t. img."""
    l1 = torch.where(gt != 0, (img - gt).abs(), torch.zeros_like(img)).mean()            # loss_utils.py:17-20
    x = torch.where(gt != 0, img, torch.zeros_like(img))                                  # :37
    g = torch.tensor([np.exp(-(i - 5) ** 2 / (2 * 1.5 ** 2)) for i in range(11)], dtype=torch.float32, device=img.device)
    g = g / g.sum()                                                                       # :27-29
    w = (g[:, None] @ g[None, :]).expand(3, 1, 11, 11).contiguous()                       # :31-35
    conv = lambda t: F.conv2d(t[None], w, padding=5, groups=3)[0]
    mu1, mu2 = conv(x), conv(gt)                                                          # :49-50
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(gt * gt) - mu2 * mu2, conv(x * gt) - mu1 * mu2   # :56-58
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))).mean()  # :63-66
    return (1 - lam) * l1 + lam * (1 - ssim), l1, ssim


def feature_loss(objects, gt, weight, bias):
    """TEST INFRASTRUCTURE. The segmentation-feature loss of the mapping step, restated without torch.nn:
    logits = Conv2d(F, C, 1)(objects) = weight (C,F) @ objects (F,HW) + bias; CrossEntropyLoss(reduction='none') =
    logsumexp_c(logits) - logits[gt] (0 where gt == ignore_index -100), mean over ALL pixels, / ln C
    (scripts/gaussian_splatting.py:215-216, 798-803). numpy float64.
    -> (loss_obj, mean_ce, d loss_obj / d objects (F,H,W), argmax (H,W))."""
    Fd, H, W = objects.shape
    Cn = weight.shape[0]
    x = objects.reshape(Fd, -1).astype(np.float64)
    z = weight.reshape(Cn, Fd).astype(np.float64) @ x + bias.astype(np.float64)[:, None]
    m = z.max(0)
    e = np.exp(z - m)
    s = e.sum(0)
    g = gt.reshape(-1).astype(np.int64)
    use = (g >= 0) & (g < Cn)
    gi = np.where(use, g, 0)
    ce = np.where(use, np.log(s) + m - z[gi, np.arange(z.shape[1])], 0.0)
    mean_ce = ce.mean()
    d = e / s
    d[gi, np.arange(z.shape[1])] -= 1.0
    d *= use[None, :] / (z.shape[1] * np.log(Cn))
    dobj = weight.reshape(Cn, Fd).astype(np.float64).T @ d
    return mean_ce / np.log(Cn), mean_ce, dobj.reshape(Fd, H, W), z.argmax(0).reshape(H, W)


def cls3d_loss(xyz, objects, weight, bias, subset, sample, k=5, lambda_val=2.0):
    """TEST INFRASTRUCTURE. loss_cls_3d (utils/loss_utils.py:71-112) on prob = softmax(classifier(objects)) as called at
    scripts/gaussian_splatting.py:812-814, restated in torch float64 with the two random index sets given (subset: indices
    into the P points or None; sample: indices into the subset). Exact pairwise distances (the reference's torch.cdist may
    use the |a|^2 + |b|^2 - 2ab form; neighbour sets agree except for near-ties).
    -> (loss, d loss / d objects (P,F), neighbour indices (S,k) into the subset)."""
    x = torch.as_tensor(xyz, dtype=torch.float64)
    o = torch.as_tensor(objects, dtype=torch.float64).reshape(x.shape[0], -1).clone().requires_grad_(True)
    W = torch.as_tensor(weight, dtype=torch.float64).reshape(-1, o.shape[1])
    b = torch.as_tensor(bias, dtype=torch.float64)
    prob = torch.softmax(o @ W.T + b, dim=1)
    if subset is not None:
        sub = torch.as_tensor(subset, dtype=torch.int64)
        x, prob = x[sub], prob[sub]
    smp = torch.as_tensor(sample, dtype=torch.int64)
    d = ((x[smp][:, None, :] - x[None, :, :]) ** 2).sum(-1)
    nb = d.topk(k, largest=False).indices
    sp, npr = prob[smp], prob[nb]
    kl = sp[:, None, :] * (torch.log(sp[:, None, :] + 1e-10) - torch.log(npr + 1e-10))
    loss = lambda_val * kl.sum(-1).mean() / prob.shape[1]
    loss.backward()
    return float(loss), o.grad.numpy(), nb.numpy()
