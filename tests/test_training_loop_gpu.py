"""The pieces compose into the reference's mapping iteration (scripts/gaussian_splatting.py:792-833: render ->
photometric + segmentation-feature loss -> backward -> optimizer.step()). Loop A is the reference's own structure -- torch activations, the
drop-in `diff_gaussian_rasterization` autograd Function, the loss restated in torch, torch.optim.Adam. Loop B is the fused
path -- FusedGaussianAdam's activated tensors -> C-ABI forward -> gsr_photometric_loss + gsr_feature_loss -> C-ABI backward into the gradient
bucket -> gsr_adam_step. After three iterations both must hold the same parameters."""
import pytest
import torch

import helpers as h
from helpers import scenes
from gsraster_b200 import capi, losses, mapping, optim
from oracle import pyloss

pytestmark = pytest.mark.gpu


def test_fused_mapping_iteration_matches_the_reference_style_loop():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    W, H, P, F, deg, lam = 320, 256, 6000, 16, 3, 0.2
    M = (deg + 1) ** 2
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=31, W=W, H=H, scale_mul=2.0)
    cam = scenes.make_camera(W, H, 0)
    d = h.to_torch(sc, cam)
    raw0 = dict(xyz=d["means3D"].clone(), f_dc=d["shs"][:, :1, :].contiguous(), f_rest=d["shs"][:, 1:, :].contiguous(),
                opacity=torch.logit(d["opacities"].clamp(1e-4, 1 - 1e-4)), scaling=torch.log(d["scales"]),
                rotation=d["rotations"] * 1.7, objects=d["sh_objs"].reshape(P, 1, F).clone())
    g = torch.Generator(device="cuda").manual_seed(8)
    gt = torch.rand(3, H, W, generator=g, device="cuda")
    gt[:, :20, :30] = 0.0                       # exercises the gt != 0 masks
    torch.manual_seed(5)
    classifier = torch.nn.Conv2d(F, 100, kernel_size=1).cuda()          # scripts/gaussian_splatting.py:215-216
    cls_criterion = torch.nn.CrossEntropyLoss(reduction='none')
    gt_obj = torch.randint(0, 100, (H, W), generator=g, device="cuda")
    torch.backends.cudnn.allow_tf32 = False                             # loop A's 1x1 convolution in fp32
    lr = dict(optim.DEFAULT_LR, xyz=1e-4)
    iters = 3

    # ---- loop A: the reference's structure ----
    pa = {k: v.clone().requires_grad_(True) for k, v in raw0.items()}
    opt = torch.optim.Adam([{"params": [pa[k]], "lr": lr[k]} for k in optim.GROUPS], lr=0.0, eps=1e-15)
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d["bg"], 1.0, d["viewmatrix"], d["projmatrix"], deg,
                                       d["campos"], False, False)
    losses_a = []
    for _ in range(iters):
        opt.zero_grad()
        depth, color, radii, objects, is_used = GaussianRasterizer(rs)(
            means3D=pa["xyz"], means2D=torch.zeros_like(pa["xyz"]), opacities=torch.sigmoid(pa["opacity"]),
            shs=torch.cat((pa["f_dc"], pa["f_rest"]), dim=1), sh_objs=pa["objects"],
            scales=torch.exp(pa["scaling"]), rotations=torch.nn.functional.normalize(pa["rotation"]))
        loss_rgb, _, _ = pyloss.photometric_loss(color, gt, lam)
        logits = classifier(objects)                                                         # :798-803
        loss_obj = cls_criterion(logits.unsqueeze(0), gt_obj.unsqueeze(0)).squeeze().mean() / torch.log(torch.tensor(100.0))
        loss = 1.0 * loss_rgb + 1.0 * loss_obj                                               # :820-821
        loss.backward()
        opt.step()
        losses_a.append(float(loss.detach()))

    # ---- loop B: the fused path ----
    pb = {k: v.clone() for k, v in raw0.items()}
    fa = optim.FusedGaussianAdam(pb, lr=lr)
    bucket = mapping.GradBucket(P, M, F, "cuda")
    ploss = losses.PhotometricLoss(H, W, lam)
    floss = losses.FeatureLoss(classifier)
    e = torch.empty(0, device="cuda")
    zd = torch.zeros(1, H, W, device="cuda")
    losses_b = []
    for _ in range(iters):
        a = fa.rasterizer_inputs()
        res = capi.rasterize_gaussians(d["bg"], a["means3D"], e, a["opacities"], a["scales"], a["rotations"], 1.0, e,
                                       d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, a["shs"],
                                       a["sh_objs"], deg, d["campos"], False, False)
        out3, grad_c = ploss(res[2], gt)
        out2, grad_o, _ = floss(res[3], gt_obj)
        capi.rasterize_gaussians_backward(d["bg"], a["means3D"], res[4], e, a["scales"], a["rotations"], 1.0, e,
                                          d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, zd, grad_c, a["shs"],
                                          grad_o, a["sh_objs"], deg, d["campos"], res[6], res[0], res[7], res[8], False,
                                          out=bucket.views)
        fa.step(bucket.views)
        losses_b.append(float(out3[0]) + float(out2[0]))

    for la, lb in zip(losses_a, losses_b):
        assert abs(la - lb) <= 2e-5 * abs(la), (losses_a, losses_b)
    assert losses_b[-1] < losses_b[0]                       # it actually descends
    for k in optim.GROUPS:
        upd_a, upd_b = pa[k].detach() - raw0[k], pb[k] - raw0[k]
        scale = float(upd_a.abs().max())
        assert scale > 0, k
        if scale > 0:
            # Adam's first steps are lr * sign-like ratios: compare the updates, not the parameters
            bad = ((upd_a - upd_b).abs() > 2e-2 * scale).float().mean()
            assert float(bad) < 1e-3, (k, float(bad))
