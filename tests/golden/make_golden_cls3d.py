"""Generates tests/golden/cls3d_*.npz by importing the REFERENCE's own loss_cls_3d
(/root/reference/src/gaussian_splatting/gaussian_splatting/utils/loss_utils.py:71-112, pure torch, CPU) and calling it as
scripts/gaussian_splatting.py:812-814 does (classifier = Conv2d(16, C, 1) on _objects_dc.permute(2,0,1), softmax over
classes). The function draws its random index sets with torch.randperm; the generator is seeded before the call and the
same two draws are replayed afterwards to record which indices it used. Run in the build container."""
import importlib.util
import os

import numpy as np
import torch

REF = "/root/reference/src/gaussian_splatting/gaussian_splatting/utils/loss_utils.py"
spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)
here = os.path.dirname(os.path.abspath(__file__))
for name, (P, F, C, max_points, sample_size, seed) in dict(a=(3000, 16, 100, 300000, 200, 1),
                                                          b=(5000, 16, 100, 2000, 150, 2),
                                                          c=(800, 16, 37, 300000, 1000, 3)).items():
    torch.manual_seed(seed)
    classifier = torch.nn.Conv2d(F, C, kernel_size=1)
    xyz = torch.randn(P, 3) * torch.tensor([2.0, 1.0, 3.0])
    objects_dc = torch.randn(P, 1, F).requires_grad_(True)
    logits3d = classifier(objects_dc.permute(2, 0, 1))
    prob_obj3d = torch.softmax(logits3d, dim=0).squeeze().permute(1, 0)
    torch.manual_seed(1000 + seed)
    loss = ref.loss_cls_3d(xyz.squeeze().detach(), prob_obj3d, 5, 2, max_points, sample_size)
    loss.backward()
    torch.manual_seed(1000 + seed)                                  # replay the draws of loss_utils.py:86-95
    subset = None
    n = P
    if P > max_points:
        subset = torch.randperm(P)[:max_points]
        n = max_points
    sample = torch.randperm(n)[:sample_size]
    np.savez_compressed(os.path.join(here, f"cls3d_{name}.npz"), xyz=xyz.numpy(), objects=objects_dc.detach().numpy().reshape(P, F),
                        weight=classifier.weight.detach().numpy().reshape(C, F), bias=classifier.bias.detach().numpy(),
                        subset=(subset.numpy().astype(np.int32) if subset is not None else np.zeros(0, np.int32)),
                        sample=sample.numpy().astype(np.int32), loss=np.float32(loss.item()),
                        grad=objects_dc.grad.numpy().reshape(P, F))
    print(name, P, C, float(loss))
