"""Generates tests/golden/featloss_*.npz with the exact torch modules and call sequence of the reference's mapping
step (scripts/gaussian_splatting.py:215-216 `classifier = torch.nn.Conv2d(num_objects, num_classes, kernel_size=1)`,
`cls_criterion = torch.nn.CrossEntropyLoss(reduction='none')`; :798-803 logits / argmax / loss_obj / log(num_classes)).
The script file itself cannot be imported (ROS node), so the five lines are replayed here on CPU fp32; the gradient
comes from torch autograd. Run in the build container."""
import os

import numpy as np
import torch

here = os.path.dirname(os.path.abspath(__file__))
for name, (H, W, F, C, seed, ignore_frac) in dict(a=(24, 40, 16, 100, 1, 0.0), b=(19, 23, 16, 100, 2, 0.1),
                                                  c=(16, 24, 32, 37, 3, 0.0)).items():
    torch.manual_seed(seed)
    classifier = torch.nn.Conv2d(F, C, kernel_size=1)
    cls_criterion = torch.nn.CrossEntropyLoss(reduction='none')
    objects = (2.0 * torch.randn(F, H, W)).requires_grad_(True)
    gt_obj = torch.randint(0, C, (H, W))
    gt_obj[torch.rand(H, W) < ignore_frac] = -100
    logits = classifier(objects)
    pred_obj = torch.argmax(logits, dim=0)
    loss_obj = cls_criterion(logits.unsqueeze(0), gt_obj.unsqueeze(0)).squeeze().mean()
    mean_ce = float(loss_obj)
    loss_obj = loss_obj / torch.log(torch.tensor(C))
    loss_obj.backward()
    np.savez_compressed(os.path.join(here, f"featloss_{name}.npz"), objects=objects.detach().numpy(), gt=gt_obj.numpy(),
                        weight=classifier.weight.detach().numpy().reshape(C, F), bias=classifier.bias.detach().numpy(),
                        loss=np.float32(loss_obj.item()), mean_ce=np.float32(mean_ce), grad=objects.grad.numpy(),
                        pred=pred_obj.numpy().astype(np.int32))
    print(name, H, W, F, C, float(loss_obj))
