"""GPU diagnostic (not a pytest): ONE fused mapping iteration of the headline shape between cudaProfilerStart/Stop, for
  ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/full python tests/gpu_profile_step.py
forward -> photometric loss (lambda 0.2) -> classifier + cross-entropy loss -> backward -> dense Adam. Three untimed
iterations first. Without ncu it just runs (exit 0) and prints the iteration time."""
import sys

import torch

import helpers  # noqa: F401  (sys.path)
from gsraster_b200 import capi, losses, mapping, optim, scenes

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
W, H, deg, F, C = 1280, 1024, 3, 16, 100
M = (deg + 1) ** 2
sc = scenes.make_scene(P, deg, F, 1234, W, H)
cam = scenes.make_camera(W, H, 0)
d = helpers.to_torch(sc, cam)
raw = dict(xyz=d["means3D"].clone(), f_dc=d["shs"][:, :1, :].contiguous(), f_rest=d["shs"][:, 1:, :].contiguous(),
           opacity=torch.logit(d["opacities"].clamp(1e-4, 1 - 1e-4)), scaling=torch.log(d["scales"]),
           rotation=d["rotations"].clone(), objects=d["sh_objs"].reshape(P, 1, F).clone())
fa = optim.FusedGaussianAdam(raw)
bucket = mapping.GradBucket(P, M, F, "cuda")
ploss = losses.PhotometricLoss(H, W, 0.2)
torch.manual_seed(3)
floss = losses.FeatureLoss(torch.nn.Conv2d(F, C, kernel_size=1).cuda())
gt = torch.rand(3, H, W, device="cuda")
gt_obj = torch.randint(0, C, (H, W), device="cuda")
e = torch.empty(0, device="cuda")
zd = torch.zeros(1, H, W, device="cuda")


def iteration():
    a = fa.rasterizer_inputs()
    res = capi.rasterize_gaussians(d["bg"], a["means3D"], e, a["opacities"], a["scales"], a["rotations"], 1.0, e,
                                   d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, a["shs"], a["sh_objs"],
                                   deg, d["campos"], False, False)
    out3, gc = ploss(res[2], gt)
    out2, go, _ = floss(res[3], gt_obj, need_pred=True)
    capi.rasterize_gaussians_backward(d["bg"], a["means3D"], res[4], e, a["scales"], a["rotations"], 1.0, e,
                                      d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, zd, gc, a["shs"], go,
                                      a["sh_objs"], deg, d["campos"], res[6], res[0], res[7], res[8], False,
                                      out=bucket.views)
    fa.step(bucket.views)
    return out3, out2


for _ in range(3):
    iteration()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
e0.record()
o3, o2 = iteration()
e1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("one fused mapping iteration: %.3f ms, loss_rgb %.5f loss_obj %.5f" % (e0.elapsed_time(e1), float(o3[0]), float(o2[0])))
