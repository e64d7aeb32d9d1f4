"""GPU diagnostic (not a pytest): where does a fwd+bwd step spend its time? Host timestamps and CUDA
events around forward / backward of the headline workload, plus the per-stage event times.
Usage: python tests/gpu_timeline.py [P] [steps]"""
import sys
import time

import torch

import helpers  # noqa: F401  (sys.path)
from gsraster_b200 import capi, scenes

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
W, H, deg, F = 1280, 1024, 3, 16
sc = scenes.make_scene(P, deg, F, 1234, W, H)
cam = scenes.make_camera(W, H, 0)
d = helpers.to_torch(sc, cam)
g = [torch.from_numpy(x).cuda() for x in scenes.make_upstream_grads(W, H, F)]
e = torch.empty(0, device="cuda")


timer = None
ext = None


def fwd():
    return capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                    d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"],
                                    d["sh_objs"], deg, d["campos"], False, False, ext=ext)


def bwd(res):
    R, depth, color, objects, radii, is_used, gb, bb, ib = res
    return capi.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                             d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, g[0], g[1],
                                             d["shs"], g[2], d["sh_objs"], deg, d["campos"], gb, R, bb, ib, False, ext=ext)


for _ in range(3):
    bwd(fwd())
torch.cuda.synchronize()
timer = capi.Timer()
ext = capi.Ext(timer=timer)
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
host = []
fstats = []
t00 = time.perf_counter()
for s in range(steps):
    t0 = time.perf_counter()
    ev[s][0].record()
    res = fwd()
    t1 = time.perf_counter()
    fstats.append(capi.last_forward_stats())
    ev[s][1].record()
    bwd(res)
    t2 = time.perf_counter()
    ev[s][2].record()
    host.append((t1 - t0, t2 - t1))
torch.cuda.synchronize()
wall = time.perf_counter() - t00
st, gp = timer.read()
print(f"P={P} R={res[0]} wall/step {1e3 * wall / steps:.3f} ms")
for s in range(steps):
    print(f" step {s}: host fwd {1e3 * host[s][0]:.3f} bwd {1e3 * host[s][1]:.3f} ms | gpu fwd "
          f"{ev[s][0].elapsed_time(ev[s][1]):.3f} bwd {ev[s][1].elapsed_time(ev[s][2]):.3f} ms | sync/alloc/launch us "
          f"{fstats[s]['sync_us']}/{fstats[s]['alloc_us']}/{fstats[s]['launch_us']}")
print(" stage ms/step:", {k: round(v / steps, 4) for k, v in st.items()}, "sum", round(sum(st.values()) / steps, 4))
print(" idle before stage, ms/step:", {k: round(v / steps, 4) for k, v in gp.items()}, "sum", round(sum(gp.values()) / steps, 4))
# forward only (render FPS)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(steps):
    fwd()
torch.cuda.synchronize()
print(f" forward only: {1e3 * (time.perf_counter() - t0) / steps:.3f} ms -> {steps / (time.perf_counter() - t0):.1f} FPS")
