"""GPU diagnostic (not a pytest): stage times of the forward pass on keyframe 0 of the headline trajectory with the depth
sort on the MAIN stream (GSR_NO_SIDE_SORT=1: stage 1 = sort + coverage / offsets, not their maximum), for A/B runs of
the sort variants (GSR_SORT_VARIANT, GSR_SORT_V1). 30 calls after 5 warm-up; also checks the sorted order against torch."""
import os
import sys

os.environ.setdefault("GSR_NO_SIDE_SORT", "1")
import torch

import helpers
from gsraster_b200 import capi, scenes

W, H, P, F = 1280, 1024, int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, 16
sc = scenes.make_trajectory_scene(P, 16, sh_degree=3, F=F, seed=1234, W=W, H=H)
cam = scenes.make_camera(W, H, 0)
d = helpers.to_torch(sc, cam)
e = torch.empty(0, device="cuda")
timer = capi.Timer()
ext = capi.Ext(timer=timer)


def fwd():
    return capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                    d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"], d["sh_objs"],
                                    3, d["campos"], False, False, ext=ext)


for _ in range(5):
    res = fwd()
timer.read()
n = 30
for _ in range(n):
    res = fwd()
st, gap = timer.read()
print("variant %s v1 %s: " % (os.environ.get("GSR_SORT_VARIANT", "0"), os.environ.get("GSR_SORT_V1", "-")) +
      " ".join("%s %.4f" % (k, v / n) for k, v in st.items() if v > 0), flush=True)
print("checksum color %.6f R %d" % (float(res[2].double().sum()), int(res[0])))
