"""GPU diagnostic (not a pytest): per-Gaussian backward (stage 5) under the gsr_ext output modes on two workloads."""
import copy
import sys

import torch

import helpers  # noqa: F401
import bench
from gsraster_b200 import capi, mapping, scenes

sys.argv = ["bench.py"]
args = bench.parse_args()
dev = torch.device("cuda", 0)
for name in ("non_saturating", "survey8d_view0"):
    a = copy.copy(args)
    a.P, a.pool = scenes.WORKLOADS[name]["P"], 1
    wl = bench.Workload(a, dev, scene=scenes.make_workload_scene(name, 3, 16, 1280, 1024), num_poses=1)
    bucket = mapping.GradBucket(wl.P, wl.M, a.F, dev)
    mask = torch.zeros(wl.P, dtype=torch.uint8, device=dev)
    ws = capi.new_row_workspace(wl.P, a.F, dev)
    res = wl.fwd(0)
    for label, kw in (("dense", {}), ("dense+ws", dict(row_workspace=ws)), ("dense+mask", dict(touched_mask=mask)),
                      ("sparse", dict(touched_mask=mask, sparse_rows=True)),
                      ("sparse+ws", dict(touched_mask=mask, sparse_rows=True, row_workspace=ws)),
                      ("sparse+ws+acc", dict(touched_mask=mask, sparse_rows=True, row_workspace=ws, accumulate=True))):
        tm = capi.Timer()
        for i in range(13):
            if i == 3:
                tm.read()
            mask.zero_()
            wl.bwd(0, res, out=bucket.views, ext=capi.Ext(timer=tm, **kw))
        ms, _ = tm.read()
        print("%-16s %-14s K5 %.4f ms  K7 %.4f ms" % (name, label, ms["render_bwd"] / 10, ms["preprocess_bwd"] / 10))
