"""Multi-GPU check of the sparse gradient exchange (torchrun, one rank per GPU): every rank renders its own V views
forward + backward, accumulating sparse rows into its bucket (gsr_ext: touched flags, sparse_rows, accumulate, row
workspace); the summed gradients from SparseGradExchange (NCCL transport with the blend-done overlap at V = 1, then
the peer-memory transport) must equal a dense all-reduce of the ranks' per-view dense gradients.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
      tests/gpu_exchange_check.py [P]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "segment-any-gaussian-splatting-slam_b200"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import torch
import torch.distributed as dist

import helpers as h
from helpers import scenes
from gsraster_b200 import capi, mapping


def run_check(P=200000, W=640, H=480, F=16, deg=3, steps=3, peer=False, V=1):
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    M = (deg + 1) ** 2
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=77, W=W, H=H)
    cams = [scenes.make_camera(W, H, rank * V + v) for v in range(V)]
    cam = cams[0]
    ds = [h.to_torch(sc, c, device=dev) for c in cams]
    d = ds[0]
    gd, gc, go = [torch.from_numpy(x).to(dev) for x in scenes.make_upstream_grads(W, H, F, seed=5 + rank)]
    e = torch.empty(0, device=dev)
    bucket = mapping.GradBucket(P, M, F, dev, symmetric=peer, copies=2 if peer else 1)
    ex = mapping.SparseGradExchange(bucket, peer_memory=peer)
    ref = mapping.GradBucket(P, M, F, dev)
    tmp = mapping.GradBucket(P, M, F, dev)
    worst = 0.0
    for it in range(steps):
        ex.begin_step()
        ext = ex.ext(accumulate=V > 1, overlap=(V == 1))
        ref.flat.zero_()
        for v in range(V):
            d = ds[v]
            res = capi.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                           d["viewmatrix"], d["projmatrix"], cam.tanfovx, cam.tanfovy, H, W, d["shs"],
                                           d["sh_objs"], deg, d["campos"], False, False)
            args = (d["bg"], d["means3D"], res[4], e, d["scales"], d["rotations"], 1.0, e, d["viewmatrix"], d["projmatrix"],
                    cam.tanfovx, cam.tanfovy, gd, gc, d["shs"], go, d["sh_objs"], deg, d["campos"], res[6], res[0], res[7],
                    res[8], False)
            capi.rasterize_gaussians_backward(*args, out=bucket.views, ext=ext)     # sparse rows, accumulated
            capi.rasterize_gaussians_backward(*args, out=tmp.views)                  # the reference-contract dense rows
            ref.flat.add_(tmp.flat)
        ex.exchange()
        idx, views = ex.finish()
        dist.all_reduce(ref.flat)
        for k, n in ex.sizes:
            full = ref.views[k].view(P, n)
            sel = full.index_select(0, idx.to(torch.int64))
            scale = float(full.abs().max()) + 1e-30
            worst = max(worst, float((sel - views[k]).abs().max()) / scale)
            mask = torch.ones(P, dtype=torch.bool, device=dev)
            mask[idx.to(torch.int64)] = False
            assert float(full[mask].abs().max() if mask.any() else 0.0) == 0.0, (k, "non-zero row outside the union")
    t = torch.tensor([worst], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t), ex.last_rows, ex.fallbacks


if __name__ == "__main__":
    lr = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    bad = False
    for peer, V in ((False, 1), (False, 2), (True, 1), (True, 3)):
        worst, rows, fb = run_check(P=P, peer=peer, V=V)
        if dist.get_rank() == 0:
            print("%s exchange check, %d view(s) per rank: world %d, union rows %d, fallbacks %d, worst |sparse - dense| / max = %.3g -> %s" % (
                "peer-memory" if peer else "NCCL", V, dist.get_world_size(), rows, fb, worst,
                "OK" if worst <= 1e-5 else "FAIL"), flush=True)
        bad = bad or worst > 1e-5
    dist.destroy_process_group()
    sys.exit(1 if bad else 0)
