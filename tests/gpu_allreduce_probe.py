"""GPU diagnostic (not a pytest): NCCL all-reduce time for the gradient bucket size.
torchrun --nproc-per-node N tests/gpu_allreduce_probe.py [MB]"""
import os
import sys

import torch
import torch.distributed as dist

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 300
lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
x = torch.ones(mb * 250_000, device="cuda")
for _ in range(5):
    dist.all_reduce(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
dist.barrier()
e0.record()
for _ in range(20):
    dist.all_reduce(x)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
n = dist.get_world_size()
if dist.get_rank() == 0:
    print(f"allreduce {mb} MB x{n}: {ms:.3f} ms  busbw {2 * (n - 1) / n * mb / ms:.1f} GB/s")
# symmetric-memory variants (torch plumbing; measures what NVLS multimem / P2P two-shot reach on this box)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(mb * 250_000, device="cuda")
    t.fill_(1.0)
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    if dist.get_rank() == 0:
        print("symm: multicast_ptr", hex(hdl.multicast_ptr) if hdl.multicast_ptr else None, "world", hdl.world_size)
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
        op = getattr(torch.ops.symm_mem, name, None)
        if op is None or (name.startswith("multimem") and not hdl.multicast_ptr):
            continue
        for _ in range(3):
            op(t, "sum", dist.group.WORLD.group_name)
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        for _ in range(20):
            op(t, "sum", dist.group.WORLD.group_name)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        if dist.get_rank() == 0:
            print(f"symm_mem.{name} {mb} MB x{n}: {ms:.3f} ms  busbw {2 * (n - 1) / n * mb / ms:.1f} GB/s")
except Exception as ex:  # diagnostic only
    if dist.get_rank() == 0:
        print("symm_mem probe failed:", repr(ex)[:300])
dist.destroy_process_group()
