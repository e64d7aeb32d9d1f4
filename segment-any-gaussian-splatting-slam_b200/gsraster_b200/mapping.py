"""View-data-parallel keyframe mapping step (no reference equivalent: the reference
optimises one view per step on one GPU, scripts/gaussian_splatting.py:792-833).

Gaussian parameters are replicated on every rank; a batch of V keyframe views is split across
ranks; each rank runs rasterizer forward+backward for its views; the Gaussian-parameter gradients
are summed over ranks with ONE all-reduce over a single flat fp32 bucket that the backward kernels
write in place (no packing pass). Batch gradient = sum of per-view gradients.

Host logic only (sharding, bucket layout, collective); works with the gloo backend on CPU for the
tests, NCCL over NVLink on the GPUs.
"""
from typing import Callable, Dict, List, Sequence

import torch

# (key in capi.rasterize_gaussians_backward's `out`, floats per Gaussian as a function of (M, F))
BUCKET_LAYOUT = (("means3D", lambda M, F: 3), ("sh", lambda M, F: 3 * M), ("objects", lambda M, F: F),
                 ("opacity", lambda M, F: 1), ("scales", lambda M, F: 3), ("rotations", lambda M, F: 4))


def shard_views(num_views: int, rank: int, world: int) -> List[int]:
    """Views owned by `rank`: contiguous blocks, sizes differing by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(num_views, world)
    start = rank * base + min(rank, rem)
    return list(range(start, start + base + (1 if rank < rem else 0)))


class GradBucket:
    """One flat fp32 buffer of 4*(11 + 3M + F) bytes per Gaussian (300 B at M=16, F=16) holding the six
    parameter-gradient tensors back to back (tensor-major), each exposed as a contiguous view."""

    def __init__(self, P: int, M: int, F: int, device):
        self.P, self.M, self.F = P, M, F
        sizes = [(k, fn(M, F)) for k, fn in BUCKET_LAYOUT]
        self.floats_per_gaussian = sum(n for _, n in sizes)
        # every segment starts 16-byte aligned (float4 stores in the kernels): pad P up to a multiple of 4
        Ppad = (P + 3) // 4 * 4
        self.flat = torch.zeros(Ppad * self.floats_per_gaussian, dtype=torch.float32, device=device)
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for k, n in sizes:
            self.views[k] = self.flat[off:off + P * n]
            off += Ppad * n

    @property
    def nbytes(self) -> int:
        return self.P * self.floats_per_gaussian * 4

    def shaped(self) -> Dict[str, torch.Tensor]:
        P, M, F = self.P, self.M, self.F
        shp = dict(means3D=(P, 3), sh=(P, M, 3), objects=(P, F), opacity=(P, 1), scales=(P, 3), rotations=(P, 4))
        return {k: v.view(shp[k]) for k, v in self.views.items()}


def mapping_step(view_ids: Sequence[int], grad_fn: Callable[[int, Dict[str, torch.Tensor], bool], None],
                 bucket: GradBucket, group=None, world: int = 1) -> GradBucket:
    """Runs grad_fn(view, out_views, accumulate) for every local view -- grad_fn must WRITE the view's
    parameter gradients into `out_views` when accumulate is False and ADD when True -- then sums the
    bucket over ranks. Returns the bucket (identical on every rank)."""
    if len(view_ids) == 0:
        bucket.flat.zero_()
    for i, v in enumerate(view_ids):
        grad_fn(v, bucket.views, i > 0)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(bucket.flat, op=dist.ReduceOp.SUM, group=group)
    return bucket


class SparseGradExchange:
    """Sum of the per-rank gradient buckets restricted to the Gaussians that ANY rank touched.

    A keyframe's backward pass gives non-zero gradients only to the Gaussians its blend actually reached (pixels
    saturate after a few hundred list entries, and a SLAM map is much larger than one view's frustum), so the dense
    4(11 + 3M + F) B/Gaussian all-reduce mostly moves zeros. The backward blend flags the Gaussians it wrote
    (capi.set_touched_mask); the exchange is then, with no host synchronisation in between,
      1. all-gather of the P-byte flags, union over ranks
      2. inclusive scan of the union -> sorted index list idx[0:count], count stays on the device
      3. gather rows idx[0:cap] of every bucket tensor into a compact tensor-major buffer (cap rows)
      4. all-reduce(SUM) of cap * (11 + 3M + F) floats
    `cap` is the capacity guessed from the previous step (1.25 x its count, multiple of 1024; P on the first step).
    `finish()` reads count back; if it exceeded cap the step is redone with a dense all-reduce of the bucket, so the
    result never depends on the guess. Rows outside the union are exactly zero on every rank, so the result equals the
    dense all-reduce up to the summation order of the collective. Result: (idx[:count], compact views[:count]);
    `scatter_dense` expands it to the dense (P, k) tensors. Works with gloo on CPU (tests) and NCCL on the GPUs."""

    def __init__(self, bucket: GradBucket, group=None, world: int = None):
        import torch.distributed as dist
        self.bucket, self.group = bucket, group
        self.world = world if world is not None else dist.get_world_size(group)
        dev = bucket.flat.device
        self.cuda = dev.type == "cuda"
        P = bucket.P
        self.Ppad = (P + 3) // 4 * 4
        self._touched = torch.zeros(self.Ppad, dtype=torch.uint8, device=dev)
        self.touched = self._touched[:P]
        self.gathered = torch.empty(self.world * self.Ppad, dtype=torch.uint8, device=dev)
        self.compact = torch.empty(P * bucket.floats_per_gaussian, dtype=torch.float32, device=dev)
        self.sizes = [(k, fn(bucket.M, bucket.F)) for k, fn in BUCKET_LAYOUT]
        self.count_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.count_host = torch.zeros(1, dtype=torch.int32)
        if self.cuda:
            from . import capi
            self.capi = capi
            self.count_host = self.count_host.pin_memory()
            self.idx_buf = torch.zeros(P + 1, dtype=torch.int32, device=dev)
            self.scratch = torch.empty(capi.exchange_scratch_bytes(P) + 256, dtype=torch.uint8, device=dev)
        else:  # host-logic path for the gloo tests: the same steps with torch ops
            self.arange = torch.arange(P, dtype=torch.int64, device=dev)
            self.idx_buf = torch.zeros(P + 1, dtype=torch.int64, device=dev)
        self.cap = P          # first step: no guess yet
        self.last_rows = 0
        self.fallbacks = 0
        self._pending = None

    def begin_step(self):
        """Zero the flags; the caller then runs the backward passes with capi.set_touched_mask(self.touched)."""
        self._touched.zero_()

    def _compact_views(self, cap):
        views, off = {}, 0
        for k, n in self.sizes:
            views[k] = self.compact[off:off + cap * n].view(cap, n)
            off += cap * n
        return views, off

    def exchange(self):
        """Queues the whole exchange; call finish() to get (idx, views)."""
        import torch.distributed as dist
        P, cap = self.bucket.P, self.cap
        dist.all_gather_into_tensor(self.gathered, self._touched, group=self.group)
        views, off = self._compact_views(cap)
        idx = self.idx_buf[:cap]
        if self.cuda:
            # csrc/exchange.cu: union + scan + ranked index writes (3 small launches), one row-gather launch
            self.capi.exchange_compact(self.gathered, self.world, P, self.Ppad, cap, self.idx_buf, self.count_dev,
                                       self.scratch)
            self.capi.exchange_rows(False, [self.bucket.views[k] for k, _ in self.sizes], [views[k] for k, _ in self.sizes],
                                    [n for _, n in self.sizes], self.idx_buf, self.count_dev, cap)
        else:
            union = self.gathered.view(self.world, self.Ppad)[:, :P].amax(0)
            pos = torch.cumsum(union, 0, dtype=torch.int32)
            self.count_dev.copy_(pos[-1:])
            # slot of Gaussian g = pos[g] - 1 if it is in the union and fits, else the spill slot `cap`
            slot = torch.where((union != 0) & (pos <= cap), pos - 1, cap).to(torch.int64)
            full = self.idx_buf[:cap + 1]
            full.zero_()
            full.scatter_(0, slot, self.arange)
            for k, n in self.sizes:
                torch.index_select(self.bucket.views[k].view(P, n), 0, idx, out=views[k])
        if cap:
            dist.all_reduce(self.compact[:off], op=dist.ReduceOp.SUM, group=self.group)
        self.count_host.copy_(self.count_dev, non_blocking=True)
        self._pending = (idx, views)

    def finish(self):
        """-> (idx, views) of the queued exchange: U sorted Gaussian indices and the summed rows, U = union size."""
        import torch.distributed as dist
        idx, views = self._pending
        self._pending = None
        if self.bucket.flat.is_cuda:
            torch.cuda.current_stream(self.bucket.flat.device).synchronize()
        U = int(self.count_host[0])
        self.last_rows = U
        P = self.bucket.P
        if U > self.cap:  # the guess was too small (every rank sees the same count): dense exchange for this step
            self.fallbacks += 1
            dist.all_reduce(self.bucket.flat, op=dist.ReduceOp.SUM, group=self.group)
            idx = torch.nonzero(self.gathered.view(self.world, self.Ppad)[:, :P].amax(0)).squeeze(1)
            views = {k: self.bucket.views[k].view(P, n).index_select(0, idx) for k, n in self.sizes}
            idx = idx.to(self.idx_buf.dtype)
        else:
            idx, views = idx[:U], {k: v[:U] for k, v in views.items()}
        self.cap = min(P, max(1024, (U + U // 4 + 1023) // 1024 * 1024))
        return idx, views

    def scatter_dense(self, idx, views):
        """Expand (idx, compact views) into the bucket's dense views (zeros elsewhere)."""
        self.bucket.flat.zero_()
        for k, n in self.sizes:
            self.bucket.views[k].view(self.bucket.P, n).index_copy_(0, idx.to(torch.int64), views[k])
        return self.bucket
