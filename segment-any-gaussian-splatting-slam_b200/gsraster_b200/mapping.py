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

    def __init__(self, P: int, M: int, F: int, device, symmetric: bool = False):
        """symmetric=True allocates the buffer as torch symmetric memory (CUDA VMM, mappable into the other ranks of the
        node) so that SparseGradExchange(peer_memory=True) can read it over NVLink; every rank must then construct it."""
        self.P, self.M, self.F = P, M, F
        sizes = [(k, fn(M, F)) for k, fn in BUCKET_LAYOUT]
        self.floats_per_gaussian = sum(n for _, n in sizes)
        # every segment starts 16-byte aligned (float4 stores in the kernels): pad P up to a multiple of 4
        Ppad = (P + 3) // 4 * 4
        self.symmetric = bool(symmetric)
        if self.symmetric:
            import torch.distributed._symmetric_memory as symm
            self.flat = symm.empty(Ppad * self.floats_per_gaussian, dtype=torch.float32, device=torch.device(device))
            self.flat.zero_()
        else:
            self.flat = torch.zeros(Ppad * self.floats_per_gaussian, dtype=torch.float32, device=device)
        self.views: Dict[str, torch.Tensor] = {}
        self.offsets: Dict[str, int] = {}
        off = 0
        for k, n in sizes:
            self.views[k] = self.flat[off:off + P * n]
            self.offsets[k] = off
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

    def __init__(self, bucket: GradBucket, group=None, world: int = None, peer_memory: bool = False):
        """peer_memory=True (CUDA, one process per GPU of one NVLink node, bucket built with symmetric=True): no NCCL on
        the data path. The flags and the bucket live in symmetric memory; after a cross-rank barrier every rank reads its
        peers' flags in place, builds the same index list and sums the union's rows straight out of the peers' buckets
        in rank order (gsr_exchange_compact_peers / gsr_exchange_sum_rows_peers) -- 2 barrier kernels + 4 small kernels
        instead of two collectives, bitwise identical sums on all ranks."""
        import torch.distributed as dist
        self.bucket, self.group = bucket, group
        self.world = world if world is not None else dist.get_world_size(group)
        dev = bucket.flat.device
        self.cuda = dev.type == "cuda"
        P = bucket.P
        self.Ppad = (P + 3) // 4 * 4
        self.peer = bool(peer_memory)
        if self.peer:
            if not (self.cuda and bucket.symmetric):
                raise RuntimeError("peer_memory exchange needs a CUDA GradBucket(symmetric=True)")
            if self.world > 16:
                raise RuntimeError("peer_memory exchange supports at most 16 ranks (one NVLink node)")
            import torch.distributed._symmetric_memory as symm
            gname = (group if group is not None else dist.group.WORLD).group_name
            self._touched = symm.empty(self.Ppad, dtype=torch.uint8, device=dev)
            self._touched.zero_()
            self.flag_hdl = symm.rendezvous(self._touched, gname)
            self.bucket_hdl = symm.rendezvous(bucket.flat, gname)
            self.flag_ptrs = [int(p) for p in self.flag_hdl.buffer_ptrs]
            self.bucket_ptrs = [int(p) for p in self.bucket_hdl.buffer_ptrs]
        else:
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
            # overlap of the flag exchange with the per-Gaussian backward (see exchange()): the library records
            # blend_done after the backward blend; the side stream takes the all-gather + compaction from there
            self.side = torch.cuda.Stream(device=dev)
            self.blend_done = torch.cuda.Event()
            self.blend_done.record(torch.cuda.current_stream(dev))   # creates the handle
            self.idx_ready = torch.cuda.Event()
            self.overlap = False
        else:  # host-logic path for the gloo tests: the same steps with torch ops
            self.arange = torch.arange(P, dtype=torch.int64, device=dev)
            self.idx_buf = torch.zeros(P + 1, dtype=torch.int64, device=dev)
        self.cap = P          # first step: no guess yet
        self.last_rows = 0
        self.fallbacks = 0
        self._pending = None

    def arm(self, overlap=True):
        """Registers the flags (and, with overlap, the blend-done event) with the rasterizer library for this thread."""
        self.capi.set_touched_mask(self.touched)
        self.overlap = bool(overlap)
        self.capi.set_blend_done_event(self.blend_done if self.overlap else None)

    def disarm(self):
        self.capi.set_touched_mask(None)
        self.capi.set_blend_done_event(None)
        self.overlap = False

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
        P, cap = self.bucket.P, (self.bucket.P if self.peer else self.cap)
        views, off = self._compact_views(cap)
        idx = self.idx_buf[:cap]
        if self.peer:
            # capacity = P: the list cannot overflow, nothing here needs the count on the host (the optimizer takes
            # idx / count / views as device pointers), so the step stays asynchronous; finish() is optional
            self.flag_hdl.barrier(channel=0, timeout_ms=20000)   # every rank's backward pass is done: flags and buckets final
            self.capi.exchange_compact_peers(self.flag_ptrs, P, cap, self.idx_buf, self.count_dev, self.scratch)
            self.capi.exchange_sum_rows_peers(self.bucket_ptrs, [self.bucket.offsets[k] for k, _ in self.sizes],
                                              [views[k] for k, _ in self.sizes], [n for _, n in self.sizes], self.idx_buf,
                                              self.count_dev, cap)
            self.flag_hdl.barrier(channel=0, timeout_ms=20000)   # nobody still reads this rank's bucket / flags
            self.count_host.copy_(self.count_dev, non_blocking=True)
            self._pending = (idx, views)
            return
        if self.cuda and self.overlap:
            # The flags are final once the LAST backward blend of the step is done (the library recorded blend_done
            # there); the flag all-gather and the union compaction run on the side stream underneath the per-Gaussian
            # backward, and the main stream joins before it gathers the rows.
            cur = torch.cuda.current_stream(self.bucket.flat.device)
            self.side.wait_event(self.blend_done)
            with torch.cuda.stream(self.side):
                dist.all_gather_into_tensor(self.gathered, self._touched, group=self.group)
                self.capi.exchange_compact(self.gathered, self.world, P, self.Ppad, cap, self.idx_buf, self.count_dev,
                                           self.scratch)
                self.idx_ready.record(self.side)
            cur.wait_event(self.idx_ready)
        else:
            dist.all_gather_into_tensor(self.gathered, self._touched, group=self.group)
            if self.cuda:
                # csrc/exchange.cu: union + scan + ranked index writes (3 small launches)
                self.capi.exchange_compact(self.gathered, self.world, P, self.Ppad, cap, self.idx_buf, self.count_dev,
                                           self.scratch)
        if self.cuda:
            # one row-gather launch
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
        if self._pending is None:
            raise RuntimeError("finish() without a queued exchange()")
        idx, views = self._pending
        self._pending = None
        if self.bucket.flat.is_cuda:
            torch.cuda.current_stream(self.bucket.flat.device).synchronize()
        U = int(self.count_host[0])
        self.last_rows = U
        P = self.bucket.P
        if U > self.cap:  # the guess was too small (every rank sees the same count): dense exchange for this step
            self.fallbacks += 1
            if self.peer:  # the NCCL path needs the gathered flags (every rank takes this branch: same count everywhere)
                dist.all_gather_into_tensor(self.gathered, self._touched, group=self.group)
            dist.all_reduce(self.bucket.flat, op=dist.ReduceOp.SUM, group=self.group)
            idx = torch.nonzero(self.gathered.view(self.world, self.Ppad)[:, :P].amax(0)).squeeze(1)
            views = {k: self.bucket.views[k].view(P, n).index_select(0, idx) for k, n in self.sizes}
            idx = idx.to(self.idx_buf.dtype)
        else:
            idx, views = idx[:U], {k: v[:U] for k, v in views.items()}
        self.cap = min(P, max(1024, (U + U // 4 + 1023) // 1024 * 1024))
        return idx, views

    def result_device(self):
        """Peer-memory mode, no host synchronisation: (idx_buf, count_dev, compact views at capacity P) of the queued
        exchange; rows r < count are valid. What gsr_adam_step / FusedGaussianAdam.step(grads, idx, count) consume."""
        if not self.peer:
            raise RuntimeError("result_device() needs peer_memory=True (the NCCL path sizes its all-reduce from a guess)")
        idx, views = self._pending
        return idx, self.count_dev, views

    def scatter_dense(self, idx, views):
        """Expand (idx, compact views) into the bucket's dense views (zeros elsewhere)."""
        self.bucket.flat.zero_()
        for k, n in self.sizes:
            self.bucket.views[k].view(self.bucket.P, n).index_copy_(0, idx.to(torch.int64), views[k])
        return self.bucket
