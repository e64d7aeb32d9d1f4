"""View-data-parallel keyframe mapping step (no reference equivalent: the reference
optimises one view per step on one GPU, scripts/gaussian_splatting.py:792-833, 887-935).

Gaussian parameters are replicated on every rank; a batch of keyframe views is split across the
ranks (longest-processing-time-first on the views' measured cost, so that ranks finish together);
each rank runs rasterizer forward+backward for its views, accumulating into ONE gradient bucket
that the backward kernels write in place; the Gaussian-parameter gradients are then summed over
ranks. Batch gradient = sum of per-view gradients.

Host logic only (view assignment, bucket layout, exchange protocol); works with the gloo backend on
CPU for the tests, peer memory over NVLink (default) or NCCL on the GPUs.
"""
from typing import Callable, Dict, List, Optional, Sequence

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


def assign_views_lpt(costs: Sequence[float], world: int) -> List[List[int]]:
    """Longest-processing-time-first assignment of the batch's views to ranks: views in descending cost order (ties by
    index), each to the currently least-loaded rank (ties by rank). Deterministic, so every rank computes the same
    assignment from the same (all-gathered) costs; `costs[v]` = the view's device time or instance count R of the
    previous iteration. A synchronous step ends with its slowest rank: with unequal views a contiguous split waits for
    max(view cost), this one for ~mean(cost) * views per rank. -> assignment[rank] = list of view indices."""
    if world <= 0:
        raise ValueError("bad world size")
    order = sorted(range(len(costs)), key=lambda v: (-float(costs[v]), v))
    load = [0.0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for v in order:
        r = min(range(world), key=lambda q: (load[q], q))
        out[r].append(v)
        load[r] += float(costs[v])
    return out


class GradBucket:
    """Flat fp32 buffer of 4*(11 + 3M + F) bytes per Gaussian (300 B at M=16, F=16) holding the six
    parameter-gradient tensors back to back (tensor-major), each exposed as a contiguous view.
    copies=2: two such buffers used alternately (flip()), so that the peers may still read the buffer of step s
    while step s+1 is being written (SparseGradExchange(peer_memory=True))."""

    def __init__(self, P: int, M: int, F: int, device, symmetric: bool = False, copies: int = 1):
        """symmetric=True allocates the buffer as torch symmetric memory (CUDA VMM, mappable into the other ranks of the
        node) so that SparseGradExchange(peer_memory=True) can read it over NVLink; every rank must then construct it."""
        self.P, self.M, self.F = P, M, F
        self.sizes = [(k, fn(M, F)) for k, fn in BUCKET_LAYOUT]
        self.floats_per_gaussian = sum(n for _, n in self.sizes)
        # every segment starts 16-byte aligned (float4 stores in the kernels): pad P up to a multiple of 4
        self.Ppad = (P + 3) // 4 * 4
        self.copy_floats = self.Ppad * self.floats_per_gaussian
        self.copies = int(copies)
        self.symmetric = bool(symmetric)
        if self.symmetric:
            import torch.distributed._symmetric_memory as symm
            self.storage = symm.empty(self.copies * self.copy_floats, dtype=torch.float32, device=torch.device(device))
            self.storage.zero_()
        else:
            self.storage = torch.zeros(self.copies * self.copy_floats, dtype=torch.float32, device=device)
        self.offsets: Dict[str, int] = {}
        off = 0
        for k, n in self.sizes:
            self.offsets[k] = off
            off += self.Ppad * n
        self.cur = 0
        self._select(0)

    def _select(self, c: int):
        self.cur = c
        self.flat = self.storage[c * self.copy_floats:(c + 1) * self.copy_floats]
        self.views: Dict[str, torch.Tensor] = {k: self.flat[self.offsets[k]:self.offsets[k] + self.P * n]
                                               for k, n in self.sizes}

    def flip(self):
        """Switch to the other copy (copies=2)."""
        self._select((self.cur + 1) % self.copies)

    @property
    def nbytes(self) -> int:
        return self.P * self.floats_per_gaussian * 4

    def shaped(self) -> Dict[str, torch.Tensor]:
        P, M, F = self.P, self.M, self.F
        shp = dict(means3D=(P, 3), sh=(P, M, 3), objects=(P, F), opacity=(P, 1), scales=(P, 3), rotations=(P, 4))
        return {k: v.view(shp[k]) for k, v in self.views.items()}


def mapping_step(view_ids: Sequence[int], grad_fn: Callable[[int, Dict[str, torch.Tensor], bool], None],
                 bucket: GradBucket, group=None, world: int = 1) -> GradBucket:
    """Dense variant: runs grad_fn(view, out_views, accumulate) for every local view -- grad_fn must WRITE the view's
    parameter gradients into `out_views` when accumulate is False and ADD when True -- then sums the
    bucket over ranks with one all-reduce. Returns the bucket (identical on every rank)."""
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
    4(11 + 3M + F) B/Gaussian all-reduce mostly moves zeros. Per step:
      begin_step()             zero this rank's step flags
      backward passes          capi.rasterize_gaussians_backward(..., out=bucket.views, ext=ex.ext()): sparse rows --
                               only the rows of the Gaussians a view reached are written (first view of the step) or
                               added to (later views), the flags say which rows are valid; no zero stream
      exchange()               union of the ranks' flags -> ascending index list idx[0:count] (count stays on the device)
                               -> for every union row the sum over the ranks that flagged it -> compact tensor-major rows
    Result: (idx[:count], compact views[:count]); `scatter_dense` expands it to dense (P, k) tensors. Rows outside the
    union are exactly zero on every rank, so the result equals the dense all-reduce up to summation order.

    Three transports:
      peer_memory=True (CUDA, one process per GPU of one NVLink node; default of bench.py): flags and bucket live in
        symmetric memory, double-buffered by step parity. ONE handshake kernel per step (gsr_exchange_signal_wait:
        publish "my step s is written", wait for every peer's), then every rank reads its peers' flags in place and
        sums the union's rows straight out of the peers' buckets in rank order (gsr_exchange_compact_peers /
        gsr_exchange_sum_rows_peers): no collective library call, no host read-back, bitwise identical sums on all
        ranks. Observing step s of a peer also proves it finished reading step s-1 (same stream), and step s+1 writes
        the other buffer, so no trailing barrier is needed.
      NCCL (CUDA, peer_memory=False): all-gather of the flags, compaction, row gather masked by this rank's own flags,
        all-reduce(SUM) of cap rows; `cap` is guessed from the previous step (1.25 x, multiple of 1024; P first) and
        finish() redoes the exchange at cap = P if the guess was too small.
      gloo (CPU tensors, tests): the same steps with torch ops."""

    def __init__(self, bucket: GradBucket, group=None, world: Optional[int] = None, peer_memory: bool = False,
                 rank: Optional[int] = None, timeout_ms: int = 20000):
        import torch.distributed as dist
        self.bucket, self.group = bucket, group
        self.world = world if world is not None else dist.get_world_size(group)
        self.rank = rank if rank is not None else (dist.get_rank(group) if dist.is_initialized() else 0)
        dev = bucket.storage.device
        self.cuda = dev.type == "cuda"
        P = bucket.P
        self.Ppad = bucket.Ppad
        self.peer = bool(peer_memory)
        self.timeout_ms = int(timeout_ms)
        self.seq = 0
        if self.peer:
            if not (self.cuda and bucket.symmetric and bucket.copies == 2):
                raise RuntimeError("peer_memory exchange needs a CUDA GradBucket(symmetric=True, copies=2)")
            if self.world > 16:
                raise RuntimeError("peer_memory exchange supports at most 16 ranks (one NVLink node)")
            import torch.distributed._symmetric_memory as symm
            gname = (group if group is not None else dist.group.WORLD).group_name
            self._touched_all = symm.empty(2 * self.Ppad, dtype=torch.uint8, device=dev)
            self._touched_all.zero_()
            self._pad = symm.empty(64, dtype=torch.int32, device=dev)   # signal pad: one uint32 slot per peer
            self._pad.zero_()
            self.flag_hdl = symm.rendezvous(self._touched_all, gname)
            self.bucket_hdl = symm.rendezvous(bucket.storage, gname)
            self.pad_hdl = symm.rendezvous(self._pad, gname)
            self.flag_base = [int(p) for p in self.flag_hdl.buffer_ptrs]
            self.bucket_base = [int(p) for p in self.bucket_hdl.buffer_ptrs]
            self.pad_ptrs = [int(p) for p in self.pad_hdl.buffer_ptrs]
            self.status = torch.zeros(1, dtype=torch.int32, device=dev)
            self.flag_hdl.barrier(channel=0, timeout_ms=self.timeout_ms)   # every rank's buffers are zeroed and mapped
        else:
            self._touched_all = torch.zeros(self.Ppad, dtype=torch.uint8, device=dev)
        self._select_flags(0)
        self.gathered = torch.empty(self.world * self.Ppad, dtype=torch.uint8, device=dev)
        self.compact = torch.empty(P * bucket.floats_per_gaussian, dtype=torch.float32, device=dev)
        self.sizes = list(bucket.sizes)
        self.count_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.count_host = torch.zeros(1, dtype=torch.int32)
        self.workspace = None
        if self.cuda:
            from . import capi
            self.capi = capi
            self.count_host = self.count_host.pin_memory()
            self.idx_buf = torch.zeros(P + 1, dtype=torch.int32, device=dev)
            self.scratch = torch.empty(capi.exchange_scratch_bytes(P) + 256, dtype=torch.uint8, device=dev)
            # persistent backward accumulation workspace: no 129-byte-per-Gaussian memset per backward call
            self.workspace = capi.new_row_workspace(P, bucket.F, dev)
            # overlap of the NCCL flag exchange with the per-Gaussian backward (see exchange()): the library records
            # blend_done after the backward blend; the side stream takes the all-gather + compaction from there
            self.side = torch.cuda.Stream(device=dev)
            self.blend_done = torch.cuda.Event()
            self.blend_done.record(torch.cuda.current_stream(dev))   # creates the handle
            self.idx_ready = torch.cuda.Event()
        else:  # host-logic path for the gloo tests: the same steps with torch ops
            self.arange = torch.arange(P, dtype=torch.int64, device=dev)
            self.idx_buf = torch.zeros(P + 1, dtype=torch.int64, device=dev)
        self.overlap = False
        self.cap = P          # NCCL transport: first step, no guess yet
        self.last_rows = 0
        self.fallbacks = 0
        self._pending = None
        self._ext_cache = {}

    def _select_flags(self, c: int):
        self.flag_cur = c
        self._touched = self._touched_all[c * self.Ppad:(c + 1) * self.Ppad] if self.peer else self._touched_all
        self.touched = self._touched[:self.bucket.P]

    def ext(self, accumulate: bool = True, overlap: bool = False, timer=None):
        """capi.Ext for the backward passes of the CURRENT step (call after begin_step()): touched flags of this step,
        sparse rows, accumulation across the views of the step, the persistent row workspace. overlap (NCCL transport,
        one view per step): also the blend-done event, so that exchange() can run the flag all-gather and the
        compaction beside the per-Gaussian backward."""
        self.overlap = bool(overlap) and not self.peer and not accumulate
        key = (self.flag_cur, bool(accumulate), self.overlap, id(timer))
        e = self._ext_cache.get(key)
        if e is None:
            e = self.capi.Ext(touched_mask=self.touched, sparse_rows=True, accumulate=accumulate,
                              blend_done=self.blend_done if self.overlap else None, row_workspace=self.workspace,
                              timer=timer)
            self._ext_cache[key] = e
        return e

    def begin_step(self):
        """Select this step's buffers (peer transport: the other copy of bucket and flags) and zero the step flags.
        The gradient bucket is NOT zeroed: its rows are valid exactly where the flags are set."""
        if self.peer:
            self.bucket.flip()
            self._select_flags(self.bucket.cur)
        self._touched.zero_()

    def _compact_views(self, cap):
        views, off = {}, 0
        for k, n in self.sizes:
            views[k] = self.compact[off:off + cap * n].view(cap, n)
            off += cap * n
        return views, off

    def exchange(self, cap: Optional[int] = None):
        """Queues the whole exchange; call finish() to get (idx, views) or result_device() for device-side consumers."""
        import torch.distributed as dist
        P = self.bucket.P
        cap = P if self.peer else (self.cap if cap is None else cap)
        views, off = self._compact_views(cap)
        idx = self.idx_buf[:cap]
        if self.peer:
            # capacity = P: the list cannot overflow, nothing here needs the count on the host (the optimizer takes
            # idx / count / views as device pointers), so the step stays asynchronous; finish() is optional
            self.seq += 1
            c = self.bucket.cur
            self.capi.exchange_signal_wait(self.pad_ptrs, self.rank, self.seq, self.timeout_ms, self.status)
            flags = [b + c * self.Ppad for b in self.flag_base]
            buckets = [b + c * self.bucket.copy_floats * 4 for b in self.bucket_base]
            self.capi.exchange_compact_peers(flags, P, cap, self.idx_buf, self.count_dev, self.scratch)
            self.capi.exchange_sum_rows_peers(buckets, [self.bucket.offsets[k] for k, _ in self.sizes],
                                              [views[k] for k, _ in self.sizes], [n for _, n in self.sizes], self.idx_buf,
                                              self.count_dev, cap, flags=flags)
            self.count_host.copy_(self.count_dev, non_blocking=True)
            self._pending = (idx, views)
            return
        if self.cuda and self.overlap:
            # The flags are final once the backward blend of the step's (only) view is done (the library recorded
            # blend_done there); the flag all-gather and the union compaction run on the side stream underneath the
            # per-Gaussian backward, and the main stream joins before it gathers the rows.
            cur = torch.cuda.current_stream(self.bucket.storage.device)
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
            # one row-gather launch; union rows this rank did not reach hold stale data in its bucket -> zeros
            self.capi.exchange_rows(False, [self.bucket.views[k] for k, _ in self.sizes], [views[k] for k, _ in self.sizes],
                                    [n for _, n in self.sizes], self.idx_buf, self.count_dev, cap, own_flags=self._touched)
        else:
            union = self.gathered.view(self.world, self.Ppad)[:, :P].amax(0)
            pos = torch.cumsum(union, 0, dtype=torch.int32)
            self.count_dev.copy_(pos[-1:])
            # slot of Gaussian g = pos[g] - 1 if it is in the union and fits, else the spill slot `cap`
            slot = torch.where((union != 0) & (pos <= cap), pos - 1, cap).to(torch.int64)
            full = self.idx_buf[:cap + 1]
            full.zero_()
            full.scatter_(0, slot, self.arange)
            own = self.touched.to(torch.bool)
            for k, n in self.sizes:
                rows = torch.index_select(self.bucket.views[k].view(P, n), 0, idx)
                views[k].copy_(torch.where(own[idx].unsqueeze(1), rows, torch.zeros_like(rows)))
        if cap:
            dist.all_reduce(self.compact[:off], op=dist.ReduceOp.SUM, group=self.group)
        self.count_host.copy_(self.count_dev, non_blocking=True)
        self._pending = (idx, views)

    def finish(self):
        """-> (idx, views) of the queued exchange: U sorted Gaussian indices and the summed rows, U = union size."""
        if self._pending is None:
            raise RuntimeError("finish() without a queued exchange()")
        idx, views = self._pending
        self._pending = None
        if self.cuda:
            torch.cuda.current_stream(self.bucket.storage.device).synchronize()
            if self.peer and int(self.status.item()) != 0:
                raise RuntimeError("peer-memory exchange: rank %d did not arrive within %d ms"
                                   % (int(self.status.item()) - 1, self.timeout_ms))
        U = int(self.count_host[0])
        self.last_rows = U
        P = self.bucket.P
        if (not self.peer) and U > self.cap:
            # the capacity guess was too small (every rank sees the same count): redo the exchange at capacity P
            self.fallbacks += 1
            self.overlap = False
            self.exchange(cap=P)
            idx, views = self._pending
            self._pending = None
            if self.cuda:
                torch.cuda.current_stream(self.bucket.storage.device).synchronize()
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

    def scatter_dense(self, idx, views, out: Optional[GradBucket] = None):
        """Expand (idx, compact views) into dense (P, k) views (zeros elsewhere) of `out` (default: the bucket)."""
        b = out if out is not None else self.bucket
        b.flat.zero_()
        for k, n in self.sizes:
            b.views[k].view(b.P, n).index_copy_(0, idx.to(torch.int64), views[k])
        return b


class ViewPipeline:
    """The views of one mapping step alternate between two CUDA streams, so that the forward pass of view j + 1 -- its
    latency-bound preprocess / depth sort / binning stages and its host read-back of num_rendered -- runs beside the
    backward pass of view j (both blend kernels leave issue slots free). Backward passes are chained by an event: they
    share the step's flags, the gradient bucket and the row workspace. The step still begins and ends on the caller's
    stream (begin() / end()), so nothing of step s + 1 overlaps step s: the optimiser step sits between them.

        vp.begin()
        for pose in views:
            with vp.view():                 # torch.cuda.stream(next stream)
                res = forward(pose)
                vp.before_backward()
                backward(pose, res, ext=accumulating ext)
                vp.after_backward()
        vp.end()
    """

    def __init__(self, device, nstreams: int = 2):
        import torch
        self.torch = torch
        self.device = device
        self.streams = [torch.cuda.Stream(device=device) for _ in range(max(1, nstreams))]
        self.ev_start = torch.cuda.Event()
        self.ev_bwd = [torch.cuda.Event() for _ in range(4)]
        self.ev_end = [torch.cuda.Event() for _ in self.streams]
        self.k = 0

    def begin(self):
        self.main = self.torch.cuda.current_stream(self.device)
        self.ev_start.record(self.main)
        self.joined = [False] * len(self.streams)
        self.prev = None
        self.k = 0

    def view(self):
        i = self.k % len(self.streams)
        self.cur, self.cur_i = self.streams[i], i
        if not self.joined[i]:
            self.cur.wait_event(self.ev_start)
            self.joined[i] = True
        return self.torch.cuda.stream(self.cur)

    def before_backward(self):
        if self.prev is not None:
            self.cur.wait_event(self.prev)

    def after_backward(self):
        ev = self.ev_bwd[self.k % len(self.ev_bwd)]
        ev.record(self.cur)
        self.prev = ev
        self.k += 1

    def end(self):
        for i, s in enumerate(self.streams):
            if self.joined[i]:
                self.ev_end[i].record(s)
                self.main.wait_event(self.ev_end[i])
