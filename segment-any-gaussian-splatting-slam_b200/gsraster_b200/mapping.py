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
    (capi.set_touched_mask); the exchange is then
      1. all-reduce(MAX) of the P-byte flags           -> union over ranks
      2. idx = nonzero(union)                           -> U sorted indices (the one host sync: U sizes step 4)
      3. gather rows idx of every bucket tensor into a compact tensor-major buffer of U rows
      4. all-reduce(SUM) of U * (11 + 3M + F) floats
    Result: (idx, compact views); `scatter_dense` expands it to the dense (P, k) tensors. Rows outside the union are
    exactly zero on every rank, so the result equals the dense all-reduce up to the summation order of the collective.
    Works with gloo on CPU (tests) and NCCL on the GPUs."""

    def __init__(self, bucket: GradBucket, group=None):
        self.bucket, self.group = bucket, group
        dev = bucket.flat.device
        self.touched = torch.zeros(bucket.P, dtype=torch.uint8, device=dev)
        self.compact = torch.empty(bucket.P * bucket.floats_per_gaussian, dtype=torch.float32, device=dev)
        self.sizes = [(k, fn(bucket.M, bucket.F)) for k, fn in BUCKET_LAYOUT]
        self.last_rows = 0

    def begin_step(self):
        """Zero the flags; the caller then runs the backward passes with capi.set_touched_mask(self.touched)."""
        self.touched.zero_()

    def exchange(self):
        import torch.distributed as dist
        dist.all_reduce(self.touched, op=dist.ReduceOp.MAX, group=self.group)
        idx = torch.nonzero(self.touched).squeeze(1)  # host sync: U
        U = int(idx.numel())
        self.last_rows = U
        views, off = {}, 0
        for k, n in self.sizes:
            dst = self.compact[off:off + U * n].view(U, n)
            torch.index_select(self.bucket.views[k].view(self.bucket.P, n), 0, idx, out=dst)
            views[k] = dst
            off += U * n
        if U:
            dist.all_reduce(self.compact[:off], op=dist.ReduceOp.SUM, group=self.group)
        return idx, views

    def scatter_dense(self, idx, views):
        """Expand (idx, compact views) into the bucket's dense views (zeros elsewhere)."""
        self.bucket.flat.zero_()
        for k, n in self.sizes:
            self.bucket.views[k].view(self.bucket.P, n).index_copy_(0, idx, views[k])
        return self.bucket
