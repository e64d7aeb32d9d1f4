"""Capacity-managed Gaussian map (csrc/exchange.cu through the C ABI).

The reference grows the map every frame by `torch.cat` of all seven parameter tensors and their fourteen Adam moment
tensors (`add_from_pcd2_tensor` -> `densification_postfix` -> `cat_tensors_to_optimizer`, scene/gaussian_model.py:180-229,
:472-534: a full copy of the map per frame) and prunes it every 200 iterations by boolean-mask indexing of the same 21
tensors (`prune_large_and_transparent` -> `prune_points` -> `_prune_optimizer`, :624-636, :443-470). Here the 21 tensors
live in buffers of a fixed capacity:
  append   writes the new rows behind the current count (nothing else moves), zero moments for them;
  prune    flags (gsr_prune_flags) -> ascending list of the kept rows (gsr_exchange_compact) -> one row-gather launch
           per 7 tensors into the second buffer set (gsr_exchange_rows), then the two sets swap. Order-preserving, so the
           result equals the reference's `tensor[mask]` bit for bit.
There is no CPU fallback."""
import ctypes as C
from typing import Dict

import torch

from . import capi
from .optim import GROUPS


def row_shapes(M, F):
    return dict(xyz=(3,), f_dc=(1, 3), f_rest=(M - 1, 3), opacity=(1,), scaling=(3,), rotation=(4,), objects=(1, F))


class GaussianStore:
    def __init__(self, capacity: int, M: int, F: int, device="cuda"):
        self.capacity, self.M, self.F, self.count = int(capacity), int(M), int(F), 0
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("gsraster_b200.store: CUDA only (no CPU fallback)")
        self.shapes = row_shapes(M, F)
        alloc = lambda: {k: torch.zeros((self.capacity,) + s, dtype=torch.float32, device=self.device)
                         for k, s in self.shapes.items()}
        # [set][kind][group]; kind: 0 parameter, 1 exp_avg, 2 exp_avg_sq
        self.sets = [[alloc(), alloc(), alloc()], [alloc(), alloc(), alloc()]]
        self.front = 0
        Ppad = (self.capacity + 3) // 4 * 4
        self.keep = torch.zeros(Ppad, dtype=torch.uint8, device=self.device)
        self.idx = torch.zeros(self.capacity + 1, dtype=torch.int32, device=self.device)
        self.count_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.scratch = torch.empty(capi.exchange_scratch_bytes(self.capacity) + 256, dtype=torch.uint8, device=self.device)

    def _views(self, kind) -> Dict[str, torch.Tensor]:
        return {k: t[:self.count] for k, t in self.sets[self.front][kind].items()}

    @property
    def params(self):
        """The raw parameters `_xyz, _features_dc, ...` of the current map: views of the first `count` rows."""
        return self._views(0)

    @property
    def exp_avg(self):
        return self._views(1)

    @property
    def exp_avg_sq(self):
        return self._views(2)

    def append(self, new: Dict[str, torch.Tensor]):
        """New Gaussians (dict keyed like GROUPS, n rows each) behind the current ones; their Adam moments start at zero
        (cat_tensors_to_optimizer, gaussian_model.py:472-496). Nothing already in the map is copied."""
        n = int(new["xyz"].shape[0])
        if self.count + n > self.capacity:
            raise RuntimeError(f"gsraster_b200.store: capacity {self.capacity} exceeded ({self.count} + {n})")
        cur = self.sets[self.front]
        for k in GROUPS:
            src = new[k].reshape((n,) + self.shapes[k])
            cur[0][k][self.count:self.count + n].copy_(src)
            cur[1][k][self.count:self.count + n].zero_()
            cur[2][k][self.count:self.count + n].zero_()
        self.count += n
        return self.count

    def prune_flags(self, min_opacity, extent=None):
        """keep flags (uint8, device) of prune_large_and_transparent(min_opacity, extent) for the current rows."""
        L = capi.lib()
        L.gsr_prune_flags.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        p = self.sets[self.front][0]
        with torch.cuda.device(self.device):
            capi._check(L.gsr_prune_flags(p["opacity"].data_ptr(), p["scaling"].data_ptr(), self.count, float(min_opacity),
                                          -1.0 if extent is None else 0.1 * float(extent), self.keep.data_ptr(),
                                          capi._stream_ptr()))
        return self.keep[:self.count]

    def compact(self, keep=None):
        """Keeps the rows whose flag is non-zero (`keep`: uint8 device tensor of `count` elements, default: the flags of
        the last prune_flags call), order preserved, for all 21 tensors. One host read-back (the new count), as the
        reference's boolean-mask indexing has. -> new count."""
        P = self.count
        if P == 0:
            return 0
        if keep is not None and keep.data_ptr() != self.keep.data_ptr():
            self.keep[:P].copy_(keep.to(torch.uint8))
        Ppad = (P + 3) // 4 * 4
        self.keep[P:Ppad].zero_()
        capi.exchange_compact(self.keep, 1, P, Ppad, P, self.idx, self.count_dev, self.scratch)
        src, dst = self.sets[self.front], self.sets[self.front ^ 1]
        widths = [int(torch.tensor(self.shapes[k]).prod()) for k in GROUPS]
        for kind in range(3):
            capi.exchange_rows(2, [src[kind][k] for k in GROUPS], [dst[kind][k] for k in GROUPS], widths, self.idx,
                               self.count_dev, P)
        self.front ^= 1
        self.count = int(self.count_dev.item())
        return self.count

    def prune_large_and_transparent(self, min_opacity, extent=None):
        """scene/gaussian_model.py:624-636 on this store. -> new count."""
        self.prune_flags(min_opacity, extent)
        return self.compact()
