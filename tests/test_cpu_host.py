"""CPU-side tests (no GPU): the C-ABI library loads and exports every symbol the header declares,
the host logic (scene generator, scratch sizing, view sharding, gradient bucket) and the N>1
mapping step over gloo with world_size 2."""
import ctypes as C
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch

import helpers as h
from helpers import scenes

HEADER = os.path.join(h.ROOT, "include", "gsraster_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", src)) - {"gsr_alloc_fn"})


def test_library_exports_every_declared_symbol():
    from gsraster_b200 import capi
    assert os.path.exists(capi.LIB_PATH), "run __graft_entry__.build() first"
    L = C.CDLL(capi.LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(L, s), s
    L.gsr_version.restype = C.c_char_p
    assert b"sm_100a" in L.gsr_version()


def test_scratch_sizes_are_pure_functions_and_monotonic():
    from gsraster_b200 import capi
    L = capi.lib()
    assert L.gsr_geom_bytes(1000) == L.gsr_geom_bytes(1000)
    assert L.gsr_geom_bytes(2000) > L.gsr_geom_bytes(1000) > 0
    assert L.gsr_image_bytes(1280, 1024) >= 1280 * 1024 * 8 + 5120 * 8
    assert L.gsr_binning_bytes(10**6) >= 10**6 * 4
    assert L.gsr_binning_bytes(0) > 0
    L.gsr_binning_bytes_bound.restype = C.c_size_t
    L.gsr_binning_bytes_bound.argtypes = [C.c_int64, C.c_int, C.c_int, C.c_int]
    # the bound covers the supertile lists (<= min(R, P * supertiles) entries of 8 B) on top of point list + masks
    assert L.gsr_binning_bytes_bound(10**6, 10**5, 1280, 1024) >= L.gsr_binning_bytes(10**6) + 8 * 10**6
    assert L.gsr_binning_bytes_bound(10**9, 1000, 1280, 1024) <= L.gsr_binning_bytes(10**9) + 8 * 1000 * 80 + 4 * 4 * 80 + 4096
    L.gsr_row_workspace_bytes.restype = C.c_size_t
    assert L.gsr_row_workspace_bytes(1000, 16) >= 1000 * (32 * 4 + 1) and L.gsr_row_workspace_bytes(1000, 32) >= 1000 * (64 * 4 + 1)
    assert L.gsr_knn_scratch_bytes(5_000_000) >= 5_000_000 * 32


def test_argument_validation_without_gpu():
    """Status codes + gsr_last_error for calls that must fail before touching the device."""
    from gsraster_b200 import capi
    L = capi.lib()
    prm = capi.Params(10, 0, 1, 7, 64, 48, 0.5, 0.4, 1.0, 0, 0, None)  # F=7 unsupported
    cnt = C.c_int(0)
    null_alloc = capi.ALLOC_FN(lambda ctx, n: None)
    rc = L.gsr_forward(C.byref(prm), null_alloc, None, null_alloc, None, null_alloc, None, *([None] * 17), C.byref(cnt))
    assert rc == -3 and b"F must be" in L.gsr_last_error()
    prm.F = 16
    rc = L.gsr_forward(C.byref(prm), null_alloc, None, null_alloc, None, null_alloc, None, *([None] * 17), C.byref(cnt))
    assert rc == -1 and b"NULL" in L.gsr_last_error()
    assert L.gsr_mark_visible(-1, None, None, None, None, None) == -1
    assert L.gsr_knn(0, None, None, None, None) == 0
    with pytest.raises(RuntimeError):  # python wrapper refuses CPU tensors: no CPU fallback
        capi.rasterize_gaussians(torch.zeros(3), torch.zeros(4, 3), torch.empty(0), torch.zeros(4, 1), torch.zeros(4, 3),
                                 torch.zeros(4, 4), 1.0, torch.empty(0), torch.eye(4), torch.eye(4), 0.5, 0.4, 48, 64,
                                 torch.zeros(4, 1, 3), torch.zeros(4, 1, 16), 0, torch.zeros(3), False, False)


def test_scene_generator_is_seeded_and_shaped():
    a = scenes.make_scene(1000, sh_degree=3, F=16, seed=42)
    b = scenes.make_scene(1000, sh_degree=3, F=16, seed=42)
    assert all((getattr(a, k) == getattr(b, k)).all() for k in ("means3D", "scales", "rotations", "opacities", "shs", "sh_objs"))
    assert a.shs.shape == (1000, 16, 3) and a.sh_objs.shape == (1000, 1, 16) and a.rotations.shape == (1000, 4)
    assert np.allclose(np.linalg.norm(a.rotations, axis=1), 1, atol=1e-5)
    assert 0.02 < (a.means3D[:, 2] < 0).mean() < 0.09
    cam = scenes.make_camera(1280, 1024, 0)
    assert abs(cam.tanfovx - 0.4947550) < 1e-6 and abs(cam.tanfovy - 0.3958817) < 1e-6
    cam3 = scenes.make_camera(1280, 1024, 3)
    V = cam3.viewmatrix.T.astype(np.float64)
    assert np.allclose(V[:3, :3] @ V[:3, :3].T, np.eye(3), atol=1e-6)
    assert np.allclose(-V[:3, :3].T @ V[:3, 3], cam3.campos, atol=1e-5)


def test_shard_views_partitions():
    from gsraster_b200 import mapping
    for V in (0, 1, 7, 8, 9, 64):
        for world in (1, 2, 3, 8):
            got = [v for r in range(world) for v in mapping.shard_views(V, r, world)]
            assert got == list(range(V))
            sizes = [len(mapping.shard_views(V, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        mapping.shard_views(4, 2, 2)


def test_assign_views_lpt_balances_and_is_deterministic():
    from gsraster_b200 import mapping
    costs = [1.7, 2.6, 1.9, 2.2, 2.7, 1.8, 2.0, 2.4, 1.75, 2.65, 1.95, 2.25, 2.75, 1.85, 2.05, 2.45]
    a = mapping.assign_views_lpt(costs, 8)
    assert a == mapping.assign_views_lpt(list(costs), 8)
    assert sorted(v for r in a for v in r) == list(range(16))
    loads = [sum(costs[v] for v in r) for r in a]
    assert max(loads) <= 1.06 * (sum(costs) / 8)          # contiguous pairs would be at 1.25x the mean
    assert mapping.assign_views_lpt([3.0, 1.0, 1.0, 1.0], 2) == [[0], [1, 2, 3]]
    assert mapping.assign_views_lpt([], 3) == [[], [], []]
    with pytest.raises(ValueError):
        mapping.assign_views_lpt([1.0], 0)


def test_grad_bucket_layout():
    from gsraster_b200 import mapping
    P, M, F = 1001, 16, 16
    b = mapping.GradBucket(P, M, F, "cpu")
    assert b.floats_per_gaussian == 11 + 3 * M + F == 75 and b.nbytes == P * 300
    sh = b.shaped()
    assert sh["sh"].shape == (P, M, 3) and sh["rotations"].shape == (P, 4)
    for k, v in b.views.items():
        assert v.data_ptr() % 16 == 0 and v.is_contiguous(), k
        v.fill_(1.0)
    # views do not overlap and cover everything but the alignment padding
    assert int(b.flat.sum()) == P * 75
    b2 = mapping.GradBucket(P, M, F, "cpu", copies=2)   # double-buffered: flip() switches flat / views to the other copy
    p0 = b2.views["sh"].data_ptr()
    b2.views["sh"].fill_(2.0)
    b2.flip()
    assert b2.views["sh"].data_ptr() != p0 and float(b2.flat.sum()) == 0.0
    b2.flip()
    assert b2.views["sh"].data_ptr() == p0 and float(b2.flat.sum()) == 2.0 * P * 48


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from gsraster_b200 import mapping
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    P, M, F, V = 257, 4, 16, 5
    bucket = mapping.GradBucket(P, M, F, "cpu")

    def grad_fn(view, out, accumulate):  # deterministic fake per-view gradient
        for j, (k, t) in enumerate(out.items()):
            val = torch.full_like(t, float((view + 1) * (j + 1)))
            t.add_(val) if accumulate else t.copy_(val)

    mapping.mapping_step(mapping.shard_views(V, rank, world), grad_fn, bucket, world=world)
    q.put((rank, {k: float(v[0]) for k, v in bucket.views.items()}, float(bucket.flat.sum())))
    dist.destroy_process_group()


def test_mapping_step_gloo_world2():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    views_sum = sum(v + 1 for v in range(5))  # 15
    for rank, first, total in res:
        for j, k in enumerate(("means3D", "sh", "objects", "opacity", "scales", "rotations")):
            assert first[k] == views_sum * (j + 1), (rank, k)
    assert res[0][2] == res[1][2]


def _sparse_worker(rank, world, port, q):
    import torch.distributed as dist
    from gsraster_b200 import mapping
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    P, M, F = 5000, 4, 16
    g = torch.Generator().manual_seed(100 + rank)
    bucket = mapping.GradBucket(P, M, F, "cpu")
    ex = mapping.SparseGradExchange(bucket)
    out = []
    for step, nrows in enumerate((150, 150, 2000)):   # step 0: cap = P; step 1: guessed cap; step 2: overflow -> dense fallback
        ex.begin_step()
        # sparse-row bucket: rows are valid only where this rank's step flags are set, everything else is STALE data
        # (here: garbage on purpose) that the exchange must not pick up
        bucket.flat.fill_(777.0)
        rows = torch.randperm(P, generator=g)[:nrows + 50 * rank]  # each rank touches its own subset
        dense = torch.zeros_like(bucket.flat)
        dviews = {k: dense[bucket.offsets[k]:bucket.offsets[k] + P * n].view(P, n) for k, n in ex.sizes}
        for v in range(2):                               # two views per rank accumulate into the same rows
            sub = rows if v == 0 else rows[::2]
            for k, n in ex.sizes:
                val = torch.randn(len(sub), n, generator=g)
                tgt = bucket.views[k].view(P, n)
                first = ex.touched[sub] == 0             # first touch of the step writes, later views add
                tgt[sub] = torch.where(first.unsqueeze(1), val, tgt[sub] + val)
                dviews[k][sub] += val
            ex.touched[sub] = 1
        dist.all_reduce(dense)                       # the dense exchange of the true per-rank sums is the oracle
        ex.exchange()
        idx, views = ex.finish()
        got = mapping.GradBucket(P, M, F, "cpu")
        ex.scatter_dense(idx, views, out=got)
        out.append((int(idx.numel()), float((got.flat - dense).abs().max()), float(dense.abs().max()),
                    bool((idx[1:] > idx[:-1]).all()), ex.fallbacks, ex.cap))
    q.put((rank, out))
    dist.destroy_process_group()


def test_sparse_grad_exchange_matches_dense_allreduce_gloo_world2():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sparse_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    a, b = sorted(res)[0][1], sorted(res)[1][1]
    for sa, sb in zip(a, b):
        assert sa[0] == sb[0]                       # same union on both ranks
    assert 300 <= a[0][0] <= 350 and a[2][0] > 2500
    assert [s_[4] for s_ in a] == [0, 0, 1]         # only the third step overflowed its guessed capacity
    assert a[0][5] == 1024                          # the capacity guess shrank after the first step
    for out in (a, b):
        for U, err, mx, sorted_, _, _ in out:
            assert sorted_ and mx > 0 and err <= 1e-6 * mx, (err, mx)


def test_dropin_modules_import_and_validate_like_the_reference():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians  # noqa: F401
    import simple_knn._C as knn
    assert GaussianRasterizationSettings._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg",
                                                     "scale_modifier", "viewmatrix", "projmatrix", "sh_degree",
                                                     "campos", "prefiltered", "debug")
    assert callable(knn.distCUDA2)
    rs = GaussianRasterizationSettings(48, 64, 0.5, 0.4, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                       torch.zeros(3), False, False)
    r = GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair"):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m)
    with pytest.raises(RuntimeError):  # CPU tensors: no CPU fallback
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), sh_objs=torch.zeros(4, 1, 16),
          scales=m, rotations=torch.zeros(4, 4))


def test_libtorch_binding_loads_with_reference_names():
    from gsraster_b200 import torch_ext
    assert torch_ext.available(), "run __graft_entry__.build()"
    m = torch_ext.load()
    for n in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible", "distCUDA2"):
        assert hasattr(m, n), n
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.mark_visible(torch.zeros(4, 3), torch.eye(4), torch.eye(4))
