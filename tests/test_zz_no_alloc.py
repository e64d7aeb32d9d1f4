"""GPU test of the allocation-free entry point gsr_knn_static (include/gsraster_b200.h; no counterpart in the reference):
same kernels as gsr_knn with the scratch in a file-scope device array.
Collected last on purpose. Called through the C ABI with raw pointers; outputs are
compared with gsr_knn's on the same points -- bit-exact, the bar of the other kNN tests -- on two calls in a row with
different sizes (the shared array keeps the previous call's contents), and the size limit is checked."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_POINTS = 262144   # GSR_KNN_STATIC_MAX_POINTS


def _lib():
    import helpers  # noqa: F401  (sys.path)
    from gsraster_b200 import capi
    L = capi.lib()
    L.gsr_knn_static.restype = C.c_int
    L.gsr_knn_static.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gsr_last_error.restype = C.c_char_p
    return L, capi


def _both(L, capi, pts):
    import torch
    P = pts.shape[0]
    out = torch.full((P,), -1.0, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.gsr_knn_static(P, C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(stream))
    assert rc == 0, L.gsr_last_error()
    ref = capi.distCUDA2(pts)   # gsr_knn with caller-provided scratch
    torch.cuda.synchronize()
    return out, ref


def test_knn_static_matches_knn_two_calls_in_a_row():
    import torch
    from gsraster_b200 import scenes
    L, capi = _lib()
    # second call smaller than the first, third at the limit: stale contents of the shared array must not matter
    for P, seed in ((100_000, 11), (2_500, 12), (MAX_POINTS, 13), (3, 14)):
        pts = torch.from_numpy(scenes.make_points(P, seed=seed)).cuda().contiguous()
        out, ref = _both(L, capi, pts)
        assert torch.equal(out.view(torch.int32), ref.view(torch.int32)), P
        assert float(out.min()) >= 0.0
    # the same input twice in a row gives the same bits
    pts = torch.from_numpy(scenes.make_points(50_000, seed=15)).cuda().contiguous()
    a, _ = _both(L, capi, pts)
    b, ref = _both(L, capi, pts)
    assert torch.equal(a.view(torch.int32), b.view(torch.int32)) and torch.equal(a.view(torch.int32), ref.view(torch.int32))


def test_knn_static_allocates_nothing_and_rejects_large_sizes():
    import torch
    from gsraster_b200 import scenes
    L, capi = _lib()
    pts = torch.from_numpy(scenes.make_points(40_000, seed=21)).cuda().contiguous()
    out = torch.empty(40_000, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    assert L.gsr_knn_static(40_000, C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(stream)) == 0  # set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        assert L.gsr_knn_static(40_000, C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(stream)) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_reserved() == reserved0 and free1 == free0   # neither torch's nor the driver's pool moved
    # beyond the limit: refused with a clear message, nothing launched
    big = torch.zeros((MAX_POINTS + 1, 3), dtype=torch.float32, device="cuda")
    outb = torch.empty(MAX_POINTS + 1, dtype=torch.float32, device="cuda")
    rc = L.gsr_knn_static(MAX_POINTS + 1, C.c_void_p(big.data_ptr()), C.c_void_p(outb.data_ptr()), C.c_void_p(stream))
    assert rc != 0 and b"static scratch limit" in L.gsr_last_error()
    assert L.gsr_knn_static(-1, None, None, None) != 0
    assert L.gsr_knn_static(0, None, None, None) == 0


def test_knn_static_under_graph_capture():
    """The reason the entry point exists: it can be captured (no allocation, no synchronisation inside the call)."""
    import torch
    from gsraster_b200 import scenes
    L, capi = _lib()
    pts = torch.from_numpy(scenes.make_points(40_000, seed=21)).cuda().contiguous()
    out = torch.empty(40_000, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    assert L.gsr_knn_static(40_000, C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(stream)) == 0  # set-up
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            rc = L.gsr_knn_static(40_000, C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(s.cuda_stream))
    assert rc == 0, L.gsr_last_error()
    out.fill_(-1.0)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.int32), capi.distCUDA2(pts).view(torch.int32))
