"""ctypes binding of include/gsraster_b200.h for torch CUDA tensors.

Mirrors RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA / markVisible of the reference's
LibTorch binding (submodules/diff-gaussian-rasterization/rasterize_points.cu:35-237) and distCUDA2
(submodules/simple-knn/spatial.cu:15-26): same argument order, same returned tuples. There is no
CPU fallback: a missing library or a non-CUDA tensor raises.
"""
import ctypes as C
import sys
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libgsraster_b200.so")

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)


class Stats(C.Structure):
    """gsr_stats: OUT record of one call."""
    _fields_ = [("num_rendered", C.c_ulonglong), ("l1_total", C.c_ulonglong), ("sync_us", C.c_ulonglong),
                ("alloc_us", C.c_ulonglong), ("launch_us", C.c_ulonglong), ("launches", C.c_int)]


class _ExtStruct(C.Structure):
    """gsr_ext (include/gsraster_b200.h): the per-call extensions. There are no mode switches in the library: what a
    call does differently from the reference is passed to that call."""
    _fields_ = [("touched_mask", C.c_void_p), ("touched_len", C.c_int), ("blend_done_event", C.c_void_p),
                ("tracking", C.c_int), ("sparse_rows", C.c_int), ("accumulate", C.c_int), ("touched_idx", C.c_void_p),
                ("touched_count", C.c_void_p), ("row_workspace", C.c_void_p), ("timer", C.c_void_p),
                ("stats", C.POINTER(Stats))]


class Params(C.Structure):
    _fields_ = [("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("F", C.c_int), ("width", C.c_int),
                ("height", C.c_int), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
                ("scale_modifier", C.c_float), ("prefiltered", C.c_int), ("debug", C.c_int),
                ("stream", C.c_void_p), ("ext", C.POINTER(_ExtStruct))]


class Timer:
    """gsr_timer: per-stage CUDA-event spans of the calls it is handed to (Ext(timer=...))."""

    def __init__(self):
        L = lib()
        L.gsr_timer_create.restype = C.c_void_p
        L.gsr_timer_destroy.argtypes = [C.c_void_p]
        L.gsr_timer_read.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        self.h = L.gsr_timer_create()
        if not self.h:
            raise RuntimeError("gsr_timer_create failed")

    def read(self):
        """-> (stage_ms, gap_ms): milliseconds per stage, and idle stream time before each stage, accumulated since the
        previous read (synchronises the recorded events)."""
        ms, gap = (C.c_float * len(STAGES))(), (C.c_float * len(STAGES))()
        _check(lib().gsr_timer_read(self.h, ms, gap))
        return dict(zip(STAGES, [float(v) for v in ms])), dict(zip(STAGES, [float(v) for v in gap]))

    def __del__(self):
        try:
            if self.h:
                lib().gsr_timer_destroy(self.h)
                self.h = None
        except Exception:
            pass


class Ext:
    """Per-call extensions of rasterize_gaussians / rasterize_gaussians_backward (gsr_ext). Keeps the tensors it
    points to alive. Fields (all optional):
      touched_mask   CUDA uint8 (P,) zeroed by the caller: flags of the Gaussians the backward blend reached
      blend_done     torch.cuda.Event (recorded once before): recorded right after the backward blend
      tracking       camera-tracking backward: only the camera gradients (camera_grads=True), no per-Gaussian outputs
      sparse_rows    backward writes only the rows of touched Gaussians (needs touched_mask), nothing else is written
      accumulate     sparse rows of Gaussians already flagged are added to (views of one step into one bucket)
      touched_idx / touched_count   CUDA int32 (P,) / (1,): list of the Gaussians written
      row_workspace  persistent CUDA uint8 buffer of row_workspace_bytes(P, F), all zero (kept zero by the library)
      timer          capi.Timer: per-stage events of this call
    After the call `ext.stats` holds the gsr_stats record of it."""

    def __init__(self, touched_mask=None, blend_done=None, tracking=False, sparse_rows=False, accumulate=False,
                 touched_idx=None, touched_count=None, row_workspace=None, timer=None):
        self.stats = Stats()
        self._keep = (touched_mask, blend_done, touched_idx, touched_count, row_workspace, timer)
        s = _ExtStruct()
        if touched_mask is not None:
            if touched_mask.dtype != torch.uint8 or not touched_mask.is_cuda or not touched_mask.is_contiguous():
                raise RuntimeError("touched mask must be a contiguous CUDA uint8 tensor")
            s.touched_mask, s.touched_len = touched_mask.data_ptr(), touched_mask.numel()
        if blend_done is not None:
            h = blend_done.cuda_event
            h = getattr(h, "value", h)
            if not h:
                raise RuntimeError("blend-done event has no CUDA handle yet: record it once before passing it")
            s.blend_done_event = h
        s.tracking, s.sparse_rows, s.accumulate = int(bool(tracking)), int(bool(sparse_rows)), int(bool(accumulate))
        for name, t in (("touched_idx", touched_idx), ("touched_count", touched_count)):
            if t is not None:
                if t.dtype != torch.int32 or not t.is_cuda or not t.is_contiguous():
                    raise RuntimeError(name + " must be a contiguous CUDA int32 tensor")
                setattr(s, name, t.data_ptr())
        if row_workspace is not None:
            if row_workspace.dtype != torch.uint8 or not row_workspace.is_cuda or not row_workspace.is_contiguous():
                raise RuntimeError("row_workspace must be a contiguous CUDA uint8 tensor")
            s.row_workspace = row_workspace.data_ptr()
        self.workspace_bytes = 0 if row_workspace is None else row_workspace.numel()
        if timer is not None:
            s.timer = timer.h
        s.stats = C.pointer(self.stats)
        self.struct = s

    def ptr(self, P=None, F=None):
        if P is not None and self.workspace_bytes and self.workspace_bytes < row_workspace_bytes(P, F):
            raise RuntimeError("row_workspace is smaller than row_workspace_bytes(P, F)")
        return C.pointer(self.struct)


def row_workspace_bytes(P, F):
    L = lib()
    L.gsr_row_workspace_bytes.restype = C.c_size_t
    L.gsr_row_workspace_bytes.argtypes = [C.c_int, C.c_int]
    return int(L.gsr_row_workspace_bytes(int(P), int(F)))


def new_row_workspace(P, F, device):
    """Zeroed persistent accumulation workspace for Ext(row_workspace=...)."""
    return torch.zeros(row_workspace_bytes(P, F) + 256, dtype=torch.uint8, device=device)



_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.gsr_last_error.restype = C.c_char_p
        L.gsr_version.restype = C.c_char_p
        for n in ("gsr_exchange_scratch_bytes", "gsr_geom_bytes", "gsr_image_bytes", "gsr_binning_bytes", "gsr_knn_scratch_bytes"):
            getattr(L, n).restype = C.c_size_t
        L.gsr_binning_bytes.argtypes = [C.c_int64]
        L.gsr_forward.argtypes = [C.POINTER(Params), ALLOC_FN, C.c_void_p, ALLOC_FN, C.c_void_p, ALLOC_FN,
                                  C.c_void_p] + [C.c_void_p] * 17 + [C.POINTER(C.c_int)]
        L.gsr_backward.argtypes = [C.POINTER(Params), C.c_int] + [C.c_void_p] * 29
        L.gsr_backward_pose.argtypes = [C.POINTER(Params), C.c_int] + [C.c_void_p] * 32
        L.gsr_mark_visible.argtypes = [C.c_int] + [C.c_void_p] * 5
        L.gsr_knn.argtypes = [C.c_int] + [C.c_void_p] * 4
        L.gsr_export_state.argtypes = [C.POINTER(Params), C.c_int] + [C.c_void_p] * 14
        L.gsr_count_pairs.argtypes = [C.POINTER(Params), C.c_int] + [C.c_void_p] * 3 + [C.POINTER(C.c_ulonglong)]
        L.gsr_exchange_scratch_bytes.argtypes = [C.c_int]
        L.gsr_exchange_compact.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p]
        L.gsr_exchange_rows.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                        C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.gsr_last_forward_stats.argtypes = [C.POINTER(C.c_ulonglong)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError(f"gsraster_b200 error {rc}: {lib().gsr_last_error().decode()}")


def _ptr(t):
    """Device pointer of a tensor, NULL for empty tensors (reference: 0-element tensor == absent)."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError("gsraster_b200: expected a CUDA tensor (no CPU fallback)")
    return t.data_ptr()


def _f32c(t):
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise RuntimeError("gsraster_b200: expected float32")
    return t.contiguous()


class _Resizer:
    """torch-backed scratch allocator (reference: resizeFunctional, rasterize_points.cu:27-33).

    The three scratch buffers of a step are hundreds of MB and are requested with the same sizes every step; taking
    them from torch's caching allocator lets smaller requests split the cached blocks, which ends in sporadic
    cudaMalloc calls (~100 ms) in the middle of a mapping loop. So each (device, role) keeps a few buffers of its own
    and hands one out again once nothing else references it (Python references and autograd saved tensors both count)."""

    _pools = {}
    _POOL_MAX = 8
    # Diagnostics (tests/gpu_sanitize.py; compute-sanitizer is closed on the GPU pool): the slack behind the requested
    # size is filled with a pattern before the kernels run and verified in take(): a write past the end of a scratch
    # buffer raises instead of corrupting a neighbour.
    debug_canary = False

    def __init__(self, device, role=""):
        self.device = device
        # keyed by the stream of the call as well: "nothing references it any more" only says that the work using a
        # buffer has been ENQUEUED, which orders it with later work on the same stream only (views of a step that
        # alternate between two streams, mapping.ViewPipeline, must not inherit each other's buffers)
        self.key = (str(device), role, torch.cuda.current_stream(device).cuda_stream)
        self.t = torch.empty(0, dtype=torch.uint8, device=device)
        self.cb = ALLOC_FN(self._alloc)

    def _alloc(self, _ctx, n):
        p = self._alloc_impl(n)
        if _Resizer.debug_canary:
            self.req = int(n)
            self.t[self.req:].fill_(0xA5)
        return p

    def _alloc_impl(self, n):
        n = int(n) + 256
        pool = _Resizer._pools.setdefault(self.key, [])
        best = None
        for i in range(len(pool)):
            t = pool[i]
            # free <=> only the pool list and the local name refer to it, and no autograd node saved it
            if sys.getrefcount(t) <= 3 and t._use_count() <= 1 and t.numel() >= n:
                if best is None or t.numel() < best.numel():
                    best = t   # the smallest free buffer that fits: a mapping loop cycles through keyframes whose R differ,
                               # and the pool settles on buffers large enough for all of them instead of re-allocating
        t = None
        if best is not None and best.numel() <= 4 * n + (64 << 20):
            self.t = best
            return best.data_ptr()
        best = None
        self.t = torch.empty(n + n // 4, dtype=torch.uint8, device=self.device)  # head-room: R varies from view to view
        pool.append(self.t)
        if len(pool) > _Resizer._POOL_MAX:
            for i in range(len(pool)):  # drop the oldest unreferenced buffer
                if sys.getrefcount(pool[i]) <= 2 and pool[i]._use_count() <= 1:
                    del pool[i]
                    break
            else:
                del pool[0]
        return self.t.data_ptr()

    def take(self):
        """Hand the buffer to the caller and break the self -> callback -> bound method -> self cycle, so that the
        buffer's life ends with the caller's reference and not at the next cyclic-GC run."""
        t, self.t, self.cb = self.t, None, None
        if _Resizer.debug_canary and getattr(self, "req", None) is not None and t is not None and t.numel():
            torch.cuda.synchronize(t.device)
            if not bool((t[self.req:] == 0xA5).all()):
                raise RuntimeError("gsraster_b200: a kernel wrote past the end of the %s scratch buffer (%d bytes requested)"
                                   % (self.key[1], self.req))
        return t


def _stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, sh_objs, degree,
                        campos, prefiltered, debug, ext=None):
    """-> (num_rendered, depth, color, objects, radii, is_used, geomBuffer, binningBuffer, imgBuffer).
    `ext` (extension): a capi.Ext with the per-call options / timer / statistics of this call."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("gsraster_b200: expected CUDA tensors (there is no CPU fallback)")
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    if sh_objs is None or sh_objs.numel() == 0:
        raise RuntimeError("sh_objs must be provided (the reference dereferences it unconditionally)")
    F = sh_objs.numel() // max(P, 1)
    out_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_objects = torch.empty((F, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    is_used = torch.empty((P,), dtype=torch.bool, device=dev)
    geo, binb, img = _Resizer(dev, "geom"), _Resizer(dev, "binning"), _Resizer(dev, "image")
    rendered = C.c_int(0)
    if P == 0:
        for t in (out_depth, out_color, out_objects):
            t.zero_()
        return 0, out_depth, out_color, out_objects, radii, is_used, geo.take(), binb.take(), img.take()
    M = sh.size(1) if (sh is not None and sh.numel() != 0) else 0
    prm = Params(P, int(degree), int(M), int(F), W, H, float(tan_fovx), float(tan_fovy), float(scale_modifier),
                 int(bool(prefiltered)), int(bool(debug)), _stream_ptr(), ext.ptr(P, F) if ext is not None else None)
    keep = [_f32c(t) for t in (background, means3D, sh, sh_objs, colors, opacity, scales, rotations, cov3D_precomp,
                               viewmatrix, projmatrix, campos)]
    with torch.cuda.device(dev):
        rc = lib().gsr_forward(C.byref(prm), geo.cb, None, binb.cb, None, img.cb, None, *[_ptr(t) for t in keep],
                               _ptr(out_depth), _ptr(out_color), _ptr(out_objects), _ptr(radii), _ptr(is_used),
                               C.byref(rendered))
    bufs = (geo.take(), binb.take(), img.take())
    _check(rc)
    return (rendered.value, out_depth, out_color, out_objects, radii, is_used) + bufs


def _alloc_grads(P, M, F, o):
    return dict(means3D=torch.empty((P, 3), **o), means2D=torch.empty((P, 3), **o), depths=torch.empty((P, 1), **o),
                colors=torch.empty((P, 3), **o), objects=torch.empty((P, F), **o), conic=torch.empty((P, 6), **o),
                opacity=torch.empty((P, 1), **o), cov3D=torch.empty((P, 6), **o), sh=torch.empty((P, M, 3), **o),
                scales=torch.empty((P, 3), **o), rotations=torch.empty((P, 4), **o))


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_depth,
                                 dL_dout_color, sh, dL_dout_objects, sh_objs, degree, campos, geomBuffer, R,
                                 binningBuffer, imageBuffer, debug, out=None, camera_grads=False, ext=None):
    """`ext` (extension): a capi.Ext with the per-call options of this call (touched flags, sparse rows, tracking, ...).
    With Ext(tracking=True) no per-Gaussian gradient is computed: the returned 9-tuple holds empty tensors and only
    `last_camera_grads` is meaningful (camera_grads must be True).
    `camera_grads` (extension, gsr_backward_pose): also compute dL/dviewmatrix (4,4), dL/dprojmatrix (4,4),
    dL/dcampos (3,) -- same flat layout as the inputs -- left in `rasterize_gaussians_backward.last_camera_grads`.
    `out` (extension): dict of preallocated contiguous fp32 tensors keyed like the local `g` below
    (e.g. views into one flat all-reduce bucket, see mapping.GradBucket) that the kernels write
    directly; missing keys are allocated here.
    -> (dL_dmeans2D, dL_dcolors, dL_dobjects, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
    dL_drotations), exactly the reference's 9-tuple (rasterize_points.cu:215)."""
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if (sh is not None and sh.numel() != 0) else 0
    F = sh_objs.numel() // max(P, 1)
    o = dict(device=dev, dtype=torch.float32)
    tracking = ext is not None and bool(ext.struct.tracking)
    if tracking:
        if not camera_grads:
            raise RuntimeError("Ext(tracking=True) computes only the camera gradients: pass camera_grads=True")
        P_alloc = 0
    else:
        P_alloc = P
    g = _alloc_grads(P_alloc, M, F, o)
    if out and not tracking:
        for k, t in out.items():
            if t.numel() != g[k].numel() or t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda:
                raise RuntimeError(f"out[{k!r}] must be a contiguous CUDA float32 tensor of {g[k].numel()} elements")
            g[k] = t.view(g[k].shape)
    if P != 0:
        prm = Params(P, int(degree), int(M), int(F), W, H, float(tan_fovx), float(tan_fovy), float(scale_modifier),
                     0, int(bool(debug)), _stream_ptr(), ext.ptr(P, F) if ext is not None else None)
        keep = [_f32c(t) for t in (background, means3D, sh, sh_objs, colors, scales, rotations, cov3D_precomp,
                                   viewmatrix, projmatrix, campos)]
        ups = [_f32c(t) for t in (dL_dout_depth, dL_dout_color, dL_dout_objects)]
        args = [C.byref(prm), int(R), *[_ptr(t) for t in keep], _ptr(radii), _ptr(geomBuffer), _ptr(binningBuffer),
                _ptr(imageBuffer), *[_ptr(t) for t in ups], _ptr(g["means2D"]), _ptr(g["conic"]), _ptr(g["opacity"]),
                _ptr(g["depths"]), _ptr(g["colors"]), _ptr(g["objects"]), _ptr(g["means3D"]), _ptr(g["cov3D"]),
                _ptr(g["sh"]), _ptr(g["scales"]), _ptr(g["rotations"])]
        with torch.cuda.device(dev):
            if camera_grads:
                cg = (torch.empty((4, 4), **o), torch.empty((4, 4), **o), torch.empty((3,), **o))
                rc = lib().gsr_backward_pose(*args, *[_ptr(t) for t in cg])
                rasterize_gaussians_backward.last_camera_grads = cg
                if ext is not None:
                    ext.camera_grads = cg
            else:
                rc = lib().gsr_backward(*args)
        _check(rc)
        if (scales is None or scales.numel() == 0) and not tracking:
            g["scales"].zero_()
            g["rotations"].zero_()
    rasterize_gaussians_backward.last_internal = (g["conic"], g["depths"])
    return (g["means2D"], g["colors"], g["objects"], g["opacity"], g["means3D"], g["cov3D"], g["sh"], g["scales"],
            g["rotations"])


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P:
        m, v, pj = _f32c(means3D), _f32c(viewmatrix), _f32c(projmatrix)
        with torch.cuda.device(means3D.device):
            _check(lib().gsr_mark_visible(P, _ptr(m), _ptr(v), _ptr(pj), _ptr(present), _stream_ptr()))
    return present


def distCUDA2(points):
    P = points.size(0)
    pts = _f32c(points)
    out = torch.zeros((P,), dtype=torch.float32, device=points.device)
    if P:
        scratch = torch.empty(lib().gsr_knn_scratch_bytes(P) + 256, dtype=torch.uint8, device=points.device)
        with torch.cuda.device(points.device):
            _check(lib().gsr_knn(P, _ptr(pts), _ptr(out), _ptr(scratch), _stream_ptr()))
    return out


def export_state(P, D, M, F, W, H, R, geomBuffer, binningBuffer, imageBuffer):
    """Parity/diagnostic dump (see gsr_export_state)."""
    dev = geomBuffer.device
    T = ((W + 15) // 16) * ((H + 15) // 16)
    o = dict(device=dev)
    st = dict(tiles_touched=torch.zeros(P, dtype=torch.int32, **o), depths=torch.zeros(P, dtype=torch.float32, **o),
              means2D=torch.zeros((P, 2), dtype=torch.float32, **o),
              conic_opacity=torch.zeros((P, 6), dtype=torch.float32, **o),
              rgb=torch.zeros((P, 3), dtype=torch.float32, **o), clamped=torch.zeros((P, 3), dtype=torch.uint8, **o),
              point_list=torch.zeros(max(R, 1), dtype=torch.int32, **o),
              sorted_keys=torch.zeros(max(R, 1), dtype=torch.int64, **o),
              ranges=torch.zeros((T, 2), dtype=torch.int32, **o), n_contrib=torch.zeros(H * W, dtype=torch.int32, **o),
              final_T=torch.zeros(H * W, dtype=torch.float32, **o))
    prm = Params(P, D, M, F, W, H, 1.0, 1.0, 1.0, 0, 0, _stream_ptr())
    with torch.cuda.device(dev):
        _check(lib().gsr_export_state(C.byref(prm), int(R), _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                                      *[_ptr(st[k]) for k in ("tiles_touched", "depths", "means2D", "conic_opacity",
                                                              "rgb", "clamped", "point_list", "sorted_keys", "ranges",
                                                              "n_contrib", "final_T")]))
    st["point_list"] = st["point_list"][:R]
    st["sorted_keys"] = st["sorted_keys"][:R]
    return st


STAGES = ("preprocess", "depth_sort_offsets", "binning", "render_fwd", "render_bwd", "preprocess_bwd")


def last_forward_stats():
    """-> dict(R, l1_total) of the last forward call on this thread."""
    c = (C.c_ulonglong * 6)()
    _check(lib().gsr_last_forward_stats(c))
    return dict(R=int(c[0]), l1_total=int(c[1]), sync_us=int(c[2]), alloc_us=int(c[3]), launch_us=int(c[4]))


def exchange_scratch_bytes(P):
    return int(lib().gsr_exchange_scratch_bytes(int(P)))


def exchange_compact(flags, nranks, P, Ppad, cap, idx, count, scratch):
    """Union + compaction of all-gathered touched flags (gsr_exchange_compact); everything stays on the device."""
    with torch.cuda.device(flags.device):
        _check(lib().gsr_exchange_compact(flags.data_ptr(), int(nranks), int(P), int(Ppad), int(cap), idx.data_ptr(),
                                          count.data_ptr(), scratch.data_ptr(), _stream_ptr()))


def exchange_rows(scatter, dense, compact, floats_per_row, idx, count, cap, own_flags=None):
    """Row gather (dense -> compact) or scatter (compact -> dense) of the bucket tensors (gsr_exchange_rows).
    own_flags (gather): uint8 (P,) -- dense rows whose flag is 0 are gathered as zeros (sparse-row buckets)."""
    n = len(dense)
    dp = (C.c_void_p * n)(*[t.data_ptr() for t in dense])
    cp = (C.c_void_p * n)(*[t.data_ptr() for t in compact])
    fp = (C.c_int * n)(*[int(v) for v in floats_per_row])
    with torch.cuda.device(idx.device):
        _check(lib().gsr_exchange_rows(int(scatter), n, dp, cp, fp, idx.data_ptr(), count.data_ptr(), int(cap),
                                       None if own_flags is None else own_flags.data_ptr(), _stream_ptr()))


def exchange_compact_peers(flag_ptrs, P, cap, idx, count, scratch):
    """gsr_exchange_compact_peers: flag_ptrs = the ranks' flag buffers as mapped in this process (ints)."""
    n = len(flag_ptrs)
    fp = (C.c_void_p * n)(*[int(p) for p in flag_ptrs])
    L = lib()
    L.gsr_exchange_compact_peers.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p]
    with torch.cuda.device(idx.device):
        _check(L.gsr_exchange_compact_peers(fp, n, int(P), int(cap), idx.data_ptr(), count.data_ptr(), scratch.data_ptr(),
                                            _stream_ptr()))


def exchange_signal_wait(pad_ptrs, rank, seq, timeout_ms, status):
    """gsr_exchange_signal_wait: publish `seq` to every peer's signal pad and wait for theirs (one launch)."""
    n = len(pad_ptrs)
    pp = (C.c_void_p * n)(*[int(p) for p in pad_ptrs])
    L = lib()
    L.gsr_exchange_signal_wait.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_int, C.c_void_p, C.c_void_p]
    with torch.cuda.device(status.device):
        _check(L.gsr_exchange_signal_wait(pp, n, int(rank), int(seq), int(timeout_ms), status.data_ptr(), _stream_ptr()))


def exchange_sum_rows_peers(bucket_ptrs, tensor_offsets, compact, floats_per_row, idx, count, cap, flags=None):
    """gsr_exchange_sum_rows_peers: compact rows = sum over ranks of the union's rows, read from the peers' buckets
    (flags: the ranks' flag buffers; a rank contributes to a row only where its flag is set)."""
    n, k = len(bucket_ptrs), len(compact)
    bp = (C.c_void_p * n)(*[int(p) for p in bucket_ptrs])
    fl = None if flags is None else (C.c_void_p * n)(*[int(p) for p in flags])
    off = (C.c_longlong * k)(*[int(o) for o in tensor_offsets])
    cp = (C.c_void_p * k)(*[t.data_ptr() for t in compact])
    fp = (C.c_int * k)(*[int(v) for v in floats_per_row])
    L = lib()
    L.gsr_exchange_sum_rows_peers.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    with torch.cuda.device(idx.device):
        _check(L.gsr_exchange_sum_rows_peers(bp, fl, n, k, off, cp, fp, idx.data_ptr(), count.data_ptr(), int(cap),
                                             _stream_ptr()))


def count_pairs(P, D, M, F, W, H, R, geomBuffer, binningBuffer, imageBuffer):
    """-> dict(evaluated, blended, warp_blended, instances) of a finished forward pass."""
    prm = Params(P, D, M, F, W, H, 1.0, 1.0, 1.0, 0, 0, _stream_ptr())
    c = (C.c_ulonglong * 4)()
    with torch.cuda.device(geomBuffer.device):
        _check(lib().gsr_count_pairs(C.byref(prm), int(R), _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer), c))
    return dict(evaluated=int(c[0]), blended=int(c[1]), warp_blended=int(c[2]), instances=int(c[3]))


def last_launch_count():
    return int(lib().gsr_last_launch_count())
