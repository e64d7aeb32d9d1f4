"""ctypes front-end of the CPU oracle (oracle/oracle.cpp). TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
LIB64 = os.path.join(HERE, "_build", "liboracle_f64.so")   # the same source with -DORC_REAL=double (pose-gradient FD checks)


def build(force=False, f64=False):
    src = os.path.join(HERE, "oracle.cpp")
    out = LIB64 if f64 else LIB
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-shared", "-fPIC"] +
                              (["-DORC_REAL=double"] if f64 else []) + [src, "-o", out])
    return out


class _Scene(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("P", "D", "M", "F", "W", "H")] + \
               [(n, C.c_float) for n in ("tan_fovx", "tan_fovy", "scale_modifier")] + [("prefiltered", C.c_int)] + \
               [(n, C.c_void_p) for n in ("background", "means3D", "shs", "sh_objs", "colors_precomp", "opacities",
                                          "scales", "rotations", "cov3D_precomp", "viewmatrix", "projmatrix",
                                          "campos")]


_STATE_FIELDS = ("depths", "clamped", "radii", "means2D", "cov3D", "conic_opacity", "rgb", "tiles_touched",
                 "point_offsets", "is_used", "final_T", "n_contrib", "out_depth", "out_color", "out_objects",
                 "ranges", "keys", "point_list")
_GRAD_FIELDS = ("dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_ddepths", "dL_dcolors", "dL_dobjects", "dL_dmean3D",
                "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot")


class _State(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _STATE_FIELDS]


class _Grads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _GRAD_FIELDS]


_lib = {}


def lib(f64=False):
    if f64 not in _lib:
        L = C.CDLL(build(f64=f64))
        L.orc_preprocess.restype = C.c_int64
        L.orc_bin.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_higher_msb.restype = C.c_uint32
        _lib[f64] = L
    return _lib[f64]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class Problem:
    """Holds contiguous host arrays + the C struct (keeps them alive)."""

    def __init__(self, scene, cam, colors_precomp=None, cov3D_precomp=None, prefiltered=False, f64=False):
        """f64: double-precision build of the oracle; the camera arrays of `cam` are then taken as float64 as given
        (a finite-difference step of 1e-6 must survive), everything else is widened from the fp32 scene."""
        self.f64 = bool(f64)
        rt = np.float64 if f64 else np.float32
        _f32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=rt)   # noqa: F811 (shadows the module helper)
        self.P = scene.means3D.shape[0]
        self.F = scene.sh_objs.shape[-1]
        self.W, self.H = cam.W, cam.H
        use_sh = colors_precomp is None
        self.M = scene.shs.shape[1] if use_sh else 0
        self.D = scene.sh_degree
        self.arr = dict(
            background=_f32(scene.bg), means3D=_f32(scene.means3D), shs=_f32(scene.shs) if use_sh else None,
            sh_objs=_f32(scene.sh_objs.reshape(self.P, self.F)), colors_precomp=_f32(colors_precomp),
            opacities=_f32(scene.opacities.reshape(-1)),
            scales=_f32(scene.scales) if cov3D_precomp is None else None,
            rotations=_f32(scene.rotations) if cov3D_precomp is None else None,
            cov3D_precomp=_f32(cov3D_precomp), viewmatrix=_f32(cam.viewmatrix), projmatrix=_f32(cam.projmatrix),
            campos=_f32(cam.campos))
        self.c = _Scene(self.P, self.D, self.M, self.F, self.W, self.H, cam.tanfovx, cam.tanfovy,
                        scene.scale_modifier, int(prefiltered), *[_p(self.arr[k]) for k in
                                                                  ("background", "means3D", "shs", "sh_objs",
                                                                   "colors_precomp", "opacities", "scales",
                                                                   "rotations", "cov3D_precomp", "viewmatrix",
                                                                   "projmatrix", "campos")])


def forward(prob):
    """Runs K1..K4 of the oracle. Returns dict of numpy arrays (state + outputs) and R."""
    L = lib(prob.f64)
    rt = np.float64 if prob.f64 else np.float32
    P, W, H, F = prob.P, prob.W, prob.H, prob.F
    T = ((W + 15) // 16) * ((H + 15) // 16)
    st = dict(depths=np.zeros(P, rt), clamped=np.zeros(3 * P, np.uint8), radii=np.zeros(P, np.int32),
              means2D=np.zeros((P, 2), rt), cov3D=np.zeros((P, 6), rt),
              conic_opacity=np.zeros((P, 6), rt), rgb=np.zeros((P, 3), rt),
              tiles_touched=np.zeros(P, np.uint32), point_offsets=np.zeros(P, np.uint32),
              is_used=np.zeros(P, np.uint8), final_T=np.zeros(H * W, rt),
              n_contrib=np.zeros(H * W, np.uint32), out_depth=np.zeros((1, H, W), rt),
              out_color=np.zeros((3, H, W), rt), out_objects=np.zeros((F, H, W), rt),
              ranges=np.zeros((T, 2), np.uint32), keys=np.zeros(0, np.uint64), point_list=np.zeros(0, np.uint32))
    cs = _State(*[_p(st[k]) for k in _STATE_FIELDS])
    R = int(L.orc_preprocess(C.byref(prob.c), C.byref(cs)))
    st["keys"] = np.zeros(max(R, 1), np.uint64)
    st["point_list"] = np.zeros(max(R, 1), np.uint32)
    cs = _State(*[_p(st[k]) for k in _STATE_FIELDS])
    L.orc_bin(C.byref(prob.c), C.byref(cs), C.c_int64(R))
    L.orc_render(C.byref(prob.c), C.byref(cs))
    st["keys"] = st["keys"][:R]
    st["point_list"] = st["point_list"][:R]
    st["R"] = R
    st["_c"] = cs
    return st


def backward(prob, st, dL_ddepth, dL_dcolor, dL_dobjects):
    L = lib(prob.f64)
    rt = np.float64 if prob.f64 else np.float32
    _f32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=rt)   # noqa: F811
    P, F, M = prob.P, prob.F, prob.M
    g = dict(dL_dmean2D=np.zeros((P, 3), rt), dL_dconic=np.zeros((P, 6), rt),
             dL_dopacity=np.zeros((P, 1), rt), dL_ddepths=np.zeros((P, 1), rt),
             dL_dcolors=np.zeros((P, 3), rt), dL_dobjects=np.zeros((P, F), rt),
             dL_dmean3D=np.zeros((P, 3), rt), dL_dcov3D=np.zeros((P, 6), rt),
             dL_dsh=np.zeros((P, max(M, 0), 3), rt), dL_dscale=np.zeros((P, 3), rt),
             dL_drot=np.zeros((P, 4), rt))
    cg = _Grads(*[_p(g[k]) for k in _GRAD_FIELDS])
    if prob.arr["shs"] is None:
        cg.dL_dsh = None
    if prob.arr["scales"] is None:
        cg.dL_dscale = None
        cg.dL_drot = None
    a, b, c = _f32(dL_ddepth), _f32(dL_dcolor), _f32(dL_dobjects)
    L.orc_render_backward(C.byref(prob.c), C.byref(st["_c"]), _p(a), _p(b), _p(c), C.byref(cg))
    L.orc_preprocess_backward(C.byref(prob.c), C.byref(st["_c"]), C.byref(cg))
    return g


def mark_visible(means3D, viewmatrix):
    m = _f32(means3D)
    v = _f32(viewmatrix)
    out = np.zeros(m.shape[0], np.uint8)
    lib().orc_mark_visible(m.shape[0], _p(m), _p(v), _p(out))
    return out.astype(bool)


def knn(points, brute=False):
    p = _f32(points)
    out = np.zeros(p.shape[0], np.float32)
    (lib().orc_knn_bruteforce if brute else lib().orc_knn)(p.shape[0], _p(p), _p(out))
    return out


def num_threads():
    return int(lib().orc_num_threads())
