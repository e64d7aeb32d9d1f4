"""TEST / BASELINE INFRASTRUCTURE ONLY -- never imported by the product package.

Thin ctypes driver of the UNMODIFIED reference CUDA rasterizer built by oracle/build_ref.sh
(oracle/_ref/libref_rasterizer*.so, shim oracle/ref_shim.cu). It reproduces what the reference's
LibTorch binding does around CudaRasterizer::Rasterizer::forward/backward
(/root/reference/submodules/diff-gaussian-rasterization/rasterize_points.cu:35-216): five
zero-filled outputs and three freshly grown scratch tensors per forward, eleven torch.zeros
gradient tensors per backward. Used by bench.py --impl reference (the "reference CUDA rasterizer
built for sm_100 on the same GPU" baseline of BASELINE.json) and by the parity tests.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)


def lib_path(F=16):
    return os.path.join(HERE, "_ref", "libref_rasterizer.so" if F == 16 else "libref_rasterizer_f%d.so" % F)


def available(F=16):
    return os.path.exists(lib_path(F))


class _Grow:
    """resizeFunctional (rasterize_points.cu:27-33): a uint8 tensor resized on demand."""

    def __init__(self, torch, device):
        self.torch, self.device = torch, device
        self.t = torch.empty(0, dtype=torch.uint8, device=device)
        self.cb = ALLOC_FN(self._alloc)

    def _alloc(self, _ctx, n):
        self.t = self.torch.empty(int(n), dtype=self.torch.uint8, device=self.device)
        return self.t.data_ptr()


class RefRasterizer:
    def __init__(self, F=16):
        L = C.CDLL(lib_path(F))
        L.ref_forward.argtypes = [ALLOC_FN, C.c_void_p] * 3 + [C.c_int] * 3 + [C.c_void_p, C.c_int, C.c_int] + \
            [C.c_void_p] * 6 + [C.c_float] + [C.c_void_p] * 5 + [C.c_float, C.c_float, C.c_int] + [C.c_void_p] * 5 + [C.c_int]
        L.ref_backward.argtypes = [C.c_int] * 4 + [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_float] + \
            [C.c_void_p] * 5 + [C.c_float, C.c_float] + [C.c_void_p] * 18 + [C.c_int]
        assert L.ref_num_objects() == F
        self.L, self.F = L, F

    def forward(self, d, W, H, tanfovx, tanfovy, degree, scale_modifier=1.0):
        """d: dict of CUDA tensors (bg, means3D, shs, sh_objs, opacities, scales, rotations, viewmatrix,
        projmatrix, campos). -> (R, depth, color, objects, radii, is_used, (geom, binning, image))."""
        import torch
        dev = d["means3D"].device
        P, F = d["means3D"].size(0), self.F
        M = d["shs"].size(1)
        f = dict(dtype=torch.float32, device=dev)
        depth = torch.full((1, H, W), 0.0, **f)
        color = torch.full((3, H, W), 0.0, **f)
        objects = torch.full((F, H, W), 0.0, **f)
        radii = torch.full((P,), 0, dtype=torch.int32, device=dev)
        is_used = torch.full((P,), False, dtype=torch.bool, device=dev)
        ga, ba, ia = _Grow(torch, dev), _Grow(torch, dev), _Grow(torch, dev)
        p = lambda t: t.data_ptr()
        R = self.L.ref_forward(ga.cb, None, ba.cb, None, ia.cb, None, P, degree, M, p(d["bg"]), W, H,
                               p(d["means3D"]), p(d["shs"]), p(d["sh_objs"]), None, p(d["opacities"]),
                               p(d["scales"]), scale_modifier, p(d["rotations"]), None, p(d["viewmatrix"]),
                               p(d["projmatrix"]), p(d["campos"]), tanfovx, tanfovy, 0, p(depth), p(color),
                               p(objects), p(radii), p(is_used), 0)
        if R < 0:
            raise RuntimeError("reference forward failed")
        return R, depth, color, objects, radii, is_used, (ga, ba, ia)

    def backward(self, d, W, H, tanfovx, tanfovy, degree, R, radii, bufs, dL_ddepth, dL_dcolor, dL_dobjects,
                 scale_modifier=1.0):
        import torch
        P, F = d["means3D"].size(0), self.F
        M = d["shs"].size(1)
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=d["means3D"].device)
        g = dict(dL_dmeans3D=z(P, 3), dL_dmeans2D=z(P, 3), dL_ddepths=z(P, 1), dL_dcolors=z(P, 3),
                 dL_dobjects=z(P, F), dL_dconic=z(P, 6), dL_dopacity=z(P, 1), dL_dcov3D=z(P, 6),
                 dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
        p = lambda t: t.data_ptr()
        ga, ba, ia = bufs
        rc = self.L.ref_backward(P, degree, M, R, p(d["bg"]), W, H, p(d["means3D"]), p(d["shs"]), p(d["sh_objs"]),
                                 None, p(d["scales"]), scale_modifier, p(d["rotations"]), None, p(d["viewmatrix"]),
                                 p(d["projmatrix"]), p(d["campos"]), tanfovx, tanfovy, p(radii), p(ga.t), p(ba.t),
                                 p(ia.t), p(dL_ddepth), p(dL_dcolor), p(dL_dobjects), p(g["dL_dmeans2D"]),
                                 p(g["dL_dconic"]), p(g["dL_dopacity"]), p(g["dL_ddepths"]), p(g["dL_dcolors"]),
                                 p(g["dL_dobjects"]), p(g["dL_dmeans3D"]), p(g["dL_dcov3D"]), p(g["dL_dsh"]),
                                 p(g["dL_dscales"]), p(g["dL_drotations"]), 0)
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return g
