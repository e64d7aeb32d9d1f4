"""Generates the golden fixtures in this directory by running the UNMODIFIED reference CUDA build
(oracle/_ref/libref_rasterizer.so, libref_knn.so; recipe oracle/build_ref.sh) on a GPU on small
seeded scenes. The reference has no CPU path and no fixtures of its own, so
this is what pins both the CPU oracle (tests/test_oracle_golden.py, CPU) and the product
(tests/test_parity_gpu.py, GPU).

Run on the GPU box:  python tests/golden/make_golden.py gpurun_out/golden
then copy gpurun_out/golden/*.npz here. Inputs are stored with the outputs so the fixtures do not
depend on numpy's random streams.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as h  # noqa: E402
from gsraster_b200 import scenes  # noqa: E402

# name -> (P, W, H, sh_degree, view, scale_mul, bg, scale_modifier, precomp)
CASES = {
    "sh3_f16": dict(P=1500, W=96, H=64, deg=3, view=0, scale_mul=3.0, bg=(0, 0, 0), mod=1.0, precomp=False),
    "sh0_bg_mod": dict(P=1200, W=80, H=48, deg=0, view=2, scale_mul=4.0, bg=(0.3, 0.1, 0.7), mod=1.7, precomp=False),
    "precomp": dict(P=1000, W=64, H=64, deg=1, view=1, scale_mul=3.0, bg=(1, 1, 1), mod=1.0, precomp=True),
}
STATE_KEYS = ("radii", "is_used", "tiles_touched", "depths", "means2D", "conic_opacity", "rgb", "clamped",
              "point_list", "sorted_keys", "ranges", "n_contrib", "final_T", "color", "depth", "objects")


def build_case(c, seed):
    sc = scenes.make_scene(c["P"], sh_degree=c["deg"], F=16, seed=seed, W=c["W"], H=c["H"], scale_mul=c["scale_mul"])
    sc.bg = np.asarray(c["bg"], np.float32)
    sc.scale_modifier = c["mod"]
    cam = scenes.make_camera(c["W"], c["H"], c["view"])
    grads = scenes.make_upstream_grads(c["W"], c["H"], 16, seed=seed + 1)
    colors = cov = None
    if c["precomp"]:
        rng = np.random.default_rng(seed + 2)
        colors = rng.uniform(0, 1, (c["P"], 3)).astype(np.float32)
        A = rng.normal(0, 0.05, (c["P"], 3, 3))
        S = A @ A.transpose(0, 2, 1) + 1e-4 * np.eye(3)
        cov = np.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).astype(np.float32)
    return sc, cam, grads, colors, cov


def save_inputs(sc, cam, grads, colors, cov):
    d = dict(in_means3D=sc.means3D, in_scales=sc.scales, in_rotations=sc.rotations, in_opacities=sc.opacities,
             in_shs=sc.shs, in_sh_objs=sc.sh_objs, in_bg=sc.bg, in_viewmatrix=cam.viewmatrix,
             in_projmatrix=cam.projmatrix, in_campos=cam.campos,
             in_meta=np.array([cam.W, cam.H, sc.sh_degree], np.int32),
             in_fmeta=np.array([cam.tanfovx, cam.tanfovy, sc.scale_modifier], np.float32),
             in_dL_ddepth=grads[0], in_dL_dcolor=grads[1], in_dL_dobjects=grads[2])
    if colors is not None:
        d["in_colors_precomp"] = colors
        d["in_cov3D_precomp"] = cov
    return d


def load_case(path):
    """-> (scene, cam, grads, colors_precomp, cov3D_precomp, reference outputs dict)."""
    z = np.load(path)
    W, H, deg = [int(v) for v in z["in_meta"]]
    tx, ty, mod = [float(v) for v in z["in_fmeta"]]
    sc = scenes.Scene(z["in_means3D"], z["in_scales"], z["in_rotations"], z["in_opacities"], z["in_shs"],
                      z["in_sh_objs"], deg, z["in_bg"], mod)
    cam = scenes.Camera(W, H, tx, ty, z["in_viewmatrix"], z["in_projmatrix"], z["in_campos"])
    grads = (z["in_dL_ddepth"], z["in_dL_dcolor"], z["in_dL_dobjects"])
    colors = z["in_colors_precomp"] if "in_colors_precomp" in z.files else None
    cov = z["in_cov3D_precomp"] if "in_cov3D_precomp" in z.files else None
    ref = {k[4:]: z[k] for k in z.files if k.startswith("ref_")}
    return sc, cam, grads, colors, cov, ref


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    assert h.have_ref(), "oracle/_ref not built"
    for i, (name, c) in enumerate(CASES.items()):
        sc, cam, grads, colors, cov = build_case(c, 500 + 10 * i)
        ref = h.run_ref(sc, cam, grads, colors_precomp=colors, cov3D_precomp=cov)
        d = save_inputs(sc, cam, grads, colors, cov)
        vis = h.npy(ref["radii"]) > 0
        for k in STATE_KEYS + h.GRAD_KEYS:
            a = h.npy(ref[k]).copy()
            if k in ("depths", "means2D", "conic_opacity", "rgb", "clamped"):
                a[~vis] = 0  # the reference leaves culled entries uninitialised (forward.cu:211, 237, 252)
            if colors is not None and k == "clamped":
                a[:] = 0     # ... and never writes `clamped` / `rgb` when colours are precomputed
            if colors is not None and k == "rgb":
                a = colors * vis[:, None]  # (forward.cu:259-265); the blend reads colors_precomp instead
            d["ref_" + k] = a
        d["ref_R"] = np.array([ref["R"]], np.int64)
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **d)
        print(name, "R", ref["R"], "visible", int(vis.sum()))
    pts = scenes.make_points(2500, seed=77)
    out = h.npy(h.run_ref_knn(pts))
    np.savez_compressed(os.path.join(outdir, "knn.npz"), in_points=pts, ref_mean_dists=out)
    few = scenes.make_points(3, seed=78)
    np.savez_compressed(os.path.join(outdir, "knn_few.npz"), in_points=few, ref_mean_dists=h.npy(h.run_ref_knn(few)))
    print("knn ok")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(h.ROOT, "gpurun_out", "golden"))
