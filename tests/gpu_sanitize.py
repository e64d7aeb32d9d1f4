"""GPU diagnostic (not a pytest): one small pass over every product kernel, meant to be run under
`compute-sanitizer --tool memcheck` (one tool per gpurun call)."""
import torch

import helpers as h
from helpers import scenes
from gsraster_b200 import capi, optim

for (P, W, H, F, deg) in ((3000, 200, 136, 16, 3), (1500, 131, 77, 32, 1), (5, 64, 64, 16, 0)):
    sc = scenes.make_scene(P, sh_degree=deg, F=F, seed=3, W=W, H=H)
    cam = scenes.make_camera(W, H, 1)
    grads = scenes.make_upstream_grads(W, H, F)
    mask = torch.zeros(P, dtype=torch.uint8, device="cuda")
    out = h.run_mine(sc, cam, grads, ext=capi.Ext(touched_mask=mask))   # forward, export (full pull), backward
    print("scene", P, W, H, F, "R", out["R"], "touched", int(mask.sum()))
    Ppad = (P + 3) // 4 * 4
    flags = torch.zeros(2, Ppad, dtype=torch.uint8, device="cuda")
    flags[0, :P] = mask
    flags[1, :P] = (torch.rand(P, device="cuda") < 0.1).to(torch.uint8)
    idx = torch.zeros(P + 1, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    scratch = torch.empty(capi.exchange_scratch_bytes(P) + 256, dtype=torch.uint8, device="cuda")
    cap = max(P // 2, 1)
    capi.exchange_compact(flags.view(-1), 2, P, Ppad, cap, idx, cnt, scratch)
    sizes = (3, 3 * (deg + 1) ** 2, F, 1, 3, 4)
    dense = [torch.randn(P, n, device="cuda") for n in sizes]
    comp = [torch.empty(cap, n, device="cuda") for n in sizes]
    capi.exchange_rows(False, dense, comp, sizes, idx, cnt, cap)
    capi.exchange_rows(True, dense, comp, sizes, idx, cnt, cap)
    M = (deg + 1) ** 2
    raw = dict(xyz=torch.randn(P, 3, device="cuda"), f_dc=torch.randn(P, 1, 3, device="cuda"),
               f_rest=torch.randn(P, M - 1, 3, device="cuda"), opacity=torch.randn(P, 1, device="cuda"),
               scaling=torch.randn(P, 3, device="cuda"), rotation=torch.randn(P, 4, device="cuda"),
               objects=torch.randn(P, 1, F, device="cuda"))
    fa = optim.FusedGaussianAdam(raw)
    g = dict(means3D=dense[0], sh=dense[1].view(P, M, 3), objects=dense[2], opacity=dense[3], scales=dense[4],
             rotations=dense[5])
    fa.step(g)
    gc = dict(means3D=comp[0], sh=comp[1], objects=comp[2], opacity=comp[3], scales=comp[4], rotations=comp[5])
    fa.step(gc, idx=idx[:cap], count=cnt)
    torch.cuda.synchronize()
pts = scenes.make_points(2000, seed=5)
capi.distCUDA2(torch.from_numpy(pts).cuda())
torch.cuda.synchronize()
print("sanitize pass done")
