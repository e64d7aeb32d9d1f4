"""GPU diagnostic for distCUDA2: product vs reference simple-knn vs oracle."""
import sys
import time

import numpy as np
import torch

from helpers import npy, run_ref_knn, scenes
from gsraster_b200 import capi
from oracle import pyoracle as po

for P in (5, 3, 1000, 20000, int(sys.argv[1]) if len(sys.argv) > 1 else 1000000):
    pts = scenes.make_points(P, seed=4321 + P)
    tp = torch.from_numpy(pts).cuda()
    mine = capi.distCUDA2(tp)
    torch.cuda.synchronize()
    ref = run_ref_knn(pts)
    bad = int((npy(mine).view(np.uint32) != npy(ref).view(np.uint32)).sum())
    print(f"P={P}: mine vs ref bit mismatches {bad}", end="")
    if P <= 20000:
        orc = po.knn(pts, brute=True)
        print(f"  vs oracle brute {int((npy(mine).view(np.uint32) != orc.view(np.uint32)).sum())}", end="")
        orc2 = po.knn(pts)
        print(f"  oracle pruned vs brute {int((orc2.view(np.uint32) != orc.view(np.uint32)).sum())}", end="")
    print()
    if P >= 100000:
        for name, fn in (("mine", lambda: capi.distCUDA2(tp)), ("ref", lambda: run_ref_knn(pts))):
            fn(); torch.cuda.synchronize(); t0 = time.time()
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            print(f"   {name} {1e3 * (time.time() - t0) / 3:.2f} ms")
