"""`simple_knn._C.distCUDA2(points) -> (P,) float32`: mean squared distance to the 3 nearest neighbours
(reference: submodules/simple-knn/ext.cpp:15-17, spatial.cu:15-26). Backed by gsr_knn (sm_100a)."""
from gsraster_b200 import torch_ext as _torch_ext

if _torch_ext.available():
    distCUDA2 = _torch_ext.load().distCUDA2
else:
    from gsraster_b200.capi import distCUDA2  # noqa: F401
