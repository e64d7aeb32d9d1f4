"""Host-side (Python) plumbing of the B200-native Gaussian-splatting rasterizer.

The product is the C-ABI library lib/libgsraster_b200.so (hand-written sm_100a CUDA, see
include/gsraster_b200.h). This package only binds it: `capi` = ctypes binding taking torch CUDA
tensors, `scenes` = seeded synthetic inputs, `mapping` = view-data-parallel mapping step.
The drop-in modules `diff_gaussian_rasterization` and `simple_knn` live next to this package.
"""
