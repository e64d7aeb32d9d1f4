"""Loader of the C++ / LibTorch host binding (csrc/torch_binding.cpp -> lib/gsraster_b200_torch.so): the
pybind module exposing rasterize_gaussians / rasterize_gaussians_backward / mark_visible / distCUDA2 with the
reference's exact signatures (diff-gaussian-rasterization/ext.cpp:15-19, simple-knn/ext.cpp:15-17)."""
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(os.path.dirname(_HERE), "lib", "gsraster_b200_torch.so")
_mod = None


def available():
    return os.path.exists(SO_PATH)


def load():
    global _mod
    if _mod is None:
        if not available():
            raise RuntimeError(f"{SO_PATH} not built: run __graft_entry__.build() (there is no CPU fallback)")
        import torch  # noqa: F401  (libtorch symbols must be loaded first)
        spec = importlib.util.spec_from_file_location("gsraster_b200_torch", SO_PATH)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod
