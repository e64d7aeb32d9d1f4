"""Builds csrc/torch_binding.cpp (C++ / LibTorch host binding with the reference's operator signatures) into
lib/gsraster_b200_torch.so, linked against lib/libgsraster_b200.so ($ORIGIN rpath). Plain g++ invocation with
torch's include / library paths: no JIT cache, the .so stays in-tree and travels to the GPU box."""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(os.path.dirname(HERE), "lib")
OUT = os.path.join(LIB, "gsraster_b200_torch.so")


def build(force=False):
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(HERE, "torch_binding.cpp")
    deps = [src, os.path.join(HERE, "..", "..", "include", "gsraster_b200.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        build_harness()
        return OUT
    inc = ce.include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-w", "-DTORCH_EXTENSION_NAME=gsraster_b200_torch",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    cmd += [f"-I{p}" for p in inc] + [src, "-o", OUT, f"-L{LIB}", "-lgsraster_b200", f"-L{tlib}", "-lc10", "-lc10_cuda",
                                      "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
                                      "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
    subprocess.check_call(cmd)
    build_harness(force=True)
    return OUT


HARNESS = os.path.join(os.path.dirname(HERE), "bin", "mapping_harness")


def build_harness(force=False):
    """csrc/mapping_harness.cpp + torch_binding.cpp -> bin/mapping_harness: a pure C++ / LibTorch process driving the
    mapping hot path through the reference's operator signatures (no Python in the process)."""
    import torch
    from torch.utils import cpp_extension as ce
    srcs = [os.path.join(HERE, "mapping_harness.cpp"), os.path.join(HERE, "torch_binding.cpp")]
    deps = srcs + [os.path.join(HERE, "..", "..", "include", "gsraster_b200.h")]
    if not force and os.path.exists(HARNESS) and all(os.path.getmtime(HARNESS) >= os.path.getmtime(d) for d in deps):
        return HARNESS
    os.makedirs(os.path.dirname(HARNESS), exist_ok=True)
    inc = ce.include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    pylib = sysconfig.get_config_var("LIBDIR")
    pyver = "python%d.%d" % sys.version_info[:2]
    cmd = ["g++", "-O2", "-std=c++17", "-w", "-DTORCH_EXTENSION_NAME=gsraster_b200_harness",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    cmd += [f"-I{p}" for p in inc] + srcs + ["-o", HARNESS, f"-L{LIB}", "-lgsraster_b200", f"-L{tlib}", "-lc10", "-lc10_cuda",
                                            "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", f"-L{pylib}",
                                            f"-l{pyver}", "-L/usr/local/cuda/lib64", "-lcudart",
                                            "-Wl,-rpath,$ORIGIN/../lib", f"-Wl,-rpath,{tlib}",
                                            "-Wl,-rpath,/usr/local/cuda/lib64"]
    # (no rpath to the system library directory: its libnccl is older than the one libtorch_cuda.so was built against)
    subprocess.check_call(cmd)
    return HARNESS


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
