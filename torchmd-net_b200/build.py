"""In-tree build of the native code (explicit nvcc / g++ commands, no JIT cache, no torch build machinery).

    python torchmd-net_b200/build.py [--force]

Outputs (git-ignored, shipped to the GPU box with the tree):
    torchmd-net_b200/torchmdnet/lib/libtmdn.so        C ABI (include/tmdn.h): all CUDA kernels, sm_100a only
    torchmd-net_b200/torchmdnet/lib/libtmdn_torch.so  thin torch op shim over the C ABI (csrc/torch_shim.cpp)
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "torchmdnet", "lib")
LIB = os.path.join(LIBDIR, "libtmdn.so")
SHIM = os.path.join(LIBDIR, "libtmdn_torch.so")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
NVCC = os.path.join(CUDA_HOME, "bin", "nvcc")

CU_SOURCES = ["scan.cu", "neighbors.cu", "csr.cu", "csr_direct.cu", "halo.cu", "pair_prior.cu", "tensornet.cu", "et.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--threads", "4"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def _deps():
    hdr = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdr.append(os.path.join(ROOT, "include", "tmdn.h"))
    return hdr


def build_lib(force=False):
    os.makedirs(LIBDIR, exist_ok=True)
    srcs = [os.path.join(CSRC, s) for s in CU_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    if force or _newer(LIB, srcs + _deps() + [os.path.abspath(__file__)]):
        _run([NVCC] + NVCC_FLAGS + ["-shared", "-o", LIB] + srcs + ["-lcudart"])
    return LIB


def build_shim(force=False):
    import torch
    from torch.utils.cpp_extension import include_paths
    src = os.path.join(CSRC, "torch_shim.cpp")
    if not (force or _newer(SHIM, [src, os.path.join(ROOT, "include", "tmdn.h"), os.path.abspath(__file__)])):
        return SHIM
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    for p in include_paths():
        cmd += ["-isystem", p]
    cmd += ["-isystem", os.path.join(CUDA_HOME, "include"), "-isystem", sysconfig.get_paths()["include"]]
    cmd += [src, "-o", SHIM, "-L" + LIBDIR, "-ltmdn", "-L" + tlib, "-ltorch", "-ltorch_cpu", "-lc10",
            "-ltorch_cuda", "-lc10_cuda", "-L" + os.path.join(CUDA_HOME, "lib64"), "-lcudart",
            "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib, "-Wl,--no-as-needed"]
    _run(cmd)
    return SHIM


def build_all(force=False):
    return build_lib(force), build_shim(force)


if __name__ == "__main__":
    build_all("--force" in sys.argv)
    print("ok")
