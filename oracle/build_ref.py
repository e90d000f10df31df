"""Builds oracle/_ref/ref_awq.so: the reference's own AWQ dequant CUDA kernel, compiled UNMODIFIED
from the sources where they lie (/root/reference/sgl-kernel/csrc/gemm/awq_kernel.cu: a
self-contained file, torch + CUDA headers only) for sm_100a, plus oracle/ref_binding.cc.

Test infrastructure.  Runs only where /root/reference exists (the build container); the GPU box
uses the prebuilt file (oracle/_ref/ is git-ignored but travels with the gpurun snapshot).
nvcc needs ~5 minutes for torch/all.h.   python -m oracle.build_ref [--force]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/sgl-kernel/csrc/gemm/awq_kernel.cu"
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "ref_awq.so")


def available() -> bool:
    return os.path.exists(OUT)


def build(force: bool = False, verbose: bool = False) -> str:
    if not os.path.exists(REF_SRC):
        return OUT if available() else ""
    binding = os.path.join(HERE, "ref_binding.cc")
    if available() and not force and os.path.getmtime(OUT) >= max(os.path.getmtime(REF_SRC), os.path.getmtime(binding)):
        return OUT
    import torch
    from torch.utils import cpp_extension as ce

    os.makedirs(OUT_DIR, exist_ok=True)
    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    abi = f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    nvcc = os.path.join(cuda_home, "bin", "nvcc")
    o1, o2 = os.path.join(OUT_DIR, "awq_kernel.o"), os.path.join(OUT_DIR, "ref_binding.o")
    cmds = [
        [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", abi, *inc,
         "-c", REF_SRC, "-o", o1],
        ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", abi, *inc, "-isystem", os.path.join(cuda_home, "include"),
         "-c", binding, "-o", o2],
        ["/usr/bin/g++", "-shared", "-o", OUT, o1, o2, "-L", tlib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda",
         "-lc10_cuda", "-L", os.path.join(cuda_home, "lib64"), "-lcudart", f"-Wl,-rpath,{tlib}"],
    ]
    for c in cmds:
        if verbose:
            print(" ".join(c[:6]), "...", flush=True)
        subprocess.run(c, check=True)
    for o in (o1, o2):
        os.remove(o)
    return OUT


def load():
    """torch.ops.ref_sgl_kernel.awq_dequantize, or None when the prebuilt library is absent."""
    if not available():
        return None
    import torch

    torch.ops.load_library(OUT)
    return torch.ops.ref_sgl_kernel.awq_dequantize


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
