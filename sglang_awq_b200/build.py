"""In-tree build of the native libraries (sm_100a only).

    python -m sglang_awq_b200.build            # build what is stale
    python -m sglang_awq_b200.build --force

Outputs (git-ignored, but they travel to the GPU box with the gpurun snapshot):
    sglang_awq_b200/lib/libawq_b200.so      C ABI of include/awq_b200.h (nvcc only, no torch)
    sglang_awq_b200/sgl_kernel/common_ops.so  TORCH_LIBRARY_FRAGMENT(sgl_kernel, ...) shim over it

nvcc cross-compiles without a GPU, so this runs in the CPU-only build container.
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "build")
LIBDIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIBDIR, "libawq_b200.so")
TORCH_LIB = os.path.join(PKG, "sgl_kernel", "common_ops.so")

CU_SOURCES = ["api.cu", "awq_dequant.cu", "awq_repack.cu", "awq_gemv.cu", "awq_gemv_direct.cu", "awq_gemv_cluster.cu", "awq_gemv_persist.cu", "awq_gemv_tc.cu", "awq_gemv_sub.cu", "awq_gemm_tcgen05.cu", "awq_allreduce.cu", "awq_moe.cu"]
HEADERS = ["common.cuh", "gemv_common.cuh", "host_common.h", "tcgen05.cuh", os.path.join(ROOT, "include", "awq_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def _run(cmd, log_path=None):
    p = subprocess.run(cmd, capture_output=True, text=True)
    if log_path:
        with open(log_path, "w") as f:
            f.write(" ".join(cmd) + "\n" + p.stdout + p.stderr)
    if p.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + p.stdout + p.stderr)
        raise RuntimeError(f"command failed: {cmd[0]} ... ({p.returncode})")
    return p


def build_c_abi(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = _nvcc()
    hdrs = [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    jobs = []
    objs = []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        _run([nvcc, *NVCC_FLAGS, "-I", os.path.join(ROOT, "include"), "-c", s, "-o", o], log_path=o + ".log")
        return o

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            for o in ex.map(compile_one, jobs):
                if verbose:
                    print("compiled", os.path.relpath(o, ROOT))
    if force or jobs or _stale(LIB, objs):
        _run([nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"])
        if verbose:
            print("linked", os.path.relpath(LIB, ROOT))
    return LIB


def build_torch_binding(force: bool = False, verbose: bool = False) -> str:
    """g++ only: the shim contains no device code, it forwards tensors' pointers to the C ABI."""
    src = os.path.join(CSRC, "torch_binding.cc")
    if not (force or _stale(TORCH_LIB, [src, LIB, os.path.join(ROOT, "include", "awq_b200.h")])):
        return TORCH_LIB
    import torch
    from torch.utils import cpp_extension as ce

    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    inc += ["-isystem", os.path.join(cuda_home, "include")]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [
        gxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=common_ops",
        f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
        "-I", os.path.join(ROOT, "include"), *inc, src, "-o", TORCH_LIB,
        "-L", LIBDIR, "-lawq_b200", "-L", tlib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda", "-lc10_cuda",
        "-L", os.path.join(cuda_home, "lib64"), "-lcudart",
        "-Wl,-rpath,$ORIGIN/../lib", f"-Wl,-rpath,{tlib}",
    ]
    _run(cmd, log_path=os.path.join(OBJ, "torch_binding.log"))
    if verbose:
        print("built", os.path.relpath(TORCH_LIB, ROOT))
    return TORCH_LIB


def build_all(force: bool = False, verbose: bool = False):
    build_c_abi(force, verbose)
    build_torch_binding(force, verbose)


def ptxas_report() -> str:
    out = []
    for src in CU_SOURCES:
        log = os.path.join(OBJ, src.replace(".cu", ".o") + ".log")
        if os.path.exists(log):
            out.append(open(log).read())
    return "\n".join(out)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--no-torch", action="store_true")
    a = ap.parse_args()
    build_c_abi(a.force, verbose=True)
    if not a.no_torch:
        build_torch_binding(a.force, verbose=True)
