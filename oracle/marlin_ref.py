"""Baseline C (VERDICT r1 "the Marlin bar"): the reference's AWQ-Marlin linear on this GPU.

Test / bench infrastructure.  build() compiles the reference's OWN Marlin sources, unmodified, where they lie
(/root/reference/sgl-kernel/csrc/gemm/marlin/gptq_marlin.cu ~16 min of nvcc, awq_marlin_repack.cu) for sm_100a plus
oracle/ref_marlin_binding.cc into oracle/_ref/ref_marlin.so (git-ignored, travels to the GPU box).  The Python half of
the reference path is restated here from python/sglang/srt/layers/quantization/marlin_utils.py:
    get_scale_perms :301-309, marlin_permute_scales :312-323, marlin_zero_points :354-375,
    awq_to_marlin_zero_points :378-398, marlin_make_workspace :260-268, should_use_atomic_add_reduce :439-461,
    apply_awq_marlin_linear :535-600 (forward_context None branch), pack_cols / unpack_cols utils.py:337-385,
and AWQMarlinLinearMethod.process_weights_after_loading awq.py:540-580.     python -m oracle.marlin_ref [--force]
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = "/root/reference/sgl-kernel/csrc/gemm/marlin"
REF_INC = "/root/reference/sgl-kernel/include"
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "ref_marlin.so")
UINT4_ID = 1125899906843648  # sgl_kernel.scalar_type.scalar_types.uint4.id (ScalarType.uint(4, None), scalar_type.py:330)


def available() -> bool:
    return os.path.exists(OUT)


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(REF_DIR, "gptq_marlin.cu"), os.path.join(REF_DIR, "awq_marlin_repack.cu")]
    if not all(os.path.exists(s) for s in srcs):
        return OUT if available() else ""
    binding = os.path.join(HERE, "ref_marlin_binding.cc")
    if available() and not force and os.path.getmtime(OUT) >= max(os.path.getmtime(p) for p in srcs + [binding]):
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
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(OUT_DIR, os.path.basename(s).replace(".cu", ".o"))
        objs.append(o)
        pre = os.environ.get("AWQ_MARLIN_PREBUILT_DIR")  # objects of an earlier identical compile (16 min saved)
        if pre and os.path.exists(os.path.join(pre, os.path.basename(o))):
            subprocess.run(["cp", os.path.join(pre, os.path.basename(o)), o], check=True)
            continue
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
               "-Xcompiler", "-fPIC", abi, *inc, "-I", REF_INC, "-I", REF_DIR, "-c", s, "-o", o]
        if verbose:
            print(" ".join(cmd[:6]), "...", os.path.basename(s), flush=True)
        procs.append(subprocess.Popen(cmd))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed on the reference Marlin sources")
    o2 = os.path.join(OUT_DIR, "ref_marlin_binding.o")
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", abi, *inc, "-isystem", os.path.join(cuda_home, "include"),
                    "-c", binding, "-o", o2], check=True)
    subprocess.run(["/usr/bin/g++", "-shared", "-o", OUT, *objs, o2, "-L", tlib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda",
                    "-lc10_cuda", "-L", os.path.join(cuda_home, "lib64"), "-lcudart", f"-Wl,-rpath,{tlib}"], check=True)
    for o in objs + [o2]:
        os.remove(o)
    return OUT


_loaded = False


def load() -> bool:
    global _loaded
    if not available():
        return False
    if not _loaded:
        import torch

        torch.ops.load_library(OUT)
        _loaded = True
    return True


def _scale_perms():
    scale_perm = [i + 8 * j for i in range(8) for j in range(8)]
    scale_perm_single = [2 * i + j for i in range(4) for j in (0, 1, 8, 9, 16, 17, 24, 25)]
    return scale_perm, scale_perm_single


def marlin_permute_scales(s, size_k: int, size_n: int, group_size: int):
    scale_perm, scale_perm_single = _scale_perms()
    if group_size < size_k and group_size != -1:
        s = s.reshape((-1, len(scale_perm)))[:, scale_perm]
    else:
        s = s.reshape((-1, len(scale_perm_single)))[:, scale_perm_single]
    return s.reshape((-1, size_n)).contiguous()


def awq_to_marlin_zero_points(q_zp_packed, size_k: int, size_n: int, num_bits: int = 4):
    import torch

    assert num_bits == 4
    dev = q_zp_packed.device
    p = q_zp_packed.cpu().numpy().astype(np.uint32)
    q = np.zeros((size_k, size_n), dtype=np.uint32)                       # unpack_cols
    for i in range(8):
        q[:, i::8] = (p >> np.uint32(4 * i)) & np.uint32(0xF)
    undo = np.argsort(np.array([0, 2, 4, 6, 1, 3, 5, 7]))
    q = q.reshape((-1, 8))[:, undo].ravel().reshape((-1, size_n))       # undo the AWQ interleave
    scale_perm, _ = _scale_perms()                                         # marlin_zero_points
    q = q.reshape((-1, len(scale_perm)))[:, scale_perm]
    q = q.reshape((-1, 8))[:, np.array([0, 2, 4, 6, 1, 3, 5, 7])].ravel().reshape((-1, size_n))
    out = np.zeros((size_k, size_n // 8), dtype=np.uint32)               # pack_cols
    for i in range(8):
        out |= q[:, i::8] << np.uint32(4 * i)
    return torch.from_numpy(out.astype(np.int32)).to(dev).contiguous()


def prepare(qweight, scales, qzeros, group_size: int = 128):
    """AWQMarlinLinearMethod.process_weights_after_loading (awq.py:540-580) on reference-layout CUDA tensors."""
    import torch

    assert load(), "oracle/_ref/ref_marlin.so is not built"
    K, N = qweight.shape[0], qweight.shape[1] * 8
    dev = qweight.device
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    return {
        "K": K, "N": N,
        "workspace": torch.zeros(sms, dtype=torch.int, device=dev),
        "qweight": torch.ops.ref_marlin.awq_marlin_repack(qweight, K, N, 4),
        "scales": marlin_permute_scales(scales, K, N, group_size),
        "qzeros": awq_to_marlin_zero_points(qzeros, K // group_size, N, 4),
        "g_idx": torch.empty(0, dtype=torch.int, device=dev), "perm": torch.empty(0, dtype=torch.int, device=dev),
    }


def linear(x, w, bias=None):
    """apply_awq_marlin_linear (marlin_utils.py:535-600)."""
    import torch

    x2 = x.reshape(-1, x.shape[-1])
    M, K, N = x2.shape[0], w["K"], w["N"]
    use_atomic_add = not (N >= 2048 or K < 2048)  # should_use_atomic_add_reduce on sm_100 (fp16/bf16 both native)
    y = torch.ops.ref_marlin.gptq_marlin_gemm(x2, w["qweight"], w["scales"], w["qzeros"], w["g_idx"], w["perm"], w["workspace"],
                                              UINT4_ID, M, N, K, use_atomic_add, True)
    if bias is not None:
        y.add_(bias)
    return y.reshape(x.shape[:-1] + (N,))


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
