"""Generate golden vectors from the REFERENCE's own AWQ dequantisation (run in the build
container only: /root/reference does not exist on the GPU box).

Two reference functions are executed, unmodified, on CPU tensors:
  (1) awq_dequantize_decomposition   python/sglang/srt/layers/quantization/awq_triton.py:342-368
      (module loaded by file path; it only needs torch + triton to import)
  (2) awq_dequantize_torch + reverse_awq_order   sgl-kernel/tests/test_awq_dequant.py:9-57
      (extracted with `ast` because the test file imports the compiled sgl_kernel)
Both must agree bit-for-bit; the output bit patterns are stored as uint16 in small .npz files.

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz
"""
import ast
import importlib.util
import os
import sys

import numpy as np
import torch

REF = os.environ.get("AWQ_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def load_decomposition():
    path = os.path.join(REF, "python/sglang/srt/layers/quantization/awq_triton.py")
    spec = importlib.util.spec_from_file_location("ref_awq_triton", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.awq_dequantize_decomposition


def load_test_oracle():
    path = os.path.join(REF, "sgl-kernel/tests/test_awq_dequant.py")
    tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef)
            and n.name in ("reverse_awq_order", "awq_dequantize_torch")]
    ns = {"torch": torch}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns["awq_dequantize_torch"]


# (name, K, N, G, dtype, full_range, signed_scales, tiny_scales)
CASES = [
    ("f16_g128_full", 512, 256, 128, torch.float16, True, True, False),
    ("bf16_g128_full", 512, 256, 128, torch.bfloat16, True, True, False),
    ("f16_gK_ref_style", 256, 128, 256, torch.float16, False, False, False),   # reference test style: G=K, words >= 0
    ("bf16_gK_ref_style", 128, 576, 128, torch.bfloat16, False, False, False),
    ("f16_g32_tail", 96, 72, 32, torch.float16, True, True, False),            # packed cols 9: not a multiple of 16
    ("f16_g64", 128, 64, 64, torch.float16, True, False, False),
    ("f16_subnormal", 256, 64, 128, torch.float16, True, True, True),          # products land in fp16 subnormals
    ("bf16_g128_wide", 256, 1024, 128, torch.bfloat16, True, True, False),
]


def main():
    deco = load_decomposition()
    test_oracle = load_test_oracle()
    g = torch.Generator().manual_seed(20261104)
    for name, K, N, G, dt, full, signed, tiny in CASES:
        lo, hi = (-(2**31), 2**31) if full else (0, 2**31 - 1)
        qweight = torch.randint(lo, hi, (K, N // 8), generator=g, dtype=torch.int64).to(torch.int32)
        qzeros = torch.randint(lo, hi, (K // G, N // 8), generator=g, dtype=torch.int64).to(torch.int32)
        scales = torch.rand(K // G, N, generator=g) * (2e-6 if tiny else 0.02) + (1e-7 if tiny else 0.005)
        if signed:
            scales = scales * (torch.randint(0, 2, scales.shape, generator=g) * 2 - 1)
        scales = scales.to(dt)
        a = deco(qweight, scales, qzeros)
        b = test_oracle(qweight, scales, qzeros, G)
        assert a.dtype == dt and b.dtype == dt, (a.dtype, b.dtype)
        abits = a.contiguous().view(torch.int16).numpy().view(np.uint16)
        bbits = b.contiguous().view(torch.int16).numpy().view(np.uint16)
        assert np.array_equal(abits, bbits), f"{name}: the two reference oracles disagree"
        np.savez_compressed(
            os.path.join(HERE, f"dequant_{name}.npz"),
            qweight=qweight.numpy(), qzeros=qzeros.numpy(),
            scales=scales.view(torch.int16).numpy().view(np.uint16),
            out=abits, group_size=np.int32(G), dtype=np.int32(0 if dt == torch.float16 else 1))
        print(f"{name}: K={K} N={N} G={G} {dt} ok, {abits.nbytes} B")

    # permutation KAT (sgl-kernel/tests/test_awq_dequant.py:11, test_marlin_repack.py:42)
    kat = []
    for p in range(8):
        v = 0xF << (4 * p)
        v = v - 2**32 if v >= 2**31 else v
        word = torch.tensor([[v]], dtype=torch.int32)
        out = deco(word, torch.ones(1, 8, dtype=torch.float16), torch.zeros(1, 1, dtype=torch.int32))
        kat.append(int(torch.nonzero(out[0]).item()))
    np.savez(os.path.join(HERE, "nibble_order_kat.npz"), nibble_to_column=np.array(kat, dtype=np.int32))
    print("nibble p -> logical column:", kat)


if __name__ == "__main__":
    sys.exit(main())
