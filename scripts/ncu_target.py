"""Small, fixed launch sequence for ncu captures (one kernel family per run).
usage: python scripts/ncu_target.py {gemv|gemm|dequant|linear_g} [M] [shape]      PROBE_GROUP=64|32 with linear_g"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "gemv"
M = int(sys.argv[2]) if len(sys.argv) > 2 else 1
SHAPES = {"gate_up": (4096, 28672), "o": (4096, 4096), "down": (14336, 4096), "qkv": (4096, 6144), "70b": (8192, 57344)}
K, N = SHAPES[sys.argv[3]] if len(sys.argv) > 3 else (4096, 28672)
dtype = torch.float16
_C.lib().awq_b200_debug_set_gemv_impl(int(os.environ.get('PROBE_IMPL', '0')), int(os.environ.get('PROBE_SPLIT', '0')), int(os.environ.get('PROBE_VARIANT', '-1')))
if os.environ.get("PROBE_GEMM_BN"):
    _C.lib().awq_b200_debug_set_gemm_plan(int(os.environ["PROBE_GEMM_BN"]), int(os.environ.get("PROBE_GEMM_ALIGN", "-1")))
G = int(os.environ.get("PROBE_GROUP", "128"))
sets = []
for i in range(2 if K * N > (1 << 28) else 4):
    g = torch.Generator(device="cuda").manual_seed(i)
    qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    qz = torch.randint(-(2**31), 2**31, (K // G, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    sc = (torch.rand(K // G, N, generator=g, device="cuda") * 0.02 + 0.005).to(dtype)
    sets.append((qw, sc, qz) if kind == "dequant" else (_C.repack_g if kind == "linear_g" else _C.repack)(qw, sc, qz))
x = torch.randn(M, K, device="cuda").to(dtype)
torch.cuda.synchronize()
for rep in range(3):
    for w in sets:
        if kind == "dequant":
            _C.dequantize(*w)
        else:
            getattr(_C, kind)(x, w, N)
torch.cuda.synchronize()
print("ok")
