"""Development aid: decode-kernel implementations against each other and the fp32 matmul over
the bit-exact dequantised weights (impl 0 = cluster, 1 = direct, 2 = legacy stream-K)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

L = _C.lib()
torch.manual_seed(0)
bad = 0
for dtype in (torch.float16, torch.bfloat16):
    for K, N in ((4096, 4096), (1024, 384), (14336, 4096), (512, 72), (3584, 8192), (128, 128)):
        g = torch.Generator(device="cuda").manual_seed(K + N)
        qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        sc = ((torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005) * (torch.rand(K // 128, N, generator=g, device="cuda") > 0.2).float().mul(2).sub(1)).to(dtype)
        w = _C.repack(qw, sc, qz)
        wd = _C.dequantize(qw, sc, qz).float()
        bias = torch.randn(N, device="cuda").to(dtype)
        for M in (1, 2, 3, 4, 7, 8, 9, 16):
            x = torch.randn(M, K, device="cuda").to(dtype)
            ref = x.float() @ wd + bias.float()
            for impl in (0, 1, 2, 3, 4):
                for variant in (-1, 0, 1, 3):
                    if impl in (0, 2) and variant >= 0:
                        continue
                    L.awq_b200_debug_set_gemv_impl(impl, 0, variant)
                    y = _C.gemv(x, w, N, bias=bias).float()
                    err = ((y - ref).norm() / ref.norm()).item()
                    lim = 1e-3 if dtype == torch.float16 else 6e-3
                    if not (err <= lim):
                        bad += 1
                        print(f"FAIL dtype={dtype} K={K} N={N} M={M} impl={impl} variant={variant} err={err:.3e}")
L.awq_b200_debug_set_gemv_impl(0, 0, -1)
print("cmp_impl: failures =", bad)
