"""Development aid: per-tile pipeline timeline of the tcgen05 decode kernel (needs the -DAWQ_TRACE
build: scripts/build_trace.sh -> sglang_awq_b200/lib/libawq_b200_trace.so).
usage: AWQ_B200_LIB=.../libawq_b200_trace.so python scripts/trace_tc.py M shape"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 1
SHAPES = {"gate_up": (4096, 28672), "o": (4096, 4096), "down": (14336, 4096), "qkv": (4096, 6144), "70b": (8192, 57344)}
K, N = SHAPES[sys.argv[2] if len(sys.argv) > 2 else "gate_up"]
L = _C.lib()
L.awq_b200_debug_set_gemv_impl(5, int(os.environ.get("TRACE_SPLIT", "0")), -1)
sets = []
for i in range(3):
    g = torch.Generator(device="cuda").manual_seed(i)
    qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).to(torch.float16)
    sets.append(_C.repack(qw, sc, qz))
x = torch.randn(M, K, device="cuda").half()
trace = torch.zeros(4 * 16 * 256, dtype=torch.int64, device="cuda")
for w in sets:
    _C.gemv(x, w, N)
torch.cuda.synchronize()
L.awq_b200_debug_set_tc_trace(ctypes.c_void_p(trace.data_ptr()))
_C.gemv(x, sets[0], N)
torch.cuda.synchronize()
L.awq_b200_debug_set_tc_trace(None)
t = trace.cpu().numpy().reshape(4, 16, 256)
names = ["prod_issue", "cvt_full", "cvt_aempty", "cvt_afull", "mma_go", "mma_commit", "drn_dfull", "drn_done"]
for cta in range(2):
    tt = t[cta]
    n = int((tt[6] > 0).sum())
    base = tt[1][0]
    print(f"CTA {cta}: {n} tiles; clocks relative to first cvt_full")
    for it in list(range(0, min(n, 24))) + list(range(max(24, n - 4), n)):
        print(f"  it={it:3d} " + " ".join(f"{nm}={(tt[r][it] - base) if tt[r][it] > 0 else -1:7d}" for r, nm in enumerate(names)))
    if n > 12:
        d = np.diff(tt[7][8:n].astype(np.float64))
        print(f"  steady-state clocks per tile (drain to drain): mean {d.mean():.0f} min {d.min():.0f} max {d.max():.0f}")
        names += ["-", "cvt_st_issued", "cvt_wait_st"]
        for a, b in ((0, 1), (1, 2), (2, 9), (9, 10), (10, 3), (3, 4), (4, 5), (5, 6), (6, 7)):
            dd = (tt[b][8:n] - tt[a][8:n]).astype(np.float64)
            print(f"  {names[a]} -> {names[b]}: mean {dd.mean():.0f}")
        for r in (1, 3, 4, 7):
            per = np.diff(tt[r][8:n].astype(np.float64))
            print(f"  {names[r]} period: mean {per.mean():.0f}  (same group, 3 tiles apart: {np.mean(tt[r][11:n] - tt[r][8:n-3]):.0f})")
