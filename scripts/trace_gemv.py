"""Development aid: per-CTA phase timeline of the decode kernel (needs the -DAWQ_TRACE build:
scripts/build_trace.sh -> sglang_awq_b200/lib/libawq_b200_trace.so).
usage: AWQ_B200_LIB=.../libawq_b200_trace.so python scripts/trace_gemv.py M shape [variant] [split]"""
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
variant = int(sys.argv[3]) if len(sys.argv) > 3 else 0
split = int(sys.argv[4]) if len(sys.argv) > 4 else 0
L = _C.lib()
L.awq_b200_debug_set_gemv_impl(0, split, variant)
sets = []
for i in range(int(os.environ.get("NSETS", "3"))):
    g = torch.Generator(device="cuda").manual_seed(i)
    qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).to(torch.float16)
    sets.append(_C.repack(qw, sc, qz))
x = torch.randn(M, K, device="cuda").half()
trace = torch.zeros(8192 * 32, dtype=torch.int64, device="cuda")
for wi, w in enumerate(sets):
    _C.gemv(x, w, N)
torch.cuda.synchronize()
L.awq_b200_debug_set_trace(ctypes.c_void_p(trace.data_ptr()))
_C.gemv(x, sets[0], N)
torch.cuda.synchronize()
L.awq_b200_debug_set_trace(None)
t = trace.cpu().numpy().reshape(-1, 2, 16)
used = t[:, 0, 0] > 0
t = t[used]
n = t.shape[0]
t0 = t[:, :, 0].min()
print(f"CTAs {n}; kernel span {(t[:, :, 8].max() - t0) / 1e3:.2f} us")
names = ["entry", "pdl_wait", "stage_done", "sync1", "grp0_done", "loop_done", "clu_wait1", "clu_bar2", "end"]
for wsel in (0, 1):
    rel = (t[:, wsel, :9] - t[:, wsel, 0:1]) / 1e3
    print(f"warp {'0' if wsel == 0 else '7'}: median us since CTA entry: " + ", ".join(f"{nm}={np.median(rel[:, i]):.2f}" for i, nm in enumerate(names)))
    print(f"         p90: " + ", ".join(f"{nm}={np.percentile(rel[:, i], 90):.2f}" for i, nm in enumerate(names)))
start = (t[:, 0, 0] - t0) / 1e3
end = (t[:, 0, 8] - t0) / 1e3
print("CTA start times us: p0/p25/p50/p75/p100 =", np.percentile(start, [0, 25, 50, 75, 100]).round(2))
print("CTA end   times us: p0/p25/p50/p75/p100 =", np.percentile(end, [0, 25, 50, 75, 100]).round(2))
smid = t[:, 0, 15]
print("CTAs per SM: min/mean/max", np.bincount(smid.astype(int)).min(), n / len(np.unique(smid)), np.bincount(smid.astype(int)).max(), "SMs used", len(np.unique(smid)))
# concurrency: how many CTAs alive at sample times
ts = np.linspace(0, end.max(), 30)
alive = [(int(((start <= q) & (end > q)).sum())) for q in ts]
print("alive CTAs over time:", alive)
