"""Development aid: phase timeline (%globaltimer, ns) of CTA 0 of the tcgen05 decode kernel over a CUDA-graph chain
of launches on rotating weight sets.   usage: python scripts/trace_tc.py M K N [impl-variant]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scripts.kernel_matrix import make_sets  # noqa: E402
from sglang_awq_b200 import _C  # noqa: E402

M, K, N = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
L = _C.lib()
L.awq_b200_debug_set_gemv_impl(5, 0, int(sys.argv[4]) if len(sys.argv) > 4 else 0)
sets = make_sets(K, N, max(2, min(16, (600 << 20) // (K * N // 2))))
x = torch.randn(M, K, device="cuda").half()
y = torch.empty(M, N, device="cuda", dtype=torch.float16)
trace = torch.zeros(16 + 64 * 16 + 4 * 160 * 8, dtype=torch.int64, device="cuda")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for t in sets:
        _C.gemv(x, t, N, out=y)
    torch.cuda.synchronize()
    L.awq_b200_debug_set_tc_trace(trace.data_ptr())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        for i in range(16):
            _C.gemv(x, sets[i % len(sets)], N, out=y)
    L.awq_b200_debug_set_tc_trace(None)
    trace.zero_()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    trace.zero_()
    e0.record(s)
    g.replay()
    e1.record(s)
    torch.cuda.synchronize()
print(f"M={M} K={K} N={N}: {e0.elapsed_time(e1) * 1e3 / 16:.2f} us per launch (with tracing on)")
t = trace.cpu().numpy()
names = ["entry", "setup", "pdl_wait", "first_A", "lastA_g0", "lastA_g1", "lastA_g2", "lastA_g3", "mma_first", "mma_last",
         "drain_first", "drain_last", "drain_done", "all_done", "dealloc"]
rows = t[16:16 + 16 * 16].reshape(16, 16)
rows = rows[rows[:, 0] > 0]
rows = rows[rows[:, 0].argsort()]
prev_end = None
for r in rows[2:10]:
    base = r[0]
    line = " ".join(f"{n}={int(r[i] - base):6d}" for i, n in enumerate(names) if r[i] > 0)
    gap = "" if prev_end is None else f" [entry - previous dealloc = {int(base - prev_end)} ns]"
    print(line + gap)
    prev_end = r[14]

import numpy as np  # noqa: E402
allc = t[16 + 64 * 16:].reshape(4, 160, 8)
for li in range(4):
    a = allc[li]
    a = a[a[:, 0] > 0]
    if not len(a):
        continue
    b0 = a[:, 0].min()
    ent, pw, dd, de = (a[:, i] - b0 for i in range(4))
    cb, ct, cs = (np.where(a[:, i] > 0, a[:, i] - b0, -1) for i in (4, 5, 6))
    q = lambda v: " ".join(f"{int(x):6d}" for x in np.percentile(v, [0, 10, 50, 90, 100]))  # noqa: E731
    print(f"launch {li + 4}: {len(a)} CTAs, ns relative to the first entry; percentiles 0/10/50/90/100")
    print(f"   entry      {q(ent)}\n   pdl_wait   {q(pw)}\n   drain_done {q(dd)}\n   dealloc    {q(de)}\n   work (dealloc - pdl_wait) {q(de - pw)}")
    print(f"   closure begin {q(cb[cb >= 0])}\n   ticket - begin {q((ct - cb)[ct >= 0])}\n   last arrivers: {int((cs >= 0).sum())}, sum+store - ticket {q((cs - ct)[cs >= 0]) if (cs >= 0).any() else ''}")
    print(f"   drain_done - closure begin {q((dd - cb)[cb >= 0])}")
    slow = np.argsort(de)[-3:]
    print("   slowest CTAs:", [(int(np.flatnonzero(allc[li][:, 3] - b0 == de[i])[0]), int(ent[i]), int(pw[i]), int(dd[i]), int(de[i])) for i in slow])
