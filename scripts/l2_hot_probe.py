"""How much of a small decode launch is HBM time?  Each shape is timed in a CUDA-graph chain of 64 launches
(a) over rotating weight sets >> L2 (HBM, what bench.py measures) and (b) over ONE weight set (L2 hits: what a
perfect cross-kernel L2 prefetch of the next linear's weights could reach).  Development aid."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scripts.kernel_matrix import algo_bytes, make_sets, time_graph  # noqa: E402
from sglang_awq_b200 import _C  # noqa: E402

SHAPES = [("70b_qkv_tp8", 8192, 1280), ("70b_o_tp8", 1024, 8192), ("70b_gate_up_tp8", 8192, 7168), ("70b_down_tp8", 3584, 8192),
          ("70b_qkv_tp4", 8192, 2560), ("70b_o_tp4", 2048, 8192), ("70b_gate_up_tp4", 8192, 14336), ("70b_down_tp4", 7168, 8192),
          ("70b_o_tp1", 8192, 8192), ("70b_qkv_tp1", 8192, 10240), ("8b_o_tp1", 4096, 4096), ("8b_qkv_tp1", 4096, 6144)]


def main():
    out = {}
    for name, K, N in SHAPES:
        nsets = max(2, min(64, (700 << 20) // (K * N // 2)))
        sets = make_sets(K, N, nsets)
        for M in (1, 4, 16):
            x = torch.randn(M, K, device="cuda").half()
            y = torch.empty(M, N, device="cuda", dtype=torch.float16)
            hbm = time_graph([lambda t=t: _C.linear(x, t, N, out=y) for t in sets], 5)
            hot = time_graph([lambda: _C.linear(x, sets[0], N, out=y) for _ in range(32)], 5)
            b = algo_bytes(M, K, N)
            out[f"{name}_M{M}"] = {"hbm_us": round(hbm, 2), "l2_hot_us": round(hot, 2), "ideal_hbm_us": round(b / 6550e3, 2)}
            print(f"{name:18s} M={M:2d}  rotating {hbm:6.2f} us   one set (L2) {hot:6.2f} us   bytes/peak {b / 6550e3:5.2f} us", flush=True)
        del sets
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/l2_hot_probe.json", "w"), indent=1)


if __name__ == "__main__":
    main()
