"""Prefill kernel (awq_b200_gemm) under every (token tile, work split) plan next to the reference
sequence (awq_dequantize + cuBLAS matmul), per model shape and M.  Development probe: its output
decides the planner in csrc/awq_gemm_tcgen05.cu (gemm_plan).
usage: python scripts/gemm_matrix.py [8b 70b qwen l2]   env: GEMM_MS=128,256,512,..."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402
from scripts.kernel_matrix import shapes_of, time_graph  # noqa: E402

PLANS = [("auto", 0, -1), ("256s", 256, 0), ("256a", 256, 1), ("128s", 128, 0), ("128a", 128, 1), ("64a", 64, 1), ("64s", 64, 0)]


def main():
    groups = sys.argv[1:] or ["8b", "l2"]
    ms = tuple(int(v) for v in os.environ.get("GEMM_MS", "128,256,512,1024,2048,4096").split(","))
    tps = tuple(int(v) for v in os.environ.get("MATRIX_TP", "1").split(","))
    L = _C.lib()
    rows = []
    print(f"# us per launch (TFLOP/s); ref = awq_dequantize + torch.matmul (cuBLAS) on the same GPU")
    print(f"{'shape':20s} {'K':>6s} {'N':>6s} {'M':>5s} | " + " | ".join(f"{p[0]:>12s}" for p in PLANS) + " |          ref | best")
    seen = set()
    for group in groups:
        for name, K, N in shapes_of(group, tps if group != "l2" else (1,)):
            if (K, N) in seen:
                continue
            seen.add((K, N))
            nsets = max(2, min(8, (300 << 20) // (K * N // 2)))
            raw = []
            for i in range(nsets):
                g = torch.Generator(device="cuda").manual_seed(i)
                qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
                qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
                sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).half()
                raw.append((qw, sc, qz))
            sets = [_C.repack(*r) for r in raw]
            for M in ms:
                x = torch.randn(M, K, device="cuda").half()
                y = torch.empty(M, N, device="cuda", dtype=torch.float16)
                flops = 2.0 * M * K * N
                rec = {"shape": name, "K": K, "N": N, "M": M, "us": {}}
                for label, bn, align in PLANS:
                    if bn and (M + bn - 1) // bn * ((N + 127) // 128) > 2 * 148 and align == 0:
                        continue  # the split plan needs segments < 2 x SMs (workspace bound)
                    L.awq_b200_debug_set_gemm_plan(bn, align)
                    try:
                        fns = [(lambda w=w: _C.gemm(x, w, N, out=y)) for w in sets]
                        rec["us"][label] = round(time_graph(fns, 5), 2)
                    except Exception as e:  # plan not runnable for this problem
                        rec["us"][label] = None
                        torch.cuda.synchronize()
                L.awq_b200_debug_set_gemm_plan(0, -1)
                fns = [(lambda r=r: torch.matmul(x, _C.dequantize(*r), out=y)) for r in raw]
                rec["us"]["ref"] = round(time_graph(fns, 5), 2)
                cells = []
                for label in [p[0] for p in PLANS] + ["ref"]:
                    us = rec["us"].get(label)
                    cells.append(f"{us:7.1f}/{flops / us / 1e6:4.0f}" if us else f"{'-':>12s}")
                best = min((v, k) for k, v in rec["us"].items() if v and k not in ("auto", "ref"))
                rec["best"] = best[1]
                print(f"{name:20s} {K:6d} {N:6d} {M:5d} | " + " | ".join(f"{c:>12s}" for c in cells) + f" | {best[1]}", flush=True)
                rows.append(rec)
            del sets, raw
            torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"rows": rows}, open("gpurun_out/gemm_matrix.json", "w"), indent=0)


if __name__ == "__main__":
    main()
