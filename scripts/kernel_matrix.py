"""Every decode kernel on every per-rank shape of the BASELINE.json models (development aid: what
the shape router in csrc/awq_gemv.cu is built from; bench.py is the contract).

    python scripts/kernel_matrix.py [groups ...]      groups: 8b 70b qwen l2 (default: all)
    env MATRIX_MS=1,2,4,8,16,32   MATRIX_TP=1,2,4,8   MATRIX_DTYPE=bf16

Distinct weight sets >> L2 per shape, CUDA-graph-captured loop, CUDA events.  Prints us per launch
and the fraction of the measured HBM peak per (shape, M, kernel); writes gpurun_out/kernel_matrix.json.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
MODELS = {"8b": (4096, 14336, 32, 8), "70b": (8192, 28672, 64, 8), "qwen": (5120, 27648, 40, 8), "l2": (4096, 11008, 32, 32)}
# (label, impl, variant)
KERNELS_ALL = [("auto", 0, -1), ("direct", 1, -1), ("persist", 3, -1), ("cluster", 4, -1), ("tc_g3d2b2", 5, 0), ("tc_g3d2b1", 5, 1), ("tc_g4d2b1", 5, 3),
           ("legacy", 2, -1)]
# MATRIX_KERNELS="label:impl:variant[:split],..." replaces the kernel list (timing experiments; split = team / cluster size)
KERNELS = ([tuple(int(f) if i else f for i, f in enumerate(k.split(":"))) for k in os.environ["MATRIX_KERNELS"].split(",")]
           if os.environ.get("MATRIX_KERNELS") else KERNELS_ALL)


def algo_bytes(M, K, N):
    return K * N // 2 + (K // 128) * N * 2 + (K // 128) * N // 2 + M * K * 2 + M * N * 2


def shapes_of(group, tps):
    hidden, inter, heads, kv = MODELS[group]
    q, kvd = heads * 128, kv * 128
    out = []
    for tp in tps:
        if heads % tp or (group == "qwen" and tp == 8):
            continue
        kv_r = max(kvd // tp, 128)
        out += [(f"{group}_qkv_tp{tp}", hidden, q // tp + 2 * kv_r), (f"{group}_o_tp{tp}", q // tp, hidden),
                (f"{group}_gate_up_tp{tp}", hidden, 2 * inter // tp), (f"{group}_down_tp{tp}", inter // tp, hidden)]
    return out


DT = torch.bfloat16 if os.environ.get("MATRIX_DTYPE") == "bf16" else torch.float16
DTC = 1 if DT == torch.bfloat16 else 0


def make_sets(K, N, nsets):
    sets = []
    for i in range(nsets):
        g = torch.Generator(device="cuda").manual_seed(i)
        qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).to(DT)
        sets.append(_C.repack(qw, sc, qz))
    return sets


def time_graph(fns, iters):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for f in fns:
            f()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for f in fns:
                f()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(iters):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * len(fns))


def main():
    groups = sys.argv[1:] or ["8b", "70b", "qwen", "l2"]
    ms = tuple(int(v) for v in os.environ.get("MATRIX_MS", "1,2,4,8,16,32").split(","))
    tps = tuple(int(v) for v in os.environ.get("MATRIX_TP", "1,2,4,8").split(","))
    L = _C.lib()
    rows = []
    print(f"# us per launch / fraction of measured HBM peak ({PEAK} GB/s); '-' = kernel cannot take the shape")
    print(f"{'shape':22s} {'K':>6s} {'N':>6s} {'M':>3s} | " + " | ".join(f"{k[0]:>12s}" for k in KERNELS) + " | best")
    seen = set()
    for group in groups:
        for name, K, N in shapes_of(group, tps if group != "l2" else (1,)):
            if (K, N) in seen:
                continue
            seen.add((K, N))
            nsets = max(2, min(24, (400 << 20) // (K * N // 2)))
            sets = make_sets(K, N, nsets)
            for M in ms:
                x = torch.randn(M, K, device="cuda").to(DT)
                y = torch.empty(M, N, device="cuda", dtype=DT)
                rec = {"shape": name, "K": K, "N": N, "M": M, "us": {}}
                for label, impl, variant, *rest in KERNELS:
                    if impl and not L.awq_b200_debug_gemv_impl_fits(impl, M, K, N, DTC):
                        continue
                    if impl == 0:
                        rec["routed"] = int(L.awq_b200_gemv_route(M, K, N, DTC))
                    L.awq_b200_debug_set_gemv_impl(impl, rest[0] if rest else 0, variant)
                    fns = [(lambda w=w: _C.gemv(x, w, N, out=y)) for w in sets]
                    rec["us"][label] = round(time_graph(fns, 10), 3)
                L.awq_b200_debug_set_gemv_impl(0, 0, -1)
                b = algo_bytes(M, K, N)
                cells = []
                for label, *_ in KERNELS:
                    us = rec["us"].get(label)
                    cells.append(f"{us:6.2f}/{b / us / 1e3 / PEAK:4.2f} " if us else f"{'-':>12s}")
                best = min((v, k) for k, v in rec["us"].items() if k != "auto")
                rec["best"] = best[1]
                print(f"{name:22s} {K:6d} {N:6d} {M:3d} | " + " | ".join(f"{c:>12s}" for c in cells) + f" | {best[1]}", flush=True)
                rows.append(rec)
            del sets
            torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"peak_gbs": PEAK, "rows": rows}, open("gpurun_out/kernel_matrix.json", "w"), indent=0)


if __name__ == "__main__":
    main()
