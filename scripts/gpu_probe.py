"""Quick kernel-level timing on the GPU box (development aid; bench.py is the contract).

Rotates over enough distinct weight sets to exceed L2 (126 MB), times with CUDA events around a
CUDA-graph-captured loop, prints algorithmic GB/s vs the measured HBM peak.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

PEAK = 6545.6
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def algo_bytes(M, K, N):
    return K * N // 2 + (K // 128) * N * 2 + (K // 128) * N // 2 + M * K * 2 + M * N * 2


def make_sets(K, N, dtype, nsets):
    sets = []
    for i in range(nsets):
        g = torch.Generator(device="cuda").manual_seed(i)
        qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).to(dtype)
        sets.append(_C.repack(qw, sc, qz))
        del qw, qz, sc
    return sets


def time_graph(fn_list, iters=20):
    """fn_list: list of callables (one per weight set). Returns avg µs per call."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for f in fn_list:
            f()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for f in fn_list:
                f()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(iters):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * len(fn_list))


MS = tuple(int(v) for v in os.environ.get('PROBE_MS', '1,4,16').split(','))


IMPL = int(os.environ.get('PROBE_IMPL', '0'))
SPLIT = int(os.environ.get('PROBE_SPLIT', '0'))


def main():
    out = {"peak_gbs": PEAK, "gemv": [], "dequant": []}
    dtype = torch.float16
    shapes = [("qkv", 4096, 6144), ("o", 4096, 4096), ("gate_up", 4096, 28672), ("down", 14336, 4096),
              ("l2_7b_mlp", 4096, 11008), ("70b_gate_up", 8192, 57344)]
    if os.environ.get("PROBE_SHAPES") == "70b_tp":  # per-rank shapes of Llama-3-70B at TP = 8 / 4 / 2
        shapes = []
        for tp in (8, 4, 2):
            shapes += [(f"qkv_tp{tp}", 8192, 10240 // tp), (f"o_tp{tp}", 8192 // tp, 8192),
                       (f"gate_up_tp{tp}", 8192, 57344 // tp), (f"down_tp{tp}", 28672 // tp, 8192)]
    kind = sys.argv[1] if len(sys.argv) > 1 else "gemv"
    if 'PROBE_DIRECT_MAX_M' in os.environ:
        _C.lib().awq_b200_set_gemv_direct_max_m(int(os.environ['PROBE_DIRECT_MAX_M']))
    for name, K, N in shapes:
        wbytes = K * N // 2
        nsets = max(2, min(24, (512 << 20) // wbytes))
        sets = make_sets(K, N, dtype, nsets)
        Ms = MS if kind == "gemv" else (1, 4, 16, 64, 256, 512, 4096)
        for M in Ms:
            x = torch.randn(M, K, device="cuda").to(dtype)
            y = torch.empty(M, N, device="cuda", dtype=dtype)
            for pdl, cps in (tuple((1, int(v)) for v in os.environ.get('PROBE_VARIANTS', '-1').split(',')) if kind == "gemv" else ((1, 0),)):
                _C.lib().awq_b200_set_pdl(pdl)
                _C.lib().awq_b200_debug_set_gemv_impl(IMPL, SPLIT, cps)
                fns = [(lambda w=w: getattr(_C, kind)(x, w, N, out=y)) for w in sets]
                us = time_graph(fns, iters=20 if M <= 64 else 3)
                gbs = algo_bytes(M, K, N) / us / 1e3
                rec = {"shape": name, "K": K, "N": N, "M": M, "pdl": pdl, "cps": cps, "us": round(us, 3), "gbs": round(gbs, 1),
                       "frac": round(gbs / PEAK, 3), "tflops": round(2.0 * M * K * N / us / 1e6, 1), "nsets": nsets}
                out["gemv"].append(rec)
                print(rec, flush=True)
        _C.lib().awq_b200_set_pdl(1)
        _C.lib().awq_b200_set_gemv_ctas_per_sm(0)
        del sets
        torch.cuda.empty_cache()
    # standalone dequant
    for name, K, N in [("l2_7b_mlp", 4096, 11008), ("gate_up", 4096, 28672)]:
        if os.environ.get("PROBE_SHAPES"):
            break
        g = torch.Generator(device="cuda").manual_seed(0)
        srcs = []
        for i in range(4):
            qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
            qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
            sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).to(dtype)
            srcs.append((qw, sc, qz, torch.empty(K, N, device="cuda", dtype=dtype)))
        L = _C.lib()

        def mk(t):
            qw, sc, qz, o = t
            return lambda: L.awq_b200_dequantize(qw.data_ptr(), sc.data_ptr(), qz.data_ptr(), o.data_ptr(), K, N, K // 128, 0,
                                                 torch.cuda.current_stream().cuda_stream)
        if kind != "gemv":
            break
        us = time_graph([mk(t) for t in srcs])
        b = K * N // 2 + (K // 128) * N * 2.5 + 2 * K * N
        rec = {"shape": name, "K": K, "N": N, "us": round(us, 2), "gbs": round(b / us / 1e3, 1), "frac": round(b / us / 1e3 / PEAK, 3)}
        out["dequant"].append(rec)
        print(rec, flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open(f"gpurun_out/probe_{kind}.json", "w"), indent=1)


if __name__ == "__main__":
    main()
