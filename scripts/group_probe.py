"""Decode timing by AWQ group size (development aid / profiles): awq_b200_linear_g at G = 128 / 64 / 32 on the
gate_up / down shapes of Llama-3-8B and Llama-3-70B, `copies` distinct weight sets per shape in one CUDA graph
(working set >> L2).  Reports us per launch and the fraction of the measured HBM peak over the packed bytes
(weights + scales + zeros of that group size)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
SHAPES = {"8b_gate_up": (4096, 28672), "8b_down": (14336, 4096), "70b_gate_up": (8192, 57344), "70b_o": (8192, 8192)}


def rand_tiles(K, N, G, g, dtype):
    qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    qz = torch.randint(-(2**31), 2**31, (K // G, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    sc = (torch.rand(K // G, N, generator=g, device="cuda") * 0.02 + 0.005).to(dtype)
    return _C.repack_g(qw, sc, qz)


def time_graph(run, iters=10):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        run()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            run()
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(iters):
            gr.replay()
        e1.record(s)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    dtype = torch.bfloat16 if "bf16" in sys.argv else torch.float16
    if os.environ.get("SUB_DECODE"):  # 0 = tcgen05 kernel only, 1 = default, 2 = per-weight kernel wherever it fits
        _C.lib().awq_b200_debug_set_sub_decode(int(os.environ["SUB_DECODE"]))
    g = torch.Generator(device="cuda").manual_seed(0)
    out = {}
    for name, (K, N) in SHAPES.items():
        for G in (128, 64, 32):
            rec = _C.lib().awq_b200_tile_bytes(G)
            nbytes = _C.lib().awq_b200_num_tiles(K, N) * rec
            copies = max(2, min(24, int(600e6 // nbytes)))
            sets = [rand_tiles(K, N, G, g, dtype) for _ in range(copies)]
            for M in (1, 2, 4, 8, 16):
                x = (torch.randn(M, K, generator=g, device="cuda") * 0.3).to(dtype)
                us = time_graph(lambda: [_C.linear_g(x, t, N) for t in sets]) / copies
                out[f"{name}_G{G}_M{M}"] = {"us": round(us, 2), "frac_hbm": round(nbytes / us / 1e3 / PEAK, 3)}
                print(f"{name:12s} G={G:3d} M={M:2d}: {us:8.2f} us  {nbytes / us / 1e3:7.0f} GB/s  {nbytes / us / 1e3 / PEAK:.3f}", flush=True)
            del sets
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open(os.environ.get("GROUP_PROBE_OUT", "gpurun_out/group_probe.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
