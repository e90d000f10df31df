"""AWQ-MoE decode timing (development aid / profiles): Mixtral-8x7B-shaped experts (hidden 4096, intermediate 14336, 8 experts,
top-2) on random-init AWQ tensors, `layers` distinct MoE layers in one CUDA graph, random routing per layer.
Reports us per MoE layer and the fraction of the measured HBM peak over the ALGORITHMIC bytes = packed weights of the
distinct experts each layer touches (an expert shared by several tokens counts once)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_awq_b200 import _C  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
E, TOPK, HID, INTER, LAYERS = 8, 2, 4096, 14336, 4


def rand_tiles(K, N, gate_up, g):
    qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    sc = (torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005).half()
    return (_C.repack_gate_up if gate_up else _C.repack)(qw, sc, qz)


def main():
    if os.environ.get("MOE_BLOCKS_PER_SM"):
        _C.lib().awq_b200_debug_set_moe_blocks_per_sm(int(os.environ["MOE_BLOCKS_PER_SM"]))
    if os.environ.get("MOE_CHUNK"):
        _C.lib().awq_b200_debug_set_moe_chunk(int(os.environ["MOE_CHUNK"]))
    g = torch.Generator(device="cuda").manual_seed(0)
    layers = [(torch.stack([rand_tiles(HID, 2 * INTER, True, g) for _ in range(E)]),
               torch.stack([rand_tiles(INTER, HID, False, g) for _ in range(E)])) for _ in range(LAYERS)]
    rec_bytes = lambda t: t.shape[1] * 8576  # noqa: E731
    out = {}
    for M in (1, 2, 4, 8, 16, 32, 64):
        x = (torch.randn(M, HID, generator=g, device="cuda") * 0.3).half()
        routes = []
        for _ in range(LAYERS):
            w, ids = torch.topk(torch.softmax(torch.randn(M, E, generator=g, device="cuda"), -1), TOPK, dim=-1)
            routes.append((w.reshape(-1).float().contiguous(), ids.reshape(-1).to(torch.int32).contiguous()))
        distinct = sum(len(set(ids.tolist())) for _, ids in routes) / LAYERS

        def run():
            for (w13, w2), (w, ids) in zip(layers, routes):
                h = _C.moe_linear(x, w13, ids, 2 * INTER, x_div=TOPK, silu=True)
                o = _C.moe_linear(h, w2, ids, HID, x_div=1, row_scale=w)
                _C.moe_sum(o, TOPK)
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
            for _ in range(10):
                gr.replay()
            e1.record(s)
            torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 10 / LAYERS
        algo = distinct * (rec_bytes(layers[0][0]) + rec_bytes(layers[0][1]))
        streamed = M * TOPK * (rec_bytes(layers[0][0]) + rec_bytes(layers[0][1]))
        out[f"M{M}"] = {"us_per_moe_layer": round(us, 1), "distinct_experts": distinct, "frac_hbm_algorithmic": round(algo / us / 1e3 / PEAK, 3),
                        "frac_hbm_streamed_per_pair": round(streamed / us / 1e3 / PEAK, 3)}
        print(f"M={M:2d}: {us:7.1f} us per MoE layer, {distinct:.1f} distinct experts, {algo / us / 1e3:7.0f} GB/s algorithmic "
              f"({algo / us / 1e3 / PEAK:.2f} of HBM peak); per-pair bytes / time = {streamed / us / 1e3:7.0f} GB/s", flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"config": "Mixtral-8x7B-shaped AWQ experts: hidden 4096, intermediate 14336, 8 experts, top-2", **out},
              open("gpurun_out/moe_probe.json", "w"), indent=1)


if __name__ == "__main__":
    main()
