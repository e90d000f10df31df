"""Latency of one static batch through the AWQ linear stack, without a server.

The reference's protocol (python/sglang/bench_one_batch.py:497-624 `latency_test_run_once`):
one prefill over batch_size x input_len tokens, then output_len - 1 decode steps of batch_size
tokens; prefill latency/throughput, the MEDIAN decode latency/throughput and the overall
throughput are reported, one result dict per (batch_size, input_len, output_len), also appended as
JSON lines to --result-filename.  `--load-format dummy` (the only format here: there is no
checkpoint reader on this path) follows loader.py:1211-1218.

What runs is the slice of the model this repository implements: every AWQ linear of every layer
(+ the row-parallel all-reduces), chained by sglang_awq_b200/srt/models/linear_stack.py.
Attention, KV cache, norms with weights, sampling and the LM head are outside the path, so the
numbers are "linears only" and say so.

    python -m sglang_awq_b200.bench_one_batch --model qwen2.5-32b --batch-size 1 --input-len 1024 --output-len 16
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 -m sglang_awq_b200.bench_one_batch \
        --model qwen2.5-32b --tp-size 4
"""
from __future__ import annotations

import argparse
import json
import os
import time
from typing import List

import numpy as np
import torch


def _rank_print(rank):
    return print if rank == 0 else (lambda *a, **k: None)


def latency_test_run_once(run_name, stack, rank_print, batch_size, input_len, output_len, device, hidden, dtype,
                          cuda_graph=True, seed=0):
    """reference bench_one_batch.py:497-624 (same keys in the returned dict)."""
    g = torch.Generator(device=device).manual_seed(seed)
    results = {"run_name": run_name, "batch_size": batch_size, "input_len": input_len, "output_len": output_len}
    x_prefill = torch.randn(batch_size * input_len, hidden, generator=g, device=device).to(dtype)
    x_decode = torch.randn(batch_size, hidden, generator=g, device=device).to(dtype)

    tot_latency = 0.0
    torch.cuda.synchronize(device)
    tic = time.perf_counter()
    with torch.no_grad():
        h = stack(x_prefill)
    torch.cuda.synchronize(device)
    prefill_latency = time.perf_counter() - tic
    tot_latency += prefill_latency
    throughput = input_len * batch_size / prefill_latency
    rank_print(f"Prefill. latency: {prefill_latency:6.5f} s, throughput: {throughput:9.2f} token/s")
    results["prefill_latency"] = prefill_latency
    results["prefill_throughput"] = throughput

    # decode runs under a CUDA graph, as the reference's model runner does by default
    graph = None
    stream = torch.cuda.current_stream(device)
    with torch.no_grad():
        h = stack(x_decode)   # every kernel has run once (module load, workspace) before capture
        torch.cuda.synchronize(device)
        if cuda_graph:
            side = torch.cuda.Stream(device=device)
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                stack(x_decode)
                torch.cuda.synchronize(device)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=side):
                    h = stack(x_decode)
            stream.wait_stream(side)
    decode_latencies: List[float] = []
    for i in range(output_len - 1):
        torch.cuda.synchronize(device)
        tic = time.perf_counter()
        if graph is not None:
            graph.replay()
        else:
            with torch.no_grad():
                h = stack(x_decode)
        torch.cuda.synchronize(device)
        latency = time.perf_counter() - tic
        tot_latency += latency
        decode_latencies.append(latency)
        if i < 5:
            rank_print(f"Decode {i}. Batch size: {batch_size}, latency: {latency:6.5f} s, "
                       f"throughput: {batch_size / latency:9.2f} token/s")
    if output_len > 1:
        med = float(np.median(decode_latencies))
        rank_print(f"Decode.  median latency: {med:6.5f} s, median throughput: {batch_size / med:9.2f} token/s")
        results["median_decode_latency"] = med
        results["median_decode_throughput"] = batch_size / med
    throughput = (input_len + output_len) * batch_size / tot_latency
    rank_print(f"Total. latency: {tot_latency:6.3f} s, throughput: {throughput:9.2f} token/s")
    results["total_latency"] = tot_latency
    results["overall_throughput"] = throughput
    results["last_hidden_abs_mean"] = float(h.float().abs().mean())
    del graph
    return results


def main(argv=None):
    from .srt import distributed as D
    from .srt.layers.quantization import AWQConfig, AWQMarlinConfig
    from .srt.model_loader import load_dummy_model
    from .srt.models import MODEL_SHAPES, AWQLinearStack

    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--model", choices=sorted(MODEL_SHAPES), default="qwen2.5-32b")
    ap.add_argument("--load-format", choices=["dummy"], default="dummy")
    ap.add_argument("--quantization", choices=["awq", "awq_marlin"], default="awq")
    ap.add_argument("--dtype", choices=["half", "bfloat16"], default="half")
    ap.add_argument("--tp-size", "--tp", type=int, default=None, help="default: WORLD_SIZE")
    ap.add_argument("--run-name", default="default")
    ap.add_argument("--batch-size", type=int, nargs="+", default=[1])
    ap.add_argument("--input-len", type=int, nargs="+", default=[1024])
    ap.add_argument("--output-len", type=int, nargs="+", default=[16])
    ap.add_argument("--num-layers", type=int, default=None, help="default: the model's layer count")
    ap.add_argument("--disable-cuda-graph", action="store_true")
    ap.add_argument("--result-filename", default="result.jsonl")
    ap.add_argument("--fuse-silu-mul", action="store_true", help="SiluAndMul in the gate_up_proj epilogue (one launch)")
    args = ap.parse_args(argv)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    tp = args.tp_size or world
    assert tp == world, "launch one process per TP rank (torchrun --nproc-per-node <tp>)"
    if not torch.cuda.is_available():
        raise SystemExit("bench_one_batch: no CUDA device; the AWQ kernels have no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        D.init_distributed_environment(world_size=world, rank=rank, local_rank=local_rank, backend="nccl")
        D.initialize_model_parallel(world)
    rank_print = _rank_print(rank)
    model = MODEL_SHAPES[args.model]
    dtype = torch.float16 if args.dtype == "half" else torch.bfloat16
    if args.quantization == "awq":
        cfg = AWQConfig.from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True})
    else:
        cfg = AWQMarlinConfig.from_config({"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True})
    with torch.device(device):
        stack = AWQLinearStack(model, cfg, num_layers=args.num_layers, params_dtype=dtype,
                               tp_rank=rank if world > 1 else 0, tp_size=world, fuse_silu_mul=args.fuse_silu_mul)
    load_dummy_model(stack)
    rank_print(f"model={args.model} quantization={args.quantization} dtype={args.dtype} tp={world} "
               f"layers={len(stack.layers)} (AWQ linears + row-parallel all-reduces only)")
    # warm up (reference: one short run before the measured ones, bench_one_batch.py:672-690)
    latency_test_run_once("warmup", stack, lambda *a, **k: None, args.batch_size[0], min(32, args.input_len[0]), 4,
                          device, model["hidden"], dtype, cuda_graph=not args.disable_cuda_graph)
    out = []
    for bs in args.batch_size:
        for il in args.input_len:
            for ol in args.output_len:
                r = latency_test_run_once(args.run_name, stack, rank_print, bs, il, ol, device, model["hidden"], dtype,
                                          cuda_graph=not args.disable_cuda_graph)
                r.update(model=args.model, tp_size=world, quantization=args.quantization, scope="awq linears only",
                         fuse_silu_mul=bool(args.fuse_silu_mul))
                out.append(r)
    if rank == 0 and args.result_filename:
        with open(args.result_filename, "a") as f:
            for r in out:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        torch.cuda.synchronize()
        D.get_tp_group().barrier()
        D.destroy_model_parallel()
        D.destroy_distributed_environment()
    return out


if __name__ == "__main__":
    main()
