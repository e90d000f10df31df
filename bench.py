#!/usr/bin/env python
"""bench.py -- headline benchmark of the AWQ W4A16 quantized-linear hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Workload (default, every N): BASELINE.json configs[4], the model its metric "tok/s at TP1/2/4/8" is
quoted on -- Llama-3-70B-AWQ, all 80 x [qkv_proj, o_proj, gate_up_proj, down_proj] linears.  One STEP
= one decode pass over every AWQ linear of the model at M = 1, 4 and 16 (960 fused W4A16 launches
over 35.6 GB of distinct packed weights >> the 126 MB L2, so HBM is measured, not L2).  Synthetic
activations, random-init AWQ tensors of the real shapes.  `--workload llama3-8b` is configs[1]
(its per-shape table is also printed in `extras` of the default run).

  value   GB/s = ALGORITHMIC bytes of the step (packed weights + scales + zeros + x + y per
          linear) / device time; inputs resident in HBM, CUDA-graph replay (decode
          runs under CUDA graphs in the reference too), CUDA events, max over ranks.
  e2e     same metric through the public layer API with HOST buffers: each step copies its
          activations from pinned host memory, replays, and copies the results back.
  decode_tokens_per_s   per M: M / (time of one pass over all linears + row-parallel all-reduces).
  N > 1   the same model tensor-parallel over N ranks (column-parallel N-splits, row-parallel
          K-splits closed by the peer-memory all-reduce): fixed total work -> "strong" scaling.
          Before timing, the all-reduce and the TP split are CHECKED (`parity` in the line).
  --impl reference   the reference's own CPU path (torch dequant + matmul, oracle/torch_port.py)
          on the host cores, bounded sample, same metric.

Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {
    # name: hidden, intermediate, q heads, kv heads, head dim, layers, qkv bias
    "llama3-8b": dict(hidden=4096, inter=14336, heads=32, kv_heads=8, head_dim=128, layers=32, qkv_bias=0),
    "llama3-70b": dict(hidden=8192, inter=28672, heads=64, kv_heads=8, head_dim=128, layers=80, qkv_bias=0),
    "qwen2.5-32b": dict(hidden=5120, inter=27648, heads=40, kv_heads=8, head_dim=128, layers=64, qkv_bias=1),
    "llama2-7b-mlp": dict(hidden=4096, inter=11008, heads=32, kv_heads=32, head_dim=128, layers=32, qkv_bias=0),
}
DECODE_MS = (1, 4, 16)
GROUP = 128
ROW_PARALLEL = ("o_proj", "down_proj")
GEMV_KERNEL_NAMES = {1: "awq::gemv_direct_kernel", 2: "awq::gemv_kernel", 3: "awq::gemv_persist_kernel",
                     4: "awq::gemv_cluster_kernel", 5: "awq::gemv_tc_kernel", 6: "awq::gemm_kernel"}

# dram__bytes_read.sum + dram__bytes_write.sum per launch from `ncu --set full` captures, keyed by
# (kernel, K, N, M).  roofline.traffic is emitted ONLY when every launch of the dominant kernel that
# was timed has an entry here for the kernel the router actually picked; otherwise it is null.
PROFILED_DRAM_TRAFFIC = {}
try:
    with open(os.path.join(ROOT, "profiles", "dram_traffic.json")) as _f:
        for _e in json.load(_f):
            PROFILED_DRAM_TRAFFIC[(_e["kernel"], _e["K"], _e["N"], _e["M"])] = _e
except Exception:
    pass


def linear_shapes(m):
    """(name, K, N) of the four AWQ linears of one decoder layer (unsharded)."""
    q = m["heads"] * m["head_dim"]
    kv = m["kv_heads"] * m["head_dim"]
    return [("qkv_proj", m["hidden"], q + 2 * kv), ("o_proj", q, m["hidden"]),
            ("gate_up_proj", m["hidden"], 2 * m["inter"]), ("down_proj", m["inter"], m["hidden"])]


def algo_bytes(M, K, N):
    """K*N/2 [qweight] + (K/G)*N*2 [scales] + (K/G)*N/2 [qzeros] + M*K*2 [x] + M*N*2 [y]."""
    return K * N // 2 + (K // GROUP) * N * 2 + (K // GROUP) * N // 2 + M * K * 2 + M * N * 2


def dequant_bytes(K, N):
    """K*N/2 + (K/G)*N*2.5 + 2*K*N."""
    return K * N // 2 + (K // GROUP) * N * 5 // 2 + 2 * K * N


def step_bytes(model, ms=None):
    ms = DECODE_MS if ms is None else ms
    return sum(algo_bytes(M, K, N) for M in ms for _, K, N in linear_shapes(model)) * model["layers"]


def load_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, 1590.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU every ~5 ms through NVML while the timed
    region runs (nvidia-smi is too slow for a 100 ms region)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index = index
        self.sm, self.reason_bits, self.max_mhz = [], 0, None
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            while not self._stop.is_set():
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                try:
                    self.reason_bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    try:
                        self.reason_bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    except Exception:
                        pass
                self._stop.wait(0.005)
        except Exception:
            pass

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        time.sleep(0.02)
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=2)

    def summary(self):
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = sorted(name for bit, name in self.REASONS.items() if self.reason_bits & bit)
        return {"sm_mhz": statistics.median(self.sm), "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.sm)}


# ---------------------------------------------------------------------------------------------
# reference arm: the reference's CPU path on the host cores (bounded sample)
# ---------------------------------------------------------------------------------------------
def cpu_reference_pass(model, ms, shapes_subset=None, threads=None, seed=0):
    """Time ONE pass of the reference CPU path over one layer's linears for every M in ms.
    Returns (seconds, algorithmic bytes)."""
    import torch

    from oracle import torch_port

    if threads:
        torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(seed)
    total_t, total_b = 0.0, 0
    for name, K, N in linear_shapes(model):
        if shapes_subset and name not in shapes_subset:
            continue
        qw = torch.randint(-(2**31), 2**31 - 1, (K, N // 8), generator=g, dtype=torch.int32)
        qz = torch.randint(-(2**31), 2**31 - 1, (K // GROUP, N // 8), generator=g, dtype=torch.int32)
        sc = (torch.rand(K // GROUP, N, generator=g) * 0.02 + 0.005).half()
        for M in ms:
            x = torch.randn(M, K, generator=g).half()
            t0 = time.perf_counter()
            torch_port.awq_linear_cpu(x, qw, sc, qz)
            total_t += time.perf_counter() - t0
            total_b += algo_bytes(M, K, N)
        del qw, qz, sc
    return total_t, total_b


def run_reference(args, model):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    # size the per-step sample so that (steps + warmup) passes end within a few minutes
    probe_t, _ = cpu_reference_pass(model, (1,), shapes_subset=("o_proj",), threads=cores)
    o_bytes = [algo_bytes(1, K, N) for n, K, N in linear_shapes(model) if n == "o_proj"][0]
    layer_bytes = sum(algo_bytes(1, K, N) for _, K, N in linear_shapes(model))
    full_est = probe_t * layer_bytes / o_bytes * len(DECODE_MS)
    budget = 150.0 / max(1, args.steps + args.warmup)
    if full_est <= budget:
        subset, ms, sample = None, DECODE_MS, "one decoder layer (qkv,o,gate_up,down) at M=1,4,16 per step"
    elif full_est / len(DECODE_MS) <= budget:
        subset, ms, sample = None, (1,), "one decoder layer (qkv,o,gate_up,down) at M=1 per step"
    else:
        subset, ms, sample = ("qkv_proj", "o_proj"), (1,), "qkv_proj + o_proj of one layer at M=1 per step"
    for _ in range(args.warmup):
        cpu_reference_pass(model, ms, subset, cores)
    t_sum, b_sum = 0.0, 0
    for i in range(args.steps):
        t, b = cpu_reference_pass(model, ms, subset, cores, seed=i)
        t_sum += t
        b_sum += b
    gbs = b_sum / t_sum / 1e9
    line = {
        "impl": "reference", "metric": "awq_w4a16_decode_hbm_gbs", "value": gbs, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_sum / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": f"{args.workload} AWQ linear stack, decode M=1/4/16 (CPU sample: {sample})",
                   "group_size": GROUP},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def build_stack(model, tp_rank, tp_size, device, dtype, seed=1234, layers=None, keep_reference_layout_of_layer0=False):
    """All AWQ linears of the model through the public layer API, random-init, repacked.
    Returns (layers, ref0): ref0 holds CPU copies of layer 0's reference-layout tensors of the
    row-parallel linears (for the N > 1 parity check) when asked for."""
    import torch

    from sglang_awq_b200.srt.layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
    from sglang_awq_b200.srt.layers.quantization import AWQConfig

    cfg = AWQConfig.from_config({"w_bit": 4, "q_group_size": GROUP, "zero_point": True})
    g = torch.Generator(device=device).manual_seed(seed + tp_rank)
    out, ref0 = [], {}
    q = model["heads"] * model["head_dim"]
    with torch.device(device):
        for li in range(model["layers"] if layers is None else layers):
            pre = f"model.layers.{li}."
            mods = {
                "qkv_proj": QKVParallelLinear(model["hidden"], model["head_dim"], model["heads"], model["kv_heads"],
                                              bias=bool(model.get("qkv_bias", 0)), quant_config=cfg, params_dtype=dtype,
                                              prefix=pre + "self_attn.qkv_proj", tp_rank=tp_rank, tp_size=tp_size),
                "o_proj": RowParallelLinear(q, model["hidden"], bias=False, quant_config=cfg, params_dtype=dtype,
                                            prefix=pre + "self_attn.o_proj", tp_rank=tp_rank, tp_size=tp_size),
                "gate_up_proj": MergedColumnParallelLinear(model["hidden"], [model["inter"]] * 2, bias=False,
                                                           quant_config=cfg, params_dtype=dtype,
                                                           prefix=pre + "mlp.gate_up_proj", tp_rank=tp_rank,
                                                           tp_size=tp_size),
                "down_proj": RowParallelLinear(model["inter"], model["hidden"], bias=False, quant_config=cfg,
                                               params_dtype=dtype, prefix=pre + "mlp.down_proj", tp_rank=tp_rank,
                                               tp_size=tp_size),
            }
            for name, mod in mods.items():
                K = mod.qweight.shape[0]
                mod.qweight.data.copy_(torch.randint(-(2**31), 2**31 - 1, mod.qweight.shape, generator=g,
                                                     device=device, dtype=torch.int64).to(torch.int32))
                mod.qzeros.data.copy_(torch.randint(-(2**31), 2**31 - 1, mod.qzeros.shape, generator=g, device=device,
                                                    dtype=torch.int64).to(torch.int32))
                # gain ~1 per linear: sigma(w - z) ~ 6.2, so scale ~ 1 / (6.2 sqrt(K_total))
                mod.scales.data.copy_(((torch.rand(mod.scales.shape, generator=g, device=device) * 0.4 + 0.8)
                                       / (6.2 * (K * (tp_size if name in ROW_PARALLEL else 1)) ** 0.5)).to(dtype))
                if mod.bias is not None:
                    mod.bias.data.copy_((torch.randn(mod.bias.shape, generator=g, device=device) * 0.1).to(dtype))
                if keep_reference_layout_of_layer0 and li == 0 and name in ROW_PARALLEL:
                    ref0[name] = {k: getattr(mod, k).data.cpu() for k in ("qweight", "qzeros", "scales")}
                mod.quant_method.process_weights_after_loading(mod)
            out.append(mods)
    return out, ref0


def tp_parity_check(layers, ref0, model, world, rank, device, dtype):
    """N > 1, before timing: one M = 4 pass through layer 0's row-parallel linears WITH the all-reduce.
      allreduce_bit_exact : the collective's result equals fp16(sum over ranks, in rank order, in fp32)
                            of the all-gathered per-rank partials, bit for bit, on every rank
      ranks_identical     : every rank holds the same bits
      tp_rel_fro          : relative Frobenius error of the all-reduced result against the UNSHARDED
                            oracle result (each rank's shard through the CPU oracle's double-accumulated
                            linear over the reference-dequantised weights, summed over ranks)."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from sglang_awq_b200.srt import distributed as D

    res = {"allreduce_bit_exact": True, "ranks_identical": True, "tp_rel_fro": 0.0, "m": 4,
           "checked": list(ROW_PARALLEL), "all_reduce_kernel": D.get_tp_group().peer_ar is not None}
    try:
        from oracle import awq_oracle
        awq_oracle.lib()
        res["reference"] = "CPU oracle (oracle/awq_oracle.c linear_f32), summed over rank shards"
    except Exception:
        awq_oracle = None
        res["reference"] = "fp32 torch.matmul over the bit-exact dequant kernel (CPU oracle library not built)"
    g = torch.Generator(device=device).manual_seed(4242 + rank)
    for name in ROW_PARALLEL:
        mod = layers[0][name]
        Kr, N = mod.awq_in_features, mod.awq_out_features
        x = torch.randn(4, Kr, generator=g, device=device).to(dtype)
        with torch.no_grad():
            partial = mod(x, skip_all_reduce=True)[0].clone()
            reduced = mod(x)[0].clone()
        torch.cuda.synchronize()
        parts = [torch.empty_like(partial) for _ in range(world)]
        dist.all_gather(parts, partial)
        acc = parts[0].float()
        for p in parts[1:]:
            acc = acc + p.float()
        expect = acc.to(dtype)
        res["allreduce_bit_exact"] &= bool(torch.equal(expect.view(torch.int16), reduced.view(torch.int16)))
        alls = [torch.empty_like(reduced) for _ in range(world)]
        dist.all_gather(alls, reduced)
        res["ranks_identical"] &= all(bool(torch.equal(a.view(torch.int16), alls[0].view(torch.int16))) for a in alls)
        # unsharded reference = sum over ranks of each rank's shard through the oracle
        r0 = ref0[name]
        if awq_oracle is not None:
            bits = lambda t: t.contiguous().view(torch.int16).cpu().numpy().view(np.uint16)  # noqa: E731
            yp = awq_oracle.linear_f32(bits(x), r0["qweight"].numpy(), bits(r0["scales"]), r0["qzeros"].numpy(),
                                       0 if dtype == torch.float16 else 1)
            yp = torch.from_numpy(yp).to(device).double()
        else:
            from sglang_awq_b200 import _C
            w = _C.dequantize(r0["qweight"].to(device), r0["scales"].to(device), r0["qzeros"].to(device))
            yp = (x.float() @ w.float()).double()
        dist.all_reduce(yp)
        err = float((reduced.double() - yp).norm() / yp.norm().clamp_min(1e-30))
        res["tp_rel_fro"] = max(res["tp_rel_fro"], err)
    flag = torch.tensor([int(res["allreduce_bit_exact"]), int(res["ranks_identical"]), int(res["tp_rel_fro"] <= 1e-3)],
                        device=device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    fro = torch.tensor([res["tp_rel_fro"]], device=device, dtype=torch.float64)
    dist.all_reduce(fro, op=dist.ReduceOp.MAX)
    res["allreduce_bit_exact"], res["ranks_identical"] = bool(flag[0].item()), bool(flag[1].item())
    res["tp_rel_fro"] = float(fro.item())
    res["ok"] = bool(flag.min().item() == 1)
    return res


def run_ours(args, model):
    import torch
    import torch.distributed as dist

    from sglang_awq_b200 import _C
    from sglang_awq_b200.srt import distributed as D

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the AWQ kernels have no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
            os.environ.pop("NCCL_DEBUG", None)  # NCCL prints its version banner on stdout at VERSION / WARN
        D.init_distributed_environment(world_size=world, rank=rank, local_rank=local_rank, backend="nccl")
        D.initialize_model_parallel(world)
    L = _C.lib()  # fail loudly if the native library is missing
    dtype = torch.float16
    hbm_peak, tf_peak, peak_src = load_peaks()

    layers, ref0 = build_stack(model, rank if world > 1 else 0, world, device, dtype,
                               keep_reference_layout_of_layer0=world > 1)
    shapes = linear_shapes(model)
    graphs = []  # every CUDA graph of this process; destroyed before the process group (teardown order)

    parity = None
    if world > 1:
        parity = tp_parity_check(layers, ref0, model, world, rank, device, dtype)
        del ref0
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": "awq_w4a16_decode_hbm_gbs", "value": None, "n_gpus": world,
                                  "parity": parity, "error": "tensor-parallel parity check failed"}), flush=True)
            sys.stdout.flush()
            os._exit(3)

    def local_kn(name, K, N):
        return (K // world, N) if name in ROW_PARALLEL else (K, N // world)

    def routed_kernel(M, K, N):
        rid = int(L.awq_b200_gemv_route(M, K, N, 0)) if L.awq_b200_linear_route(M, K, N) == 0 else 6
        return GEMV_KERNEL_NAMES.get(rid, f"kernel#{rid}")

    # per-(M, linear) device inputs (rank-local K for row-parallel layers), views of ONE buffer so that
    # the e2e arm moves them with a single pinned copy
    in_elems = {name: local_kn(name, K, N)[0] for name, K, N in shapes}
    total_in = sum(M * in_elems[name] for M in DECODE_MS for name, _, _ in shapes)
    x_dev = torch.empty(total_in, dtype=dtype, device=device)
    x_host = torch.randn(total_in, dtype=torch.float32).to(dtype).pin_memory()
    x_dev.copy_(x_host)
    xs, off = {}, 0
    for M in DECODE_MS:
        for name, _, _ in shapes:
            n = M * in_elems[name]
            xs[(M, name)] = x_dev[off:off + n].view(M, in_elems[name])
            off += n
    out_elems = sum(M * model["hidden"] for M in DECODE_MS)
    y_dev = torch.empty(out_elems, dtype=dtype, device=device)   # last layer's down_proj outputs
    y_host = torch.empty(out_elems, dtype=dtype).pin_memory()

    def one_pass():
        launches, yoff = 0, 0
        for M in DECODE_MS:
            last = None
            for mods in layers:
                for name, _, _ in shapes:
                    last, _ = mods[name](xs[(M, name)])  # row-parallel layers all-reduce inside when world > 1
                    launches += 1
            y_dev[yoff:yoff + last.numel()].copy_(last.reshape(-1))
            yoff += last.numel()
        return launches

    stream = torch.cuda.Stream(device=device)
    stream.wait_stream(torch.cuda.current_stream())

    def capture(fn):
        """run once eagerly on the bench stream, then capture into a CUDA graph"""
        with torch.cuda.stream(stream):
            fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                fn()
        graphs.append(g)
        return g

    with torch.cuda.stream(stream):
        launches = one_pass()
        torch.cuda.synchronize()
    graph = capture(one_pass)
    replay = graph.replay

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(steps):
                fn()
            e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def e2e_step():
        x_dev.copy_(x_host, non_blocking=True)
        replay()
        y_host.copy_(y_dev, non_blocking=True)

    with torch.cuda.stream(stream):
        for _ in range(max(3, args.warmup)):
            replay()
    with ClockSampler(local_rank) as clocks:
        torch.cuda.profiler.start()  # lets `ncu --profile-from-start off` see exactly the timed launches
        ms_total = timed(replay, args.steps)
        torch.cuda.profiler.stop()
        for _ in range(2):
            with torch.cuda.stream(stream):
                e2e_step()
        ms_e2e = timed(e2e_step, args.steps)
    clk = clocks.summary()

    # ---- decode tokens/s per batch size (BASELINE.json metric, "tok/s at TP1/2/4/8"): one pass over
    # every AWQ linear of the model (+ the row-parallel all-reduces) at ONE M, CUDA-graph replay.
    # Linears only: attention, norms and sampling are outside this path. ----
    tok_s = {}
    for M in DECODE_MS:
        def pass_m(M=M):
            for mods in layers:
                for name, _, _ in shapes:
                    mods[name](xs[(M, name)])
        gm = capture(pass_m)
        with torch.cuda.stream(stream):
            gm.replay()
        reps = max(3, args.steps // 2)
        t_ms = timed(gm.replay, reps) / reps
        tok_s[f"M{M}"] = {"ms_per_pass": round(t_ms, 4), "tokens_per_s": round(M / (t_ms * 1e-3), 1),
                          "gbs": round(step_bytes(model, (M,)) / (t_ms * 1e-3) / 1e9, 1),
                          "frac_hbm_per_gpu": round(step_bytes(model, (M,)) / (t_ms * 1e-3) / 1e9 / world / hbm_peak, 3)}

    ar_launches = 0
    if world > 1 and D.get_tp_group().peer_ar is not None:
        ar_launches = len(DECODE_MS) * model["layers"] * len(ROW_PARALLEL)
    bytes_step = step_bytes(model)  # whole model, all ranks together
    ms_per_step = ms_total / args.steps
    value = bytes_step / (ms_per_step * 1e-3) / 1e9
    e2e_value = bytes_step / (ms_e2e / args.steps * 1e-3) / 1e9

    def time_calls(fn, iters):
        with torch.cuda.stream(stream):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(iters):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters * 1e-3

    def per_launch_seconds(mods, x, iters=10, skip_all_reduce=True):
        """mean device time of one launch: the layers' instances of one linear (distinct weights,
        >> L2 together) back to back in a CUDA graph"""
        def fn():
            for m_ in mods:
                if hasattr(m_, "reduce_results"):
                    m_(x, skip_all_reduce=skip_all_reduce)
                else:
                    m_(x)
        g = capture(fn)
        t = time_calls(g.replay, iters) / len(mods)
        graphs.remove(g)
        del g
        return t

    def shape_table(stack, shp, ms, prefill_ms=(), iters=10):
        """per (linear, M): us per launch, algorithmic GB/s, fraction of the measured HBM peak, routed
        kernel; prefill rows in TFLOP/s; one standalone-dequant row per linear."""
        tab = {}
        gen = torch.Generator(device=device).manual_seed(7)
        # every decode row before any prefill row: a tensor-bound GEMM burst leaves the board at its power cap for a
        # while, and the HBM-bound rows measured right behind it came out 10-40 % slow
        for name, K, N in shp:
            Kl, Nl = local_kn(name, K, N)
            mods = [l[name] for l in stack]
            for M in ms:
                x = torch.randn(M, Kl, generator=gen, device=device).to(dtype)
                t = per_launch_seconds(mods, x, iters)
                b = algo_bytes(M, Kl, Nl)
                tab[f"{name}_M{M}"] = {"us": round(t * 1e6, 2), "gbs": round(b / t / 1e9, 1),
                                       "frac_hbm": round(b / t / 1e9 / hbm_peak, 3), "kernel": routed_kernel(M, Kl, Nl)}
        for name, K, N in shp:
            Kl, Nl = local_kn(name, K, N)
            mods = [l[name] for l in stack]
            for Mp in prefill_ms:
                torch.cuda.synchronize()
                time.sleep(0.3)  # every prefill row starts from an un-capped board (bursts of 8 GEMMs reach the power cap)
                xp = torch.randn(Mp, Kl, generator=gen, device=device).to(dtype)
                sub = mods[:max(2, min(8, len(mods)))]
                t = time_calls(lambda: [m_(xp, skip_all_reduce=True) if hasattr(m_, "reduce_results") else m_(xp)
                                        for m_ in sub], 3) / len(sub)
                tf = 2.0 * Mp * Kl * Nl / t / 1e12
                tab[f"{name}_M{Mp}"] = {"us": round(t * 1e6, 1), "tflops": round(tf, 1),
                                        "frac_tensor": round(tf / tf_peak, 3), "kernel": routed_kernel(Mp, Kl, Nl)}
                del xp
        return tab

    # ---- roofline of the dominant kernel: the decode kernel on gate_up_proj (largest share of the
    # step's device time, profiles/*launch_list*), measured live: all gate_up launches of one M in a
    # CUDA graph (distinct weights >> L2), CUDA events on the launching stream ----
    # (timed ALONE, like the burst peak it is compared with: MEASURED_PEAKS.json's HBM figure is a best-of-10 copy on an
    # idle board, so the board gets half a second to leave the power cap the 350 ms timed region put it in)
    torch.cuda.synchronize()
    time.sleep(0.5)
    dom_name, dom_K, dom_N = [s_ for s_ in shapes if s_[0] == "gate_up_proj"][0]
    dom_Kl, dom_Nl = local_kn(dom_name, dom_K, dom_N)
    dom_mods = [l[dom_name] for l in layers]
    dom = {"bytes": 0, "sec": 0.0, "launches": 0, "per_m": {}}
    traffic_entries = []
    for M in DECODE_MS:
        torch.cuda.synchronize()
        time.sleep(0.3)
        t = per_launch_seconds(dom_mods, xs[(M, dom_name)])
        b = algo_bytes(M, dom_Kl, dom_Nl)
        kname = routed_kernel(M, dom_Kl, dom_Nl)
        dom["bytes"] += b
        dom["sec"] += t
        dom["launches"] += 1
        dom["per_m"][f"M{M}"] = {"us": round(t * 1e6, 2), "gbs": round(b / t / 1e9, 1),
                                 "frac": round(b / t / 1e9 / hbm_peak, 3), "kernel": kname}
        traffic_entries.append(PROFILED_DRAM_TRAFFIC.get((kname, dom_Kl, dom_Nl, M)))
    dom_gbs = dom["bytes"] / dom["sec"] / 1e9
    if all(e is not None for e in traffic_entries):
        traffic = sum(e["dram_bytes"] for e in traffic_entries) / len(traffic_entries)
        traffic_src = [{"kernel": e["kernel"], "M": e["M"], "dram_bytes": e["dram_bytes"], "source": e["source"]}
                       for e in traffic_entries]
    else:
        traffic, traffic_src = None, "no ncu --set full capture of the routed kernel on this shape (profiles/dram_traffic.json)"

    # share of the step spent in the dominant kernel / in the collective (from the live timings)
    dom_share = dom["sec"] * model["layers"] / (ms_per_step * 1e-3)

    # ---- extras (rank 0, N = 1 only) ----
    extras = {}
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_extras:
        per_shape = shape_table(layers, shapes, (1, 4, 16), prefill_ms=(512, 4096))
        extras["per_shape"] = per_shape
        # standalone dequant op (kernel 1, the one bit-exact op): reference-layout tensors of each shape
        torch.cuda.synchronize()
        time.sleep(0.5)  # let the board leave its power cap after the prefill rows (the op is HBM-bound)
        deq = {}
        gen = torch.Generator(device=device).manual_seed(5)
        for name, K, N in shapes:
            sets = []
            for _ in range(max(2, min(6, (768 << 20) // dequant_bytes(K, N) + 1))):
                qw = torch.randint(-(2**31), 2**31 - 1, (K, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                qz = torch.randint(-(2**31), 2**31 - 1, (K // GROUP, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                sc = (torch.rand(K // GROUP, N, generator=gen, device=device) * 0.02 + 0.005).to(dtype)
                sets.append((qw, sc, qz, torch.empty(K, N, device=device, dtype=dtype)))

            def deq_all(sets=sets):
                for qw, sc, qz, o in sets:
                    L.awq_b200_dequantize(qw.data_ptr(), sc.data_ptr(), qz.data_ptr(), o.data_ptr(), qw.shape[0],
                                          qw.shape[1] * 8, sc.shape[0], 0, torch.cuda.current_stream().cuda_stream)
            t = time_calls(deq_all, 5) / len(sets)
            deq[name] = {"us": round(t * 1e6, 1), "gbs": round(dequant_bytes(K, N) / t / 1e9, 1),
                         "frac_hbm": round(dequant_bytes(K, N) / t / 1e9 / hbm_peak, 3)}
            del sets
            torch.cuda.empty_cache()
        extras["dequant_op"] = deq

        # group-64 / group-32 checkpoints (same reference op: group = K / scales rows): awq_b200_linear_g on the
        # dominant shape, M = 1 / 4 / 8 -> the sub-group decode kernels (awq_gemv_sub.cu), M = 16 -> the tcgen05 kernel
        grp = {}
        gname, gK, gN = max(shapes, key=lambda s: s[1] * s[2])
        for G in (64, 32):
            rec_bytes = int(L.awq_b200_tile_bytes(G))
            nbytes = int(L.awq_b200_num_tiles(gK, gN)) * rec_bytes
            sets = []
            for _ in range(max(2, min(6, (768 << 20) // nbytes + 1))):
                qw = torch.randint(-(2**31), 2**31 - 1, (gK, gN // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                qz = torch.randint(-(2**31), 2**31 - 1, (gK // G, gN // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                sc = (torch.rand(gK // G, gN, generator=gen, device=device) * 0.02 + 0.005).to(dtype)
                sets.append(_C.repack_g(qw, sc, qz))
                del qw, qz, sc
            for M in (1, 4, 8, 16):
                xg = torch.randn(M, gK, device=device).to(dtype)

                def g_all(xg=xg, sets=sets):
                    for t in sets:
                        _C.linear_g(xg, t, gN)
                gg = capture(g_all)
                t = time_calls(gg.replay, 5) / len(sets)
                grp[f"{gname}_G{G}_M{M}"] = {"us": round(t * 1e6, 2), "gbs": round(nbytes / t / 1e9, 1),
                                              "frac_hbm": round(nbytes / t / 1e9 / hbm_peak, 3),
                                              "kernel": ("awq::gemv_sub_kernel" if (M * (128 // G) <= 8 and dtype == torch.float16) else
                                                         "awq::gemv_sub_pw_kernel" if M <= 8 else "awq::gemm_kernel")}
                graphs.remove(gg)
                del gg
            del sets
            torch.cuda.empty_cache()
        extras["group_size_decode"] = grp

        # BASELINE.json configs[1] + [2]: Llama-3-8B table (decode M = 1/4/16 + prefill M = 4096)
        small = None
        if args.workload != "llama3-8b":
            m8 = MODELS["llama3-8b"]
            small, _ = build_stack(m8, 0, 1, device, dtype, seed=77)
            torch.cuda.synchronize()
            time.sleep(0.5)
            extras["llama3_8b_per_shape"] = shape_table(small, linear_shapes(m8), (1, 4, 16), prefill_ms=(512, 4096))
            pass8 = {}
            for M in DECODE_MS:
                x8 = {name: torch.randn(M, K, device=device).to(dtype) for name, K, _ in linear_shapes(m8)}

                def p8(x8=x8):
                    for mods in small:
                        for name, _, _ in linear_shapes(m8):
                            mods[name](x8[name])
                g8 = capture(p8)
                t8 = time_calls(g8.replay, 5)
                pass8[f"M{M}"] = {"ms_per_pass": round(t8 * 1e3, 4), "gbs": round(step_bytes(m8, (M,)) / t8 / 1e9, 1),
                                  "frac_hbm": round(step_bytes(m8, (M,)) / t8 / 1e9 / hbm_peak, 3)}
                graphs.remove(g8)
                del g8
            extras["llama3_8b_whole_pass"] = pass8

        # the reference's own CUDA path on THIS GPU: its dequant kernel (compiled unmodified for
        # sm_100a, oracle/_ref) -> torch.matmul (cuBLAS), AWQLinearMethod.apply of awq.py:434-451,
        # same CUDA-graph protocol, reference-layout tensors; Llama-3-8B shapes + the config-1 shape
        try:
            from oracle import build_ref

            ref_fn = build_ref.load()
        except Exception:
            ref_fn = None
        ref_shapes = linear_shapes(MODELS["llama3-8b"]) + [("llama2_7b_mlp_up", 4096, 11008)]
        ours_src = small if small is not None else layers
        if ref_fn is not None:
            ref_gpu = {}
            gen = torch.Generator(device=device).manual_seed(99)
            for name, K, N in ref_shapes:
                sets = []
                for _ in range(max(2, min(8, (600 << 20) // (K * N // 2)))):
                    qw = torch.randint(-(2**31), 2**31 - 1, (K, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                    qz = torch.randint(-(2**31), 2**31 - 1, (K // GROUP, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                    sc = (torch.rand(K // GROUP, N, generator=gen, device=device) * 0.02 + 0.005).to(dtype)
                    sets.append((qw, sc, qz))
                tiles = [_C.repack(qw, sc, qz) for qw, sc, qz in sets]
                for M in (1, 16, 512, 4096):
                    x = torch.randn(M, K, generator=gen, device=device).to(dtype)

                    def ref_all(x=x, sets=sets):
                        for qw, sc, qz in sets:
                            torch.matmul(x, ref_fn(qw, sc, qz))

                    def ours_all(x=x, tiles=tiles, N=N):
                        for tl in tiles:
                            _C.linear(x, tl, N)
                    if M <= 16:
                        gr, go = capture(ref_all), capture(ours_all)
                        tr, to = time_calls(gr.replay, 5) / len(sets), time_calls(go.replay, 5) / len(sets)
                        graphs.remove(gr), graphs.remove(go)
                        del gr, go
                    else:
                        tr, to = time_calls(ref_all, 3) / len(sets), time_calls(ours_all, 3) / len(sets)
                    row = {"reference_us": round(tr * 1e6, 2), "ours_us": round(to * 1e6, 2),
                           "speedup_ours": round(tr / to, 2)}
                    if M >= 512:
                        row["reference_tflops"] = round(2.0 * M * K * N / tr / 1e12, 1)
                        row["ours_tflops"] = round(2.0 * M * K * N / to / 1e12, 1)
                    ref_gpu[f"{name}_M{M}"] = row
                del sets, tiles
                torch.cuda.empty_cache()
            extras["reference_cuda_path_same_gpu"] = ref_gpu
        # baseline C: the reference's AWQ-Marlin path (awq_marlin_repack + gptq_marlin_gemm compiled unmodified from
        # its sources, python permutes restated in oracle/marlin_ref.py): what an AWQ checkpoint resolves to by
        # default in the reference (awq.py:267-289); same shapes and protocol
        try:
            from oracle import marlin_ref

            have_marlin = marlin_ref.load()
        except Exception:
            have_marlin = False
        if have_marlin:
            mar = {}
            gen = torch.Generator(device=device).manual_seed(98)
            for name, K, N in ref_shapes:
                sets = []
                for _ in range(max(2, min(8, (600 << 20) // (K * N // 2)))):
                    qw = torch.randint(-(2**31), 2**31 - 1, (K, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                    qz = torch.randint(-(2**31), 2**31 - 1, (K // GROUP, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                    sc = (torch.rand(K // GROUP, N, generator=gen, device=device) * 0.02 + 0.005).to(dtype)
                    sets.append((marlin_ref.prepare(qw, sc, qz), _C.repack(qw, sc, qz)))
                for M in (1, 4, 16, 64, 512, 4096):
                    x = torch.randn(M, K, generator=gen, device=device).to(dtype)

                    def mar_all(x=x, sets=sets):
                        for mw, _ in sets:
                            marlin_ref.linear(x, mw)

                    def ours_all(x=x, sets=sets, N=N):
                        for _, tl in sets:
                            _C.linear(x, tl, N)
                    if M <= 64:
                        gr, go = capture(mar_all), capture(ours_all)
                        tr, to = time_calls(gr.replay, 5) / len(sets), time_calls(go.replay, 5) / len(sets)
                        graphs.remove(gr), graphs.remove(go)
                        del gr, go
                    else:
                        tr, to = time_calls(mar_all, 3) / len(sets), time_calls(ours_all, 3) / len(sets)
                    row = {"marlin_us": round(tr * 1e6, 2), "ours_us": round(to * 1e6, 2), "speedup_ours": round(tr / to, 2)}
                    if M >= 512:
                        row["marlin_tflops"] = round(2.0 * M * K * N / tr / 1e12, 1)
                        row["ours_tflops"] = round(2.0 * M * K * N / to / 1e12, 1)
                    else:
                        row["marlin_frac_hbm"] = round(algo_bytes(M, K, N) / tr / 1e9 / hbm_peak, 3)
                        row["ours_frac_hbm"] = round(algo_bytes(M, K, N) / to / 1e9 / hbm_peak, 3)
                    mar[f"{name}_M{M}"] = row
                del sets
                torch.cuda.empty_cache()
            extras["reference_marlin_same_gpu"] = mar
        # bf16 (AWQMarlinConfig advertises it, awq.py:237-239): the same kernels' bf16 instantiations on the Llama-3-8B
        # shapes -- decode M = 1 / 16 (per-weight-scaled mma.sync path: the bf16 significand cannot hold the subnormal
        # trick) and prefill M = 4096 (tcgen05, kind::f16 with bf16 operands)
        torch.cuda.synchronize()
        time.sleep(0.5)
        bf = {}
        gen = torch.Generator(device=device).manual_seed(97)
        for name, K, N in linear_shapes(MODELS["llama3-8b"]):
            sets = []
            for _ in range(max(2, min(8, (600 << 20) // (K * N // 2)))):
                qw = torch.randint(-(2**31), 2**31 - 1, (K, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                qz = torch.randint(-(2**31), 2**31 - 1, (K // GROUP, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                sc = (torch.rand(K // GROUP, N, generator=gen, device=device) * 0.02 + 0.005).to(torch.bfloat16)
                sets.append(_C.repack(qw, sc, qz))
            for M in (1, 16, 4096):
                x = torch.randn(M, K, generator=gen, device=device).to(torch.bfloat16)

                def bf_all(x=x, sets=sets, N=N):
                    for tl in sets:
                        _C.linear(x, tl, N)
                if M <= 16:
                    gb = capture(bf_all)
                    t = time_calls(gb.replay, 5) / len(sets)
                    graphs.remove(gb)
                    del gb
                    bf[f"{name}_M{M}"] = {"us": round(t * 1e6, 2), "frac_hbm": round(algo_bytes(M, K, N) / t / 1e9 / hbm_peak, 3)}
                else:
                    time.sleep(0.3)
                    t = time_calls(bf_all, 3) / len(sets)
                    bf[f"{name}_M{M}"] = {"us": round(t * 1e6, 1), "tflops": round(2.0 * M * K * N / t / 1e12, 1),
                                          "frac_tensor": round(2.0 * M * K * N / t / 1e12 / tf_peak, 3)}
            del sets
            torch.cuda.empty_cache()
        extras["bf16_llama3_8b_per_shape"] = bf

        tfs = [v["tflops"] for k, v in (extras.get("llama3_8b_per_shape") or per_shape).items() if k.endswith("_M4096")]
        extras["prefill_m4096_tflops_mean_llama3_8b"] = round(sum(tfs) / len(tfs), 1)
        extras["prefill_frac_of_measured_bf16_peak"] = round(sum(tfs) / len(tfs) / tf_peak, 3)
        del small
        torch.cuda.empty_cache()

        # BASELINE.json configs[0] / BASELINE.md §3 (B): Llama-2-7B MLP matrix (K=4096, N=11008) at M = 1 and
        # M = 512, the GPU kernel beside the reference's CPU path (fp16 matmul and fp32-upcast matmul),
        # 3 warm-ups + 5 reps, median, host core count stated
        from oracle import torch_port

        cores = len(os.sched_getaffinity(0))
        torch.set_num_threads(cores)
        K1, N1 = 4096, 11008
        gcpu = torch.Generator().manual_seed(3)
        qw = torch.randint(-(2**31), 2**31 - 1, (K1, N1 // 8), generator=gcpu, dtype=torch.int32)
        qz = torch.randint(-(2**31), 2**31 - 1, (K1 // GROUP, N1 // 8), generator=gcpu, dtype=torch.int32)
        sc = (torch.rand(K1 // GROUP, N1, generator=gcpu) * 0.02 + 0.005).half()
        tl1 = [_C.repack(qw.to(device), sc.to(device), qz.to(device)) for _ in range(24)]  # 24 x 23.5 MB >> L2
        cfg1 = {"shape": [K1, N1], "cores": cores}
        for M in (1, 512):
            x = torch.randn(M, K1, generator=gcpu).half()
            xg = x.to(device)

            def ours1(xg=xg):
                for tl in tl1:
                    _C.linear(xg, tl, N1)
            if M <= 16:
                g1 = capture(ours1)
                tg = time_calls(g1.replay, 10) / len(tl1)
                graphs.remove(g1)
                del g1
            else:
                tg = time_calls(ours1, 5) / len(tl1)
            row = {"gpu_us": round(tg * 1e6, 2), "gpu_kernel": routed_kernel(M, K1, N1)}
            if M == 1:
                row["gpu_frac_hbm"] = round(algo_bytes(M, K1, N1) / tg / 1e9 / hbm_peak, 3)
            else:
                row["gpu_tflops"] = round(2.0 * M * K1 * N1 / tg / 1e12, 1)
                row["gpu_frac_tensor"] = round(2.0 * M * K1 * N1 / tg / 1e12 / tf_peak, 3)
            for label, mm_dtype in (("cpu_fp16_matmul", None), ("cpu_fp32_upcast_matmul", torch.float32)):
                t0 = time.perf_counter()
                torch_port.awq_linear_cpu(x, qw, sc, qz, matmul_dtype=mm_dtype)
                first = time.perf_counter() - t0
                warm, reps = (2, 5) if first < 1.5 else ((1, 3) if first < 6 else (0, 1))
                for _ in range(warm):
                    torch_port.awq_linear_cpu(x, qw, sc, qz, matmul_dtype=mm_dtype)
                ts = []
                for _ in range(reps):
                    t0 = time.perf_counter()
                    torch_port.awq_linear_cpu(x, qw, sc, qz, matmul_dtype=mm_dtype)
                    ts.append(time.perf_counter() - t0)
                row[label] = {"median_ms": round(statistics.median(ts) * 1e3, 2), "warmups": warm + 1, "reps": reps,
                              "gpu_speedup": round(statistics.median(ts) / tg, 1)}
            cfg1[f"M{M}"] = row
        extras["config1_llama2_7b_mlp_4096x11008"] = cfg1
        del tl1
        torch.cuda.empty_cache()

        # CPU baseline: the reference's CPU path (torch port) on this box's host cores, bounded sample
        t_cpu, b_cpu, n_pass = 0.0, 0, 0
        t_start = time.perf_counter()
        sample_model = MODELS["llama3-8b"]  # one decoder layer of the 8B shapes: ~2 s per pass on 16 cores
        while time.perf_counter() - t_start < 12.0 and n_pass < 6:
            t, b = cpu_reference_pass(sample_model, DECODE_MS, None, cores, seed=n_pass)
            t_cpu += t
            b_cpu += b
            n_pass += 1
        cpu_baseline = {"value": b_cpu / t_cpu / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
                        "sample": f"{n_pass} Llama-3-8B decoder layer(s) x (qkv,o,gate_up,down) x M=1,4,16 via torch "
                                  f"dequant+matmul (oracle/torch_port.py, pinned to the reference's golden vectors), "
                                  f"{t_cpu:.1f} s"}

    line = None
    if rank == 0:
        if world == 1:
            ar_desc = None
        elif ar_launches:
            ar_desc = "peer-memory all-reduce kernel (awq_b200_allreduce: awq::allreduce_ll_kernel / awq::allreduce_oneshot_kernel)"
        else:
            ar_desc = "ncclAllReduce"
        line = {
            "metric": "awq_w4a16_decode_hbm_gbs", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {
                "workload": f"{args.workload} AWQ: all {model['layers'] * 4} linears x decode M={','.join(map(str, DECODE_MS))} per step"
                            + (f", tensor-parallel TP={world} ({ar_desc} after o_proj/down_proj)" if world > 1 else ""),
                "group_size": GROUP, "launches_per_step": launches, "cuda_graph": True,
                "l2": f"distinct weights per step {step_bytes(model, (1,)) / 1e9 / world:.2f} GB per GPU >> 126 MB L2 (no flush needed)",
                "pdl": True, "all_reduce": ar_desc,
            },
            "decode_tokens_per_s": {k: v["tokens_per_s"] for k, v in tok_s.items()},
            "roofline": {"bound": "hbm", "achieved": dom_gbs, "peak": hbm_peak, "unit": "GB/s",
                         "frac": dom_gbs / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "kernel": f"decode kernels on {dom_name} (K={dom_Kl}, N={dom_Nl}"
                                   + ("" if world == 1 else "; per-rank shard")
                                   + f"), mean over its M={','.join(map(str, DECODE_MS))} launches; routed kernel per M in per_m",
                         "algorithmic_bytes_per_launch": dom["bytes"] / dom["launches"],
                         "us_per_launch": dom["sec"] / dom["launches"] * 1e6, "per_m": dom["per_m"],
                         "share_of_step": round(dom_share, 3),
                         "whole_step": {"achieved": value / world, "frac": value / world / hbm_peak,
                                        "note": f"all {launches} linear launches of the step"
                                                + (f" + {ar_launches} all-reduce launches" if ar_launches else "")
                                                + ", per GPU"}},
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": int(x_host.numel() * 2),
                    "d2h_bytes_per_step": int(y_host.numel() * 2)},
            # our kernels inside the timed region, per rank: one fused linear per layer call, plus one
            # peer-memory all-reduce per row-parallel call when N > 1 (NCCL launches are not counted)
            "gpu_launches": (launches + ar_launches) * args.steps,
            "clocks": clk,
            "decode_tokens_per_s_linears_only": tok_s,
        }
        if parity is not None:
            line["parity"] = parity
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        if extras:
            line["extras"] = extras
        print(json.dumps(line), flush=True)

    # ---- teardown: graphs first (they hold the all-reduce's staging pointers), then the peer mapping,
    # then the process group.  A watchdog ends the process if a collective destructor still blocks. ----
    sys.stdout.flush()
    if world > 1:
        def _watchdog():
            time.sleep(30)
            sys.stderr.write("[bench] teardown did not finish within 30 s; exiting\n")
            sys.stderr.flush()
            os._exit(0)
        threading.Thread(target=_watchdog, daemon=True).start()
        barrier()
        graphs.clear()
        del graph, replay
        torch.cuda.synchronize()
        D.destroy_model_parallel()
        dist.barrier()
        D.destroy_distributed_environment()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(MODELS), default="llama3-70b")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--ms", default=None, help="decode batch sizes of the step, e.g. 1,2,4,8,16,32,64 "
                                               "(default 1,4,16 = the BASELINE.json headline config; implies --no-extras)")
    args = ap.parse_args()
    if args.ms:
        global DECODE_MS
        DECODE_MS = tuple(int(v) for v in args.ms.split(","))
        assert all(1 <= m <= 4096 for m in DECODE_MS)
        args.no_extras = True
    model = MODELS[args.workload]
    if args.impl == "reference":
        run_reference(args, model)
    else:
        run_ours(args, model)


if __name__ == "__main__":
    main()
