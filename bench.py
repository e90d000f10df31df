#!/usr/bin/env python
"""bench.py -- headline benchmark of the AWQ W4A16 quantized-linear hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Workload (BASELINE.json configs[1], "Llama-3-8B-AWQ all linear shapes, decode GEMV M=1/4/16"):
one STEP = one decode pass over every AWQ linear of the model -- 32 layers x [qkv_proj, o_proj,
gate_up_proj, down_proj] -- at M = 1, 4 and 16, i.e. 384 fused W4A16 launches over 3.6 GB of
distinct packed weights (>> the 126 MB L2, so HBM is measured, not L2).  Synthetic activations,
random-init AWQ tensors of the real shapes.

  value   GB/s = ALGORITHMIC bytes of the step (packed weights + scales + zeros + x + y per
          linear) / device time; inputs resident in HBM, CUDA-graph replay (decode
          runs under CUDA graphs in the reference too), CUDA events, max over ranks.
  e2e     same metric through the public layer API with HOST buffers: each step copies its
          activations from pinned host memory, replays, and copies the results back.
  N > 1   the same model tensor-parallel over N ranks (column-parallel N-splits, row-parallel
          K-splits closed by an NCCL all-reduce): fixed total work -> "strong" scaling.
  --impl reference   the reference's own CPU path (torch dequant + matmul, oracle/torch_port.py)
          on the host cores, bounded sample, same metric.

Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {
    # name: hidden, intermediate, q heads, kv heads, head dim, layers
    "llama3-8b": dict(hidden=4096, inter=14336, heads=32, kv_heads=8, head_dim=128, layers=32),
    "llama3-70b": dict(hidden=8192, inter=28672, heads=64, kv_heads=8, head_dim=128, layers=80),
    "qwen2.5-32b": dict(hidden=5120, inter=27648, heads=40, kv_heads=8, head_dim=128, layers=64),
    "llama2-7b-mlp": dict(hidden=4096, inter=11008, heads=32, kv_heads=32, head_dim=128, layers=32),
}
DECODE_MS = (1, 4, 16)
GROUP = 128
# dram__bytes_read.sum + dram__bytes_write.sum of one gate_up_proj decode launch (ncu --set full,
# profiles/r01b_gemv_m1_gate_up.txt / r01b_gemv_m16_gate_up.txt: 61.5 - 61.6 MB read + 0.2 - 0.3 MB written)
DOMINANT_KERNEL_DRAM_BYTES_PER_LAUNCH = 61.8e6


def linear_shapes(m):
    """(name, K, N) of the four AWQ linears of one decoder layer (unsharded)."""
    q = m["heads"] * m["head_dim"]
    kv = m["kv_heads"] * m["head_dim"]
    return [("qkv_proj", m["hidden"], q + 2 * kv), ("o_proj", q, m["hidden"]),
            ("gate_up_proj", m["hidden"], 2 * m["inter"]), ("down_proj", m["inter"], m["hidden"])]


def algo_bytes(M, K, N):
    """K*N/2 [qweight] + (K/G)*N*2 [scales] + (K/G)*N/2 [qzeros] + M*K*2 [x] + M*N*2 [y]."""
    return K * N // 2 + (K // GROUP) * N * 2 + (K // GROUP) * N // 2 + M * K * 2 + M * N * 2


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
        self._nvml = None

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
    full_est = probe_t * 14.0 * len(DECODE_MS)  # one layer is ~14x o_proj's bytes
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
def build_stack(model, tp_rank, tp_size, device, dtype, seed=1234):
    """All AWQ linears of the model through the public layer API, random-init, repacked."""
    import torch

    from sglang_awq_b200.srt.layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
    from sglang_awq_b200.srt.layers.quantization import AWQConfig

    cfg = AWQConfig.from_config({"w_bit": 4, "q_group_size": GROUP, "zero_point": True})
    g = torch.Generator(device=device).manual_seed(seed + tp_rank)
    layers = []
    q = model["heads"] * model["head_dim"]
    with torch.device(device):
        for li in range(model["layers"]):
            pre = f"model.layers.{li}."
            mods = {
                "qkv_proj": QKVParallelLinear(model["hidden"], model["head_dim"], model["heads"], model["kv_heads"],
                                              bias=False, quant_config=cfg, params_dtype=dtype,
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
            for mod in mods.values():
                K = mod.qweight.shape[0]
                mod.qweight.data.copy_(torch.randint(-(2**31), 2**31 - 1, mod.qweight.shape, generator=g,
                                                     device=device, dtype=torch.int64).to(torch.int32))
                mod.qzeros.data.copy_(torch.randint(-(2**31), 2**31 - 1, mod.qzeros.shape, generator=g, device=device,
                                                    dtype=torch.int64).to(torch.int32))
                # gain ~1 per linear: sigma(w - z) ~ 6.2, so scale ~ 1 / (6.2 sqrt(K))
                mod.scales.data.copy_(((torch.rand(mod.scales.shape, generator=g, device=device) * 0.4 + 0.8)
                                       / (6.2 * K ** 0.5)).to(dtype))
                mod.quant_method.process_weights_after_loading(mod)
            layers.append(mods)
    return layers


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
    _C.lib()  # fail loudly if the native library is missing
    dtype = torch.float16
    hbm_peak, tf_peak, peak_src = load_peaks()

    layers = build_stack(model, rank if world > 1 else 0, world, device, dtype)
    shapes = linear_shapes(model)
    row_parallel = {"o_proj", "down_proj"}

    # per-(M, linear) device inputs (rank-local K for row-parallel layers), views of ONE buffer so that
    # the e2e arm moves them with a single pinned copy
    in_elems = {name: (K // world if name in row_parallel else K) for name, K, _ in shapes}
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
    graph = None
    with torch.cuda.stream(stream):
        launches = one_pass()
        torch.cuda.synchronize()
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                one_pass()
        except Exception as e:  # e.g. a collective that cannot be captured: time eager launches instead
            graph = None
            if rank == 0:
                print(f"[bench] CUDA graph capture failed ({type(e).__name__}: {e}); timing eager launches",
                      file=sys.stderr)
        torch.cuda.synchronize()

    def replay():
        if graph is not None:
            graph.replay()
        else:
            one_pass()

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
        with torch.cuda.stream(stream):
            pass_m()
            torch.cuda.synchronize()
            gm = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gm, stream=stream):
                pass_m()
            gm.replay()
        t_ms = timed(gm.replay, max(3, args.steps // 2)) / max(3, args.steps // 2)
        tok_s[f"M{M}"] = {"ms_per_pass": round(t_ms, 4), "tokens_per_s": round(M / (t_ms * 1e-3), 1),
                          "gbs": round(step_bytes(model, (M,)) / (t_ms * 1e-3) / 1e9, 1)}
        del gm

    ar_launches = 0
    if world > 1 and D.get_tp_group().peer_ar is not None:
        ar_launches = len(DECODE_MS) * model["layers"] * len(row_parallel)
    bytes_step = step_bytes(model)  # whole model, all ranks together
    ms_per_step = ms_total / args.steps
    value = bytes_step / (ms_per_step * 1e-3) / 1e9
    e2e_value = bytes_step / (ms_e2e / args.steps * 1e-3) / 1e9

    def finish():
        """N > 1: leave together, then exit without tearing the NCCL communicator down -- its
        destructor can block while CUDA graphs that captured its collectives are still alive
        (observed: the line was printed, then the ranks hung until the launcher's timeout)."""
        if world > 1:
            barrier()
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(0)

    if rank != 0:
        finish()
        return

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

    # ---- roofline of the dominant kernel: the decode kernel on gate_up_proj (largest share of the
    # step's device time, profiles/*launch_list*), measured live: the 32 gate_up launches of one
    # M in a CUDA graph (distinct weights, 1.9 GB >> L2), CUDA events on the launching stream ----
    dom_name, dom_K, dom_N = [s_ for s_ in shapes if s_[0] == "gate_up_proj"][0]
    dom_N_local = dom_N // world
    dom_mods = [l[dom_name] for l in layers]
    dom = {"bytes": 0, "sec": 0.0, "launches": 0, "per_m": {}}
    for M in DECODE_MS:
        x = xs[(M, dom_name)]
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(stream):
            for m_ in dom_mods:
                m_(x)
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=stream):
                for m_ in dom_mods:
                    m_(x)
        t = time_calls(g.replay, 10) / len(dom_mods)
        b = algo_bytes(M, dom_K, dom_N_local)
        dom["bytes"] += b
        dom["sec"] += t
        dom["launches"] += 1
        dom["per_m"][f"M{M}"] = {"us": round(t * 1e6, 2), "gbs": round(b / t / 1e9, 1), "frac": round(b / t / 1e9 / hbm_peak, 3)}
    dom_gbs = dom["bytes"] / dom["sec"] / 1e9

    # ---- extras (rank 0, N = 1 only): per-shape table, prefill TFLOP/s, CPU baseline ----
    extras = {}
    cpu_baseline = None
    if world == 1 and not args.no_extras:
        per_shape = {}
        for name, K, N in shapes:
            mods = [l[name] for l in layers]
            for M in (1, 16):
                x = xs[(M, name)]
                g = torch.cuda.CUDAGraph()
                with torch.cuda.stream(stream):
                    for m_ in mods:
                        m_(x)
                    torch.cuda.synchronize()
                    with torch.cuda.graph(g, stream=stream):
                        for m_ in mods:
                            m_(x)
                t = time_calls(g.replay, 10) / len(mods)
                per_shape[f"{name}_M{M}"] = {"us": round(t * 1e6, 2), "gbs": round(algo_bytes(M, K, N) / t / 1e9, 1),
                                             "frac_hbm": round(algo_bytes(M, K, N) / t / 1e9 / hbm_peak, 3)}
            Mp = 4096
            xp = torch.randn(Mp, K, device=device, dtype=torch.float32).to(dtype)
            t = time_calls(lambda: [m_(xp) for m_ in mods[:8]], 3) / 8
            tf = 2.0 * Mp * K * N / t / 1e12
            per_shape[f"{name}_M{Mp}"] = {"us": round(t * 1e6, 1), "tflops": round(tf, 1),
                                          "frac_tensor": round(tf / tf_peak, 3)}
            del xp
        extras["per_shape"] = per_shape
        # the reference's own CUDA path on THIS GPU: its dequant kernel (compiled unmodified for
        # sm_100a, oracle/_ref) -> torch.matmul (cuBLAS), AWQLinearMethod.apply of awq.py:434-451,
        # same CUDA-graph protocol, reference-layout tensors of the same shapes
        try:
            from oracle import build_ref

            ref_fn = build_ref.load()
        except Exception:
            ref_fn = None
        if ref_fn is not None:
            ref_gpu = {}
            gen = torch.Generator(device=device).manual_seed(99)
            for name, K, N in shapes:
                sets = []
                for _ in range(max(2, min(8, (600 << 20) // (K * N // 2)))):
                    qw = torch.randint(-(2**31), 2**31 - 1, (K, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                    qz = torch.randint(-(2**31), 2**31 - 1, (K // GROUP, N // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
                    sc = (torch.rand(K // GROUP, N, generator=gen, device=device) * 0.02 + 0.005).to(dtype)
                    sets.append((qw, sc, qz))
                for M in (1, 16):
                    x = xs[(M, name)]
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.stream(stream):
                        for qw, sc, qz in sets:
                            torch.matmul(x, ref_fn(qw, sc, qz))
                        torch.cuda.synchronize()
                        with torch.cuda.graph(g, stream=stream):
                            for qw, sc, qz in sets:
                                torch.matmul(x, ref_fn(qw, sc, qz))
                    t = time_calls(g.replay, 5) / len(sets)
                    ours_us = per_shape[f"{name}_M{M}"]["us"]
                    ref_gpu[f"{name}_M{M}"] = {"us": round(t * 1e6, 2), "speedup_ours": round(t * 1e6 / ours_us, 2)}
                del sets
                torch.cuda.empty_cache()
            extras["reference_cuda_path_same_gpu"] = ref_gpu
        tfs = [v["tflops"] for k, v in per_shape.items() if k.endswith("_M4096")]
        extras["prefill_m4096_tflops_mean"] = round(sum(tfs) / len(tfs), 1)
        extras["prefill_frac_of_measured_bf16_peak"] = round(sum(tfs) / len(tfs) / tf_peak, 3)
        # CPU baseline: the reference's CPU path (torch port) on this box's host cores, bounded sample
        cores = len(os.sched_getaffinity(0))
        t_cpu, b_cpu = 0.0, 0
        n_pass = 0
        t_start = time.perf_counter()
        while time.perf_counter() - t_start < 12.0 and n_pass < 6:
            t, b = cpu_reference_pass(model, DECODE_MS, None, cores, seed=n_pass)
            t_cpu += t
            b_cpu += b
            n_pass += 1
        cpu_baseline = {"value": b_cpu / t_cpu / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
                        "sample": f"{n_pass} decoder layer(s) x (qkv,o,gate_up,down) x M=1,4,16 via torch dequant+matmul "
                                  f"(oracle/torch_port.py), {t_cpu:.1f} s"}

    line = {
        "metric": "awq_w4a16_decode_hbm_gbs", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {
            "workload": f"{args.workload} AWQ: all {model['layers'] * 4} linears x decode M={','.join(map(str, DECODE_MS))} per step"
                        + (f", tensor-parallel TP={world} (NCCL all-reduce after o_proj/down_proj)" if world > 1 else ""),
            "group_size": GROUP, "launches_per_step": launches, "cuda_graph": graph is not None,
            "l2": f"distinct weights per step {step_bytes(model, (1,)) / 1e9:.2f} GB >> 126 MB L2 (no flush needed)",
            "pdl": True,
            "all_reduce": (None if world == 1 else
                           ("peer-memory one-shot kernel (awq_b200_allreduce)" if ar_launches else "ncclAllReduce")),
        },
        "roofline": {"bound": "hbm", "achieved": dom_gbs, "peak": hbm_peak, "unit": "GB/s",
                     "frac": dom_gbs / hbm_peak, "traffic": DOMINANT_KERNEL_DRAM_BYTES_PER_LAUNCH if world == 1 else None,
                     "peak_source": peak_src,
                     "kernel": f"decode kernel on gate_up_proj (K={dom_K}, N={dom_N_local}"
                               + ("): awq::gemv_persist_kernel" if dom_N_local >= 9600 else "; per-rank shard")
                               + f", mean over its M={','.join(map(str, DECODE_MS))} launches",
                     "algorithmic_bytes_per_launch": dom["bytes"] / dom["launches"],
                     "us_per_launch": dom["sec"] / dom["launches"] * 1e6, "per_m": dom["per_m"],
                     "whole_step": {"achieved": value / world, "frac": value / world / hbm_peak,
                                    "note": "all 384 launches of the step incl. the latency-bound small matrices"}},
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": int(x_host.numel() * 2),
                "d2h_bytes_per_step": int(y_host.numel() * 2)},
        # our kernels inside the timed region, per rank: one fused linear per layer call, plus one
        # peer-memory all-reduce per row-parallel call when N > 1 (NCCL launches are not counted)
        "gpu_launches": (launches + ar_launches) * args.steps,
        "clocks": clk,
        "decode_tokens_per_s_linears_only": tok_s,
    }
    if cpu_baseline is not None:
        line["cpu_baseline"] = cpu_baseline
    if extras:
        line["extras"] = extras
    print(json.dumps(line), flush=True)
    finish()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(MODELS), default="llama3-8b")
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
