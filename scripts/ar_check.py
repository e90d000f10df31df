"""Multi-GPU check of the one-shot peer-memory all-reduce (run under torchrun on N GPUs of one node):
bit-exactness against the fp32 rank-order sum, identical results on every rank, slot rotation
over many back-to-back calls, CUDA-graph replay, and latency next to ncclAllReduce.
Prints one JSON line on rank 0."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from sglang_awq_b200.srt import distributed as D  # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    D.init_distributed_environment(world_size=world, rank=rank, local_rank=local, backend="nccl")
    D.initialize_model_parallel(world)
    tp = D.get_tp_group()
    dev = torch.device("cuda", local)
    out = {"world": world, "peer_path": tp.peer_ar is not None, "cases": 0, "bad": []}
    if tp.peer_ar is None:
        if rank == 0:
            print(json.dumps(out))
        os._exit(0)
    ar = tp.peer_ar
    from sglang_awq_b200 import _C
    L = _C.lib()
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    def expect(x):
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x)
        s = torch.zeros_like(x, dtype=torch.float32)
        for p in parts:  # rank order, fp32: what the kernel does
            s += p.float()
        return s.to(x.dtype)

    for dtype in (torch.float16, torch.bfloat16):
        for numel in (8, 4096, 8192, 4096 * 16 + 8, 8192 * 16, 8192 * 64):
            for rep in range(8):
                L.awq_b200_debug_set_allreduce_packets((1, 0, 3, 2, 1, 3, 0, 3)[rep])  # every kernel, interleaved
                x = torch.randn(numel, generator=g, device=dev).to(dtype)
                want = expect(x)
                y = x.clone()
                got = tp.all_reduce(y)
                same = [torch.empty_like(got) for _ in range(world)]
                dist.all_gather(same, got)
                out["cases"] += 1
                if not torch.equal(got.view(torch.int16), want.view(torch.int16)):
                    out["bad"].append(["value", str(dtype), numel, rep])
                if any(not torch.equal(s.view(torch.int16), got.view(torch.int16)) for s in same):
                    out["bad"].append(["ranks differ", str(dtype), numel, rep])
    L.awq_b200_debug_set_allreduce_packets(1)
    # fused all-reduce + residual add + RMSNorm (awq_b200_allreduce_residual_rmsnorm) against the two kernels back to
    # back (bit-identical by construction) and against the fp32 torch restatement of RMSNorm.forward_native
    from sglang_awq_b200 import _C as C
    for dtype in (torch.float16, torch.bfloat16):
        for rows, hidden in ((1, 8192), (4, 8192), (16, 8192), (3, 5120), (64, 4096), (2, 1024), (5, 1000), (40, 8192)):
            for mode in (1, 2, 0, 3):
                L.awq_b200_debug_set_allreduce_packets(mode)
                x = torch.randn(rows, hidden, generator=g, device=dev).to(dtype)
                res = torch.randn(rows, hidden, generator=g, device=dev).to(dtype)
                dist.broadcast(res, 0)  # the residual stream is replicated
                w = (torch.rand(hidden, generator=g, device=dev) + 0.5).to(dtype)
                dist.broadcast(w, 0)
                a = expect(x)
                want_res = (a.float() + res.float())
                want = (want_res * torch.rsqrt(want_res.pow(2).mean(-1, keepdim=True) + 1e-6) * w.float())
                x1, r1 = a.clone(), res.clone()
                C.residual_rmsnorm(x1, r1, w, 1e-6)  # unfused second half on the exact all-reduce result
                x2, r2 = x.clone(), res.clone()
                tp.all_reduce_residual_rmsnorm(x2, r2, w, 1e-6)
                out["cases"] += 1
                if not torch.equal(r2.view(torch.int16), want_res.to(dtype).view(torch.int16)):
                    out["bad"].append(["fused norm: residual", str(dtype), rows, hidden, mode])
                if not (torch.equal(x2.view(torch.int16), x1.view(torch.int16)) and torch.equal(r2.view(torch.int16), r1.view(torch.int16))):
                    out["bad"].append(["fused norm != all-reduce then norm", str(dtype), rows, hidden, mode])
                err = float((x2.float() - want).norm() / want.norm())
                if err > (1e-3 if dtype == torch.float16 else 4e-3):
                    out["bad"].append(["fused norm: value", str(dtype), rows, hidden, mode, err])
    L.awq_b200_debug_set_allreduce_packets(1)
    # row-parallel linear + all-reduce in one call (awq_b200_linear_allreduce: the decode kernel's epilogue pushes the
    # partial sums, the collective only polls) against linear -> all_reduce: bit-identical, also under graph replay and
    # interleaved with plain all-reduces (they share the per-block epochs)
    def rand_tiles(K, N):
        qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device=dev, dtype=torch.int64).to(torch.int32)
        qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device=dev, dtype=torch.int64).to(torch.int32)
        sc = (torch.rand(K // 128, N, generator=g, device=dev) * 0.02 + 0.005).half()
        return C.repack(qw, sc, qz)
    fused_cells = []
    for K, N in ((1024, 8192), (3584, 8192), (2048, 8192), (7168, 8192), (1024, 4096), (1792, 5120)):
        tiles = rand_tiles(K, N)
        bias = torch.randn(N, generator=g, device=dev).half() if rank == 0 else None
        for M in (1, 2, 4, 8, 16):
            x = torch.randn(M, K, generator=g, device=dev).half()
            want = C.linear(x, tiles, N, bias=bias)
            want = tp.all_reduce(want)
            got = ar.linear_all_reduce(x, tiles, bias, N)
            junk = torch.randn(4096, generator=g, device=dev).half()
            tp.all_reduce(junk)                                   # a plain collective in between
            got2 = ar.linear_all_reduce(x, tiles, bias, N)
            out["cases"] += 1
            fused_cells.append((K, N, M, int(L.awq_b200_gemv_route(M, K, N, 0))))
            if not (torch.equal(got.view(torch.int16), want.view(torch.int16)) and torch.equal(got2.view(torch.int16), want.view(torch.int16))):
                out["bad"].append(["linear_allreduce", K, N, M])
    out["linear_allreduce_routes"] = sorted(set(r for _, _, _, r in fused_cells))
    # the same in bf16 (its own route table, the kernels' bf16 instantiations of the packet push)
    bf_routes = set()
    for K, N in ((1024, 8192), (3584, 8192), (2048, 4096)):
        gq = torch.Generator(device=dev).manual_seed(1000 + rank + K)
        qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=gq, device=dev, dtype=torch.int64).to(torch.int32)
        qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=gq, device=dev, dtype=torch.int64).to(torch.int32)
        sc = (torch.rand(K // 128, N, generator=gq, device=dev) * 0.02 + 0.005).bfloat16()
        tiles_bf = C.repack(qw, sc, qz)
        for M in (1, 4, 8, 16):
            x = torch.randn(M, K, generator=g, device=dev).bfloat16()
            want = tp.all_reduce(C.linear(x, tiles_bf, N))
            got = ar.linear_all_reduce(x, tiles_bf, None, N)
            got2 = ar.linear_all_reduce(x, tiles_bf, None, N)
            out["cases"] += 1
            bf_routes.add(int(L.awq_b200_gemv_route(M, K, N, 1)))
            if not (torch.equal(got.view(torch.int16), want.view(torch.int16)) and torch.equal(got2.view(torch.int16), want.view(torch.int16))):
                out["bad"].append(["linear_allreduce bf16", K, N, M])
    out["linear_allreduce_routes_bf16"] = sorted(bf_routes)
    tiles = rand_tiles(3584, 8192)
    xg = [torch.randn(4, 3584, generator=g, device=dev).half() for _ in range(10)]
    wants = [tp.all_reduce(C.linear(xx, tiles, 8192)) for xx in xg]
    sg = torch.cuda.Stream()
    sg.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(sg):
        outs = [ar.linear_all_reduce(xx, tiles, None, 8192) for xx in xg]
        torch.cuda.synchronize()
        grf = torch.cuda.CUDAGraph()
        with torch.cuda.graph(grf, stream=sg):
            outs = [ar.linear_all_reduce(xx, tiles, None, 8192) for xx in xg]
        for _ in range(3):
            grf.replay()
        torch.cuda.synchronize()
    if any(not torch.equal(a.view(torch.int16), b.view(torch.int16)) for a, b in zip(outs, wants)):
        out["bad"].append(["linear_allreduce graph replay"])
    # 200 back-to-back calls on changing data (both slots, every epoch parity), out-of-place
    x = torch.randn(8192 * 4, generator=g, device=dev).half()
    acc = torch.zeros_like(x)
    ref = torch.zeros_like(x)
    for i in range(200):
        xi = (x * (1.0 + 0.01 * i)).half()
        ref.copy_(expect(xi))
        ar.all_reduce(xi, out=acc)
        if not torch.equal(acc.view(torch.int16), ref.view(torch.int16)):
            out["bad"].append(["sequence", i])
            break
    # CUDA graph: capture 20 calls, replay 5 times
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    bufs = [torch.randn(8192, generator=g, device=dev).half() for _ in range(20)]
    wants = [expect(b) for b in bufs]
    orig = [b.clone() for b in bufs]
    with torch.cuda.stream(s):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            for b in bufs:
                tp.all_reduce(b)
        for _ in range(5):
            for b, o in zip(bufs, orig):
                b.copy_(o)
            gr.replay()
        torch.cuda.synchronize()
    if any(not torch.equal(b.view(torch.int16), w.view(torch.int16)) for b, w in zip(bufs, wants)):
        out["bad"].append(["graph replay"])
    out["error_flag"] = ar.error_flag()

    # latency: 100 calls per graph, peer kernel vs NCCL
    def time_graph(fn, numel):
        t = torch.randn(numel, generator=g, device=dev).half()
        with torch.cuda.stream(s):
            fn(t)
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                for _ in range(100):
                    fn(t)
            gr.replay()
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            for _ in range(5):
                gr.replay()
            e1.record(s)
            torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / 500

    lat = {}
    for numel in (4096, 8192, 8192 * 2, 8192 * 4, 8192 * 8, 8192 * 16, 8192 * 32, 8192 * 64):
        rec = {"peer_us": round(time_graph(lambda t: ar.all_reduce(t), numel), 2)}
        if numel * 2 <= 512 * 1024:
            L.awq_b200_debug_set_allreduce_packets(2)
            rec["peer_packets_us"] = round(time_graph(lambda t: ar.all_reduce(t), numel), 2)
        if world >= 4 and numel % (8 * world) == 0:
            L.awq_b200_debug_set_allreduce_packets(3)
            rec["peer_twoshot_us"] = round(time_graph(lambda t: ar.all_reduce(t), numel), 2)
        L.awq_b200_debug_set_allreduce_packets(0)
        rec["peer_flag_exchange_us"] = round(time_graph(lambda t: ar.all_reduce(t), numel), 2)
        L.awq_b200_debug_set_allreduce_packets(1)
        rec["nccl_us"] = round(time_graph(lambda t: dist.all_reduce(t, group=tp.device_group), numel), 2)
        if numel % 8192 == 0:
            rows = numel // 8192
            res = torch.randn(rows, 8192, generator=g, device=dev).half()
            w = torch.ones(8192, device=dev).half()
            rec["fused_ar_residual_rmsnorm_us"] = round(time_graph(
                lambda t: tp.all_reduce_residual_rmsnorm(t.view(rows, 8192), res, w, 1e-6), numel), 2)
            rec["ar_then_residual_rmsnorm_us"] = round(time_graph(
                lambda t: (ar.all_reduce(t), C.residual_rmsnorm(t.view(rows, 8192), res, w, 1e-6)), numel), 2)
        lat[f"{numel * 2 // 1024}KB"] = rec
    out["latency"] = lat
    # fused vs unfused row-parallel linear, chains of 50 calls in a graph (Llama-3-70B per-rank o_proj / down_proj shards)
    fl = {}
    for name, K, N in (("o_proj", 8192 // world, 8192), ("down_proj", 28672 // world, 8192)):
        tiles = rand_tiles(K, N)
        for M in (1, 4, 16):
            x = torch.randn(M, K, generator=g, device=dev).half()
            rec = {}
            for label, on in (("fused_us", 1), ("unfused_us", 0), ("fused_us", 1), ("unfused_us", 0)):
                L.awq_b200_debug_set_linear_allreduce_fusion(on)
                with torch.cuda.stream(s):
                    ar.linear_all_reduce(x, tiles, None, N)
                    torch.cuda.synchronize()
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr, stream=s):
                        for _ in range(50):
                            ar.linear_all_reduce(x, tiles, None, N)
                    gr.replay()
                    torch.cuda.synchronize()
                    dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(s)
                    for _ in range(5):
                        gr.replay()
                    e1.record(s)
                    torch.cuda.synchronize()
                rec[label] = round(min(rec.get(label, 1e9), e0.elapsed_time(e1) * 1e3 / 250), 2)
            L.awq_b200_debug_set_linear_allreduce_fusion(1)
            rec["route"] = int(L.awq_b200_gemv_route(M, K, N, 0))
            fl[f"{name}_M{M}"] = rec
    out["linear_allreduce_latency"] = fl
    out["ok"] = not out["bad"] and out["error_flag"] == 0
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print(json.dumps(out), flush=True)
    os._exit(0)


if __name__ == "__main__":
    main()
