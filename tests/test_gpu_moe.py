"""-m gpu: AWQ mixture-of-experts path against the CPU oracle.

Reference: AWQMoEMethod (python/sglang/srt/layers/quantization/awq.py:661-852) -> fused_marlin_moe: for every token,
out = sum_k w_k * W2_e(silu(W13gate_e x) * W13up_e x) over the router's top-k experts.  The oracle evaluates every
(token, k) pair with oracle.linear_f32 (double accumulation over the reference-dequantised expert weights), rounding
to fp16 where the reference's kernels round (after each linear, after the activation, after the weighted row, after
the top-k sum).  Bar: relative Frobenius error <= 1e-3."""
import numpy as np
import pytest
import torch

from gpu_util import from_bits, rel_fro, to_bits

from sglang_awq_b200.srt.layers.moe import FusedMoE
from sglang_awq_b200.srt.layers.quantization import AWQConfig, AWQMarlinConfig, AWQMoEMethod

pytestmark = pytest.mark.gpu


def _cfg():
    return AWQConfig.from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True})


def _checkpoint(oracle, E, hidden, inter, seed):
    """per expert: w1 (gate), w3 (up): [hidden, inter]; w2: [inter, hidden] in the reference AWQ layout"""
    ck = []
    for e in range(E):
        ck.append({n: oracle.make_inputs(K, N, 128, 0, seed=seed + 10 * e + i, signed_scales=True)
                   for i, (n, K, N) in enumerate((("w1", hidden, inter), ("w3", hidden, inter), ("w2", inter, hidden)))})
        for n in ("w1", "w3"):  # outputs of O(1) so that silu is exercised off its linear range
            s = oracle.bits_to_f32(ck[e][n]["scales"], 0) * np.float32(0.6 / np.sqrt(hidden) / 0.015)
            ck[e][n]["scales"] = oracle.f32_to_bits(s, 0)
        s = oracle.bits_to_f32(ck[e]["w2"]["scales"], 0) * np.float32(1.0 / np.sqrt(inter) / 0.09)
        ck[e]["w2"]["scales"] = oracle.f32_to_bits(s, 0)
    return ck


def _load(layer, ck):
    for e, ex in enumerate(ck):
        for shard, t in ex.items():
            for pname, key in (("qweight", "qweight"), ("qzeros", "qzeros"), ("scales", "scales")):
                prefix = "w2" if shard == "w2" else "w13"
                param = getattr(layer, f"{prefix}_{pname}")
                src = torch.from_numpy(t[key]) if key != "scales" else from_bits(t[key], 0, "cpu")
                param.weight_loader(param, src, f"experts.{e}.{shard}.{pname}", shard, e)


def _oracle_moe(oracle, ck, x16, ids, wts, tp=1, rank=0):
    M, topk = ids.shape
    hidden = x16.shape[1]
    out = np.zeros((M, hidden), dtype=np.float32)
    for t in range(M):
        for k in range(topk):
            e = int(ids[t, k])
            if e < 0:
                continue
            ex = ck[e]
            inter = ex["w1"]["qweight"].shape[1] * 8
            ipp = inter // tp
            sl, slp = slice(rank * ipp, (rank + 1) * ipp), slice(rank * ipp // 8, (rank + 1) * ipp // 8)
            xb = x16[t:t + 1].view(np.uint16)
            g = oracle.linear_f32(xb, ex["w1"]["qweight"][:, slp].copy(), ex["w1"]["scales"][:, sl].copy(), ex["w1"]["qzeros"][:, slp].copy(), 0)
            u = oracle.linear_f32(xb, ex["w3"]["qweight"][:, slp].copy(), ex["w3"]["scales"][:, sl].copy(), ex["w3"]["qzeros"][:, slp].copy(), 0)
            gu = np.concatenate([g, u], axis=1).astype(np.float16)
            hb, _ = oracle.silu_mul(gu.view(np.uint16), 0)
            rows = slice(rank * ipp, (rank + 1) * ipp)
            o = oracle.linear_f32(hb, ex["w2"]["qweight"][rows].copy(), ex["w2"]["scales"][rank * ipp // 128:(rank + 1) * ipp // 128].copy(),
                                  ex["w2"]["qzeros"][rank * ipp // 128:(rank + 1) * ipp // 128].copy(), 0)
            out[t] += (np.float32(wts[t, k]) * o[0]).astype(np.float16).astype(np.float32)
    return out.astype(np.float16)


def _routing(M, E, topk, seed, device="cuda"):
    g = torch.Generator(device=device).manual_seed(seed)
    logits = torch.randn(M, E, generator=g, device=device)
    w, ids = torch.topk(torch.softmax(logits, -1), topk, dim=-1)
    return (w / w.sum(-1, keepdim=True)).float(), ids.to(torch.int32)


# M <= 64: streaming kernel (device-side routing); above: grouped through the dense kernels
@pytest.mark.parametrize("M,E,topk", [(1, 8, 2), (4, 8, 2), (16, 8, 2), (3, 16, 4), (33, 4, 2), (64, 8, 2), (200, 8, 2), (70, 4, 2)])
def test_awq_moe_matches_oracle(cabi, oracle, M, E, topk):
    hidden, inter = 512, 768
    ck = _checkpoint(oracle, E, hidden, inter, seed=M + E)
    layer = FusedMoE(E, hidden, inter, topk, torch.float16, _cfg(), tp_rank=0, tp_size=1)
    assert isinstance(layer.quant_method, AWQMoEMethod)
    _load(layer, ck)
    layer.cuda()
    layer.quant_method.process_weights_after_loading(layer)
    layer.quant_method.process_weights_after_loading(layer)  # idempotent
    assert layer.w13_qweight.shape == (E, (2 * inter // 128) * (hidden // 128), 2144)
    g = torch.Generator(device="cuda").manual_seed(M)
    x = (torch.randn(M, hidden, generator=g, device="cuda") * 0.5).half()
    w, ids = _routing(M, E, topk, seed=M + 1)
    with torch.no_grad():
        out = layer(x, w, ids)
        out2 = layer(x, w, ids)
    assert out.shape == (M, hidden) and torch.equal(out, out2)
    ref = _oracle_moe(oracle, ck, to_bits(x).view(np.float16), ids.cpu().numpy(), w.cpu().numpy())
    assert oracle.rel_fro(out.float().cpu().numpy(), ref.astype(np.float32)) <= 1e-3


def test_awq_moe_streaming_equals_grouped_and_masks_experts(cabi, oracle):
    """the two routes agree (the grouped one rounds the unweighted row once more), ids of -1 contribute nothing
    (expert-parallel masks), and the streaming route is capturable in a CUDA graph"""
    E, topk, hidden, inter, M = 8, 2, 512, 512, 24
    ck = _checkpoint(oracle, E, hidden, inter, seed=5)
    layer = FusedMoE(E, hidden, inter, topk, torch.float16, AWQMarlinConfig.from_config(
        {"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True}), tp_rank=0, tp_size=1)
    assert isinstance(layer.quant_method, AWQMoEMethod)
    _load(layer, ck)
    layer.cuda()
    layer.quant_method.process_weights_after_loading(layer)
    x = (torch.randn(M, hidden, device="cuda") * 0.5).half()
    w, ids = _routing(M, E, topk, seed=2)
    ids[::3, 1] = -1
    qm = layer.quant_method
    with torch.no_grad():
        a = qm.apply(layer, x, w, ids)                   # streaming (24 tokens)
        b = _grouped(qm, layer, x, w, ids)             # grouped (96 tokens)
    assert rel_fro(a, b.float()) <= 1e-3
    ref = _oracle_moe(oracle, ck, to_bits(x).view(np.float16), ids.cpu().numpy(), w.cpu().numpy())
    assert oracle.rel_fro(a.float().cpu().numpy(), ref.astype(np.float32)) <= 1e-3
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.no_grad():
        qm.apply(layer, x, w, ids)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            c = qm.apply(layer, x, w, ids)
        gr.replay()
        torch.cuda.synchronize()
    assert torch.equal(c, a)


def _grouped(qm, layer, x, w, ids):
    """force the grouped route: a batch that is large for the streaming gate, made of the same rows repeated"""
    reps = 4
    out = qm.apply(layer, x.repeat(reps, 1), w.repeat(reps, 1), ids.repeat(reps, 1))
    return out[:x.shape[0]]


def test_awq_moe_tensor_parallel_partials_sum_to_unsharded(cabi, oracle):
    """w13 is column-split, w2 row-split (reference FusedMoE with is_transposed AWQ tensors): the ranks' partial outputs
    add up to the unsharded result"""
    E, topk, hidden, inter, M = 4, 2, 512, 1024, 6
    ck = _checkpoint(oracle, E, hidden, inter, seed=11)
    x = (torch.randn(M, hidden, device="cuda") * 0.5).half()
    w, ids = _routing(M, E, topk, seed=4)
    parts = []
    for r in range(2):
        layer = FusedMoE(E, hidden, inter, topk, torch.float16, _cfg(), reduce_results=False, tp_rank=r, tp_size=2)
        _load(layer, ck)
        layer.cuda()
        layer.quant_method.process_weights_after_loading(layer)
        with torch.no_grad():
            parts.append(layer(x, w, ids).float())
        ref_r = _oracle_moe(oracle, ck, to_bits(x).view(np.float16), ids.cpu().numpy(), w.cpu().numpy(), tp=2, rank=r)
        assert oracle.rel_fro(parts[-1].cpu().numpy(), ref_r.astype(np.float32)) <= 1e-3
    full = _oracle_moe(oracle, ck, to_bits(x).view(np.float16), ids.cpu().numpy(), w.cpu().numpy())
    assert oracle.rel_fro((parts[0] + parts[1]).cpu().numpy(), full.astype(np.float32)) <= 2e-3


def test_moe_linear_argument_checks(cabi):
    L = cabi.lib()
    x = torch.zeros(2, 256, dtype=torch.float16, device="cuda")
    t = torch.zeros(4 * 2144 * 4, dtype=torch.int32, device="cuda")
    ids = torch.zeros(2, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.awq_b200_moe_linear(x.data_ptr(), t.data_ptr(), ids.data_ptr(), None, x.data_ptr(), 2, 1, 250, 256, 4, 0, 0, st) != 0
    assert L.awq_b200_moe_linear(x.data_ptr(), t.data_ptr(), None, None, x.data_ptr(), 2, 1, 256, 256, 4, 0, 0, st) != 0
    assert L.awq_b200_moe_linear(x.data_ptr(), t.data_ptr(), ids.data_ptr(), None, x.data_ptr(), 2, 1, 256, 24, 4, 1, 0, st) != 0
    assert L.awq_b200_moe_linear(x.data_ptr(), t.data_ptr(), ids.data_ptr(), None, x.data_ptr(), 0, 1, 256, 256, 4, 0, 0, st) == 0
    assert L.awq_b200_moe_sum(x.data_ptr(), x.data_ptr(), 1, 2, 250, 0, st) != 0


# ---- long-K expert matrices: moe_gemv_chunk_kernel (x staged per chunk of the K slice, 8 pairs per pass) ----
def _expert_stack(cabi, E, K, N, gate_up, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    tiles, dense = [], []
    for _ in range(E):
        qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
        sign = (torch.rand(K // 128, N, generator=g, device="cuda") > 0.3).half() * 2 - 1
        sc = ((torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005) * (1.0 / (K ** 0.5) / 0.06)).half() * sign
        tiles.append((cabi.repack_gate_up if gate_up else cabi.repack)(qw, sc, qz))
        dense.append(cabi.dequantize(qw, sc, qz).float())  # bit-exact with the reference op (test_gpu_dequant)
    return torch.stack(tiles), dense


@pytest.mark.parametrize("K,N,silu", [(14336, 4096, False), (14336, 264, False), (9088, 1024, False), (7168, 4096, True),
                                      (7168, 272, True)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_moe_linear_long_k_chunked_kernel(cabi, K, N, silu, dtype):
    L = cabi.lib()
    E = 4
    tiles, dense = _expert_stack(cabi, E, K, N, silu, seed=K + N)
    if dtype == torch.bfloat16:  # same packed integers, bf16 scales
        g = torch.Generator(device="cuda").manual_seed(K + N)
        tiles, dense = [], []
        for _ in range(E):
            qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
            qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
            sc = ((torch.rand(K // 128, N, generator=g, device="cuda") * 0.02 + 0.005) * (1.0 / (K ** 0.5) / 0.06)).bfloat16()
            tiles.append((cabi.repack_gate_up if silu else cabi.repack)(qw, sc, qz))
            dense.append(cabi.dequantize(qw, sc, qz).float())
        tiles = torch.stack(tiles)
    for M, topk in ((3, 2), (16, 2), (32, 2), (9, 4)):
        g = torch.Generator(device="cuda").manual_seed(M)
        x = (torch.randn(M, K, generator=g, device="cuda") * 0.5).to(dtype)
        ids = torch.stack([torch.randperm(E, generator=g, device="cuda")[:topk] for _ in range(M)]).reshape(-1).to(torch.int32)
        if M == 16:
            ids[5] = -1      # a pair routed to another rank: its row is zero
            ids[::3] = 2     # a crowded expert: more than 8 pairs -> several rounds
        scale = None if silu else torch.rand(M * topk, generator=g, device="cuda")
        ref = torch.zeros((M * topk, N // 2 if silu else N), device="cuda")
        for p in range(M * topk):
            e = int(ids[p])
            if e < 0:
                continue
            full = x[p // topk].float() @ dense[e]
            if silu:
                gate, up = full[: N // 2].to(dtype).float(), full[N // 2:].to(dtype).float()
                ref[p] = torch.nn.functional.silu(gate) * up
            else:
                ref[p] = full * scale[p]
        try:
            L.awq_b200_debug_set_moe_chunk(2)   # the chunked kernel whatever the pair count
            y = cabi.moe_linear(x, tiles, ids, N, x_div=topk, row_scale=scale, silu=silu)
            y2 = cabi.moe_linear(x, tiles, ids, N, x_div=topk, row_scale=scale, silu=silu)
            L.awq_b200_debug_set_moe_chunk(0)
            y_old = cabi.moe_linear(x, tiles, ids, N, x_div=topk, row_scale=scale, silu=silu)
        finally:
            L.awq_b200_debug_set_moe_chunk(1)
        tol = 1e-3 if dtype == torch.float16 else 4e-3   # bf16 outputs: 8-bit significand
        assert rel_fro(y, ref) <= tol, f"K={K} N={N} M={M} topk={topk}: {rel_fro(y, ref):.2e}"
        assert torch.equal(y.view(torch.int16), y2.view(torch.int16)), "deterministic"
        assert rel_fro(y, y_old.float()) <= tol, "chunked vs whole-K kernel"
        if M == 16:
            assert not y[5].any()
