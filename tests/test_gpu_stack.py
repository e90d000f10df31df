"""-m gpu: dummy-weight full-stack parity (BASELINE.json configs[3] / north_star "last-layer logits
checked to the same tolerance on dummy-weight runs").

Per-rank shapes of Qwen2.5-32B (qkv bias) and Llama-3-70B at tp in {1, 2, 4}: the AWQ linears of
L decoder layers are chained qkv(+bias) -> o -> gate_up -> silu*mul -> down through the public
layer API with the reference's dummy-loader ORDER (python/sglang/srt/model_loader/loader.py:1211-1218:
process_weights_after_loading first, then initialize_dummy_weights,
python/sglang/srt/model_loader/weight_utils.py:1108-1137, overwrites every float parameter in
place -- here the strided scale views inside the tile records); the int32 parameters are seeded
before the hook.  All TP ranks run on this GPU in lockstep; the row-parallel partials are summed
in fp32 rank order (what the all-reduce computes).  The last hidden state is compared with the
CPU oracle run on the SAME parameters: oracle.linear_f32 (double accumulation over the
reference-dequantised weights) per linear, rounded to fp16 where the reference's torch ops round,
the same fp32 glue."""
import numpy as np
import pytest
import torch

from gpu_util import from_bits, to_bits

from sglang_awq_b200.srt.layers.quantization import AWQConfig
from sglang_awq_b200.srt.model_loader import initialize_dummy_weights, seed_integer_params
from sglang_awq_b200.srt.models import MODEL_SHAPES, AWQLinearStack, forward_tp_emulated

pytestmark = pytest.mark.gpu
LAYERS = 2


def _cfg():
    return AWQConfig.from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True})


def _ref_scales(layer):
    """[K/128, N] uint16 scale bits of a TILED layer, read back through its `scales` view."""
    K, N = layer.awq_in_features, layer.awq_out_features
    KT, NT = K // 128, (N + 127) // 128
    v = to_bits(layer.scales.data).reshape(NT, KT, 8, 16)
    e = [((j & 7) << 1) | (j >> 3) for j in range(16)]  # entry of row j inside a 16-row block (paired order)
    return np.ascontiguousarray(v[..., e].transpose(1, 0, 2, 3).reshape(KT, NT * 128)[:, :N])


def _build_rank(model, rank, tp, seed, fuse=False):
    stack = AWQLinearStack(model, _cfg(), num_layers=LAYERS, params_dtype=torch.float16, tp_rank=rank, tp_size=tp,
                           fuse_silu_mul=fuse).cuda()
    seed_integer_params(stack, seed + 1000 * rank)
    ints = [{"qweight": lin.qweight.data.cpu().numpy().copy(), "qzeros": lin.qzeros.data.cpu().numpy().copy()}
            for _, lin in stack.linears()]
    for _, module in stack.named_modules():   # loader.py:1211-1214
        qm = getattr(module, "quant_method", None)
        if qm is not None:
            qm.process_weights_after_loading(module)
    initialize_dummy_weights(stack, seed=seed)  # loader.py:1218
    params = []
    for (name, lin), iv in zip(stack.linears(), ints):
        assert lin.awq_layout == "tiled"
        assert bool(getattr(lin, "awq_gate_up", False)) == (fuse and name == "gate_up_proj")
        sc = _ref_scales(lin)
        if getattr(lin, "awq_gate_up", False):  # the views are in slab order: physical row n' = logical perm[n']
            from oracle import awq_oracle
            logical = np.empty_like(sc)
            logical[:, awq_oracle.gate_up_interleave_perm(lin.awq_out_features // 2)] = sc
            sc = logical
        params.append({"name": name, **iv, "scales": sc,
                       "bias": None if lin.bias is None else to_bits(lin.bias.data)})
    return stack, params


def _f16(v):
    return v.astype(np.float16)


def _rms(h):
    f = h.astype(np.float32)
    return _f16(f / np.sqrt((f * f).mean(-1, keepdims=True, dtype=np.float32) + np.float32(1e-6)))


def _oracle_linear(oracle, x16, p, bias_ok=True):
    y = oracle.linear_f32(x16.view(np.uint16), p["qweight"], p["scales"], p["qzeros"], 0,
                          p["bias"] if bias_ok else None)
    return _f16(y)


def _oracle_stack(oracle, ranks, h16, q_per_rank):
    tp = len(ranks)
    for li in range(LAYERS):
        parts = []
        for r in range(tp):
            qkv_p, o_p = ranks[r][4 * li], ranks[r][4 * li + 1]
            qkv = _oracle_linear(oracle, h16, qkv_p)
            parts.append(_oracle_linear(oracle, np.ascontiguousarray(qkv[:, :q_per_rank]), o_p).astype(np.float32))
        h16 = _rms(h16.astype(np.float32) + _f16(sum(parts[1:], parts[0])).astype(np.float32))
        parts = []
        for r in range(tp):
            gu_p, d_p = ranks[r][4 * li + 2], ranks[r][4 * li + 3]
            gu = _oracle_linear(oracle, h16, gu_p).astype(np.float32)
            d = gu.shape[1] // 2
            g, u = gu[:, :d], gu[:, d:]
            act = _f16(g / (np.float32(1) + np.exp(-g)) * u)
            parts.append(_oracle_linear(oracle, act, d_p).astype(np.float32))
        h16 = _rms(h16.astype(np.float32) + _f16(sum(parts[1:], parts[0])).astype(np.float32))
    return h16


# one M per (model, tp): decode kernels of every family (M = 1, 4/5, 8, 16), the mid range and the
# tcgen05 GEMM path (M = 96, 300)
CASES = [("qwen2.5-32b", 1, 1), ("qwen2.5-32b", 2, 16), ("qwen2.5-32b", 4, 5), ("qwen2.5-32b", 4, 96),
         ("llama3-70b", 1, 16), ("llama3-70b", 2, 1), ("llama3-70b", 4, 8), ("llama3-70b", 4, 300)]


FUSED_CASES = [("qwen2.5-32b", 2, 1), ("qwen2.5-32b", 4, 16), ("llama3-70b", 4, 4), ("llama3-70b", 2, 200)]


@pytest.mark.parametrize("name,tp,M,fuse", [c + (False,) for c in CASES] + [c + (True,) for c in FUSED_CASES],
                         ids=[f"{n}-tp{t}-M{m}" for n, t, m in CASES] + [f"{n}-tp{t}-M{m}-fused_silu" for n, t, m in FUSED_CASES])
def test_dummy_weight_stack_matches_oracle(cabi, oracle, name, tp, M, fuse):
    model = MODEL_SHAPES[name]
    built = [_build_rank(model, r, tp, seed=1234, fuse=fuse) for r in range(tp)]
    stacks = [b[0] for b in built]
    g = torch.Generator(device="cuda").manual_seed(M + tp)
    h0 = torch.randn(M, model["hidden"], generator=g, device="cuda").half()
    with torch.no_grad():
        h = forward_tp_emulated(stacks, h0)
    torch.cuda.synchronize()
    assert h.shape == (M, model["hidden"]) and torch.isfinite(h).all()
    q_per_rank = model["heads"] * model["head_dim"] // tp
    ref = _oracle_stack(oracle, [b[1] for b in built], to_bits(h0).view(np.float16), q_per_rank)
    err = oracle.rel_fro(h.float().cpu().numpy(), ref.astype(np.float32))
    assert err <= 1e-3, f"{name} tp={tp} M={M}: last hidden state off by {err:.2e}"
    if model.get("qkv_bias"):
        assert stacks[0].layers[0].qkv_proj.bias is not None  # Qwen2 qkv bias is on the path (qwen2.py:143-148)


def test_dummy_init_is_rank_invariant_and_lands_in_the_records(cabi):
    """initialize_dummy_weights seeds one generator per parameter, so equal-shaped shards get equal
    values on every rank (weight_utils.py:1121-1126), and writing through the strided `scales`
    view must change the dequantised weights (i.e. reach the bytes the kernels read)."""
    from sglang_awq_b200.srt.layers.quantization import AWQLinearMethod

    model = dict(MODEL_SHAPES["llama3-8b"], hidden=1024, inter=2048, heads=8, kv_heads=2)
    a, _ = _build_rank(model, 0, 2, seed=7)
    b, _ = _build_rank(model, 1, 2, seed=7)
    la, lb = a.layers[0].down_proj, b.layers[0].down_proj
    assert torch.equal(la.scales.data, lb.scales.data)
    assert not torch.equal(la.qweight.data, lb.qweight.data)  # integer words are seeded per rank
    W = AWQLinearMethod.dequantize(la)
    assert float(W.abs().max()) <= 15e-3 + 1e-6 and float(W.abs().max()) > 1e-3
