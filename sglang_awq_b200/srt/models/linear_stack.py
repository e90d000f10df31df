"""The AWQ linear stack of a decoder-only model: every quantized linear of every layer, chained.

This is the slice of the reference's model files that lies on the AWQ path
(python/sglang/srt/models/llama.py:94-106 LlamaMLP, :143-178 attention projections,
python/sglang/srt/models/qwen2.py:143-148 qkv bias): per layer
    qkv_proj (QKVParallelLinear, + bias for Qwen2) -> o_proj (RowParallelLinear, all-reduce)
    gate_up_proj (MergedColumnParallelLinear) -> SiluAndMul -> down_proj (RowParallelLinear, all-reduce)
Attention, rotary embeddings, KV cache, norms with weights, embeddings and the LM head are NOT
part of the path (out of scope).  So that the four linears can be chained and
their last output checked end to end ("last-layer logits on dummy-weight runs", BASELINE.json),
the glue between them is the simplest shape-preserving stand-in:
    attention      -> the rank's q slice of qkv, unchanged          [M, q_size / tp]
    residual+norm  -> h = rms_normalise(h + branch), unit weight    (csrc/awq_allreduce.cu residual_rmsnorm, fp32 math)
    SiluAndMul     -> silu(gate) * up                               (fused epilogue, or csrc silu_and_mul, fp32 math)
Used by tests/test_gpu_stack.py, sglang_awq_b200/bench_one_batch.py and bench.py.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from ..layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
from ..layers.quantization.base_config import QuantizationConfig

MODEL_SHAPES: Dict[str, Dict[str, int]] = {
    # name: hidden, intermediate, q heads, kv heads, head dim, layers, qkv bias (HF configs of the BASELINE.json models)
    "llama3-8b": dict(hidden=4096, inter=14336, heads=32, kv_heads=8, head_dim=128, layers=32, qkv_bias=0),
    "llama3-70b": dict(hidden=8192, inter=28672, heads=64, kv_heads=8, head_dim=128, layers=80, qkv_bias=0),
    "qwen2.5-32b": dict(hidden=5120, inter=27648, heads=40, kv_heads=8, head_dim=128, layers=64, qkv_bias=1),
    "llama2-7b-mlp": dict(hidden=4096, inter=11008, heads=32, kv_heads=32, head_dim=128, layers=32, qkv_bias=0),
}


def linear_shapes(m: Dict[str, int]) -> List[Tuple[str, int, int]]:
    """(name, K, N) of the four AWQ linears of one decoder layer, unsharded."""
    q = m["heads"] * m["head_dim"]
    kv = m["kv_heads"] * m["head_dim"]
    return [("qkv_proj", m["hidden"], q + 2 * kv), ("o_proj", q, m["hidden"]),
            ("gate_up_proj", m["hidden"], 2 * m["inter"]), ("down_proj", m["inter"], m["hidden"])]


_ONES: Dict[Tuple[torch.device, torch.dtype, int], torch.Tensor] = {}


def add_rms_normalise(h: torch.Tensor, branch: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """rms_normalise(h + branch) with unit weight, through the residual + RMSNorm kernel (fp32 sum, fp32 statistics,
    one rounding): the reference's fused_add_rmsnorm with weight = 1.  Returns the normalised tensor (written over
    `branch`); `h` receives the rounded sum like the reference's residual stream."""
    from sglang_awq_b200 import sgl_kernel

    key = (h.device, h.dtype, h.shape[-1])
    if key not in _ONES:
        _ONES[key] = torch.ones(h.shape[-1], dtype=h.dtype, device=h.device)
    branch = branch.contiguous()
    sgl_kernel.fused_add_rmsnorm(branch.view(-1, h.shape[-1]), h.contiguous().view(-1, h.shape[-1]), _ONES[key], eps)
    return branch


def silu_and_mul(gate_up: torch.Tensor) -> torch.Tensor:
    """reference python/sglang/srt/layers/activation.py SiluAndMul.forward_cuda -> sgl_kernel.silu_and_mul"""
    from sglang_awq_b200 import sgl_kernel

    return sgl_kernel.silu_and_mul(gate_up.contiguous())


class AWQDecoderLinears(torch.nn.Module):
    def __init__(self, m: Dict[str, int], quant_config: QuantizationConfig, params_dtype: torch.dtype, prefix: str,
                 tp_rank: Optional[int], tp_size: Optional[int], fuse_silu_mul: bool = False):
        super().__init__()
        q = m["heads"] * m["head_dim"]
        self.qkv_proj = QKVParallelLinear(m["hidden"], m["head_dim"], m["heads"], m["kv_heads"],
                                          bias=bool(m.get("qkv_bias", 0)), quant_config=quant_config,
                                          params_dtype=params_dtype, prefix=prefix + "self_attn.qkv_proj",
                                          tp_rank=tp_rank, tp_size=tp_size)
        self.o_proj = RowParallelLinear(q, m["hidden"], bias=False, quant_config=quant_config, params_dtype=params_dtype,
                                        prefix=prefix + "self_attn.o_proj", tp_rank=tp_rank, tp_size=tp_size)
        self.gate_up_proj = MergedColumnParallelLinear(m["hidden"], [m["inter"]] * 2, bias=False,
                                                       quant_config=quant_config, params_dtype=params_dtype,
                                                       prefix=prefix + "mlp.gate_up_proj", tp_rank=tp_rank,
                                                       tp_size=tp_size)
        self.down_proj = RowParallelLinear(m["inter"], m["hidden"], bias=False, quant_config=quant_config,
                                           params_dtype=params_dtype, prefix=prefix + "mlp.down_proj", tp_rank=tp_rank,
                                           tp_size=tp_size)
        self.q_size_per_rank = q // self.o_proj.tp_size
        # read by AWQLinearMethod.process_weights_after_loading: repack gate / up rows interleaved so that
        # SiluAndMul runs in the linear's epilogue (one launch, no [M, 2 * inter] round trip)
        self.gate_up_proj.awq_fuse_silu_mul = bool(fuse_silu_mul)
        self.fuse_silu_mul = bool(fuse_silu_mul)

    def mlp_act(self, h: torch.Tensor) -> torch.Tensor:
        """gate_up_proj -> SiluAndMul (reference llama.py:94-106), fused when the layer was built for it"""
        if self.fuse_silu_mul:
            return self.gate_up_proj.forward_silu_mul(h)
        return silu_and_mul(self.gate_up_proj(h)[0])


class AWQLinearStack(torch.nn.Module):
    """`layers` decoder layers' worth of AWQ linears with the stand-in glue described above."""

    def __init__(self, model: Dict[str, int], quant_config: QuantizationConfig, num_layers: Optional[int] = None,
                 params_dtype: torch.dtype = torch.float16, tp_rank: Optional[int] = None,
                 tp_size: Optional[int] = None, fuse_silu_mul: bool = False):
        super().__init__()
        self.model = dict(model)
        n = model["layers"] if num_layers is None else num_layers
        self.layers = torch.nn.ModuleList(
            AWQDecoderLinears(model, quant_config, params_dtype, f"model.layers.{i}.", tp_rank, tp_size, fuse_silu_mul)
            for i in range(n))

    def forward(self, h: torch.Tensor) -> torch.Tensor:
        """h: [M, hidden] (replicated on every TP rank) -> last hidden state [M, hidden]."""
        h = h.clone()  # the residual kernel updates its residual operand in place
        for lyr in self.layers:
            qkv, _ = lyr.qkv_proj(h)
            o, _ = lyr.o_proj(qkv[..., :lyr.q_size_per_rank].contiguous())
            h = add_rms_normalise(h, o)
            d, _ = lyr.down_proj(lyr.mlp_act(h))
            h = add_rms_normalise(h, d)
        return h

    def linears(self):
        for lyr in self.layers:
            yield from (("qkv_proj", lyr.qkv_proj), ("o_proj", lyr.o_proj), ("gate_up_proj", lyr.gate_up_proj),
                        ("down_proj", lyr.down_proj))


def _sum_ranks(parts: List[torch.Tensor]) -> torch.Tensor:
    """What the all-reduce computes: fp32 sum in rank order, one rounding (csrc/awq_allreduce.cu)."""
    acc = parts[0].float()
    for p in parts[1:]:
        acc = acc + p.float()
    return acc.to(parts[0].dtype)


def forward_tp_emulated(stacks: List[AWQLinearStack], h: torch.Tensor) -> torch.Tensor:
    """All TP ranks' stacks on ONE device, in lockstep: column-parallel layers run per rank, the
    row-parallel partial sums are added in rank order where the all-reduce would run."""
    h = h.clone()  # the residual kernel updates its residual operand in place
    for lyrs in zip(*(s.layers for s in stacks)):
        o = _sum_ranks([l.o_proj(l.qkv_proj(h)[0][..., :l.q_size_per_rank].contiguous(), skip_all_reduce=True)[0]
                        for l in lyrs])
        h = add_rms_normalise(h, o)
        d = _sum_ranks([l.down_proj(l.mlp_act(h), skip_all_reduce=True)[0] for l in lyrs])
        h = add_rms_normalise(h, d)
    return h
