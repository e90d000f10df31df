"""Shared fixtures for the layer-level tests: a synthetic AWQ checkpoint for one decoder layer
(q/k/v/o/gate/up/down) and helpers that load it into the TP layers the way a model's
load_weights does (reference llama.py:560-568 stacked_params_mapping)."""
import numpy as np
import torch

from sglang_awq_b200.srt.layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
from sglang_awq_b200.srt.layers.quantization import AWQConfig


def awq_cfg(group=128):
    return AWQConfig.from_config({"w_bit": 4, "q_group_size": group, "zero_point": True})


def make_matrix(oracle, K, N, G, seed, dt=0):
    inp = oracle.make_inputs(K, N, G, dt, seed=seed)
    return {k: inp[k] for k in ("qweight", "qzeros", "scales")}


def to_torch(m, dt=0):
    tdt = torch.float16 if dt == 0 else torch.bfloat16
    return {"qweight": torch.from_numpy(m["qweight"]), "qzeros": torch.from_numpy(m["qzeros"]),
            "scales": torch.from_numpy(m["scales"].view(np.int16)).view(tdt)}


def make_checkpoint(oracle, hidden=512, heads=8, kv_heads=2, head_dim=64, inter=1024, G=128, dt=0):
    shapes = {"q": (hidden, heads * head_dim), "k": (hidden, kv_heads * head_dim), "v": (hidden, kv_heads * head_dim),
              "o": (heads * head_dim, hidden), "gate": (hidden, inter), "up": (hidden, inter), "down": (inter, hidden)}
    return {name: make_matrix(oracle, K, N, G, seed=11 + i, dt=dt) for i, (name, (K, N)) in enumerate(shapes.items())}


def load_param_set(layer, mats, shard_ids):
    """mats: list of matrices, shard_ids: matching loader shard ids (None for unfused layers)."""
    for m, sid in zip(mats, shard_ids):
        t = to_torch(m)
        for pname in ("qweight", "qzeros", "scales"):
            p = getattr(layer, pname)
            if sid is None:
                p.weight_loader(p, t[pname])
            else:
                p.weight_loader(p, t[pname], sid)


def build_layers(ckpt, cfg, tp_rank, tp_size, hidden=512, heads=8, kv_heads=2, head_dim=64, inter=1024,
                 dtype=torch.float16):
    qkv = QKVParallelLinear(hidden, head_dim, heads, kv_heads, bias=False, quant_config=cfg, params_dtype=dtype,
                            prefix="model.layers.0.self_attn.qkv_proj", tp_rank=tp_rank, tp_size=tp_size)
    o = RowParallelLinear(heads * head_dim, hidden, bias=False, quant_config=cfg, params_dtype=dtype,
                          prefix="model.layers.0.self_attn.o_proj", tp_rank=tp_rank, tp_size=tp_size)
    gate_up = MergedColumnParallelLinear(hidden, [inter, inter], bias=False, quant_config=cfg, params_dtype=dtype,
                                         prefix="model.layers.0.mlp.gate_up_proj", tp_rank=tp_rank, tp_size=tp_size)
    down = RowParallelLinear(inter, hidden, bias=False, quant_config=cfg, params_dtype=dtype,
                             prefix="model.layers.0.mlp.down_proj", tp_rank=tp_rank, tp_size=tp_size)
    load_param_set(qkv, [ckpt["q"], ckpt["k"], ckpt["v"]], ["q", "k", "v"])
    load_param_set(o, [ckpt["o"]], [None])
    load_param_set(gate_up, [ckpt["gate"], ckpt["up"]], [0, 1])
    load_param_set(down, [ckpt["down"]], [None])
    return {"qkv": qkv, "o": o, "gate_up": gate_up, "down": down}


def dense(oracle, m, dt=0):
    """fp32 dense [K, N] of a checkpoint matrix via the oracle."""
    return oracle.bits_to_f32(oracle.dequantize(m["qweight"], m["scales"], m["qzeros"], dt), dt)


def layer_dense(oracle, layer, dt=0):
    """fp32 dense matrix of what a (not yet post-processed) layer holds."""
    bits = layer.scales.data.view(torch.int16).numpy().view(np.uint16)
    return oracle.bits_to_f32(oracle.dequantize(layer.qweight.data.numpy(), bits, layer.qzeros.data.numpy(), dt), dt)
