"""CPU: host-side mirror of the reference's plugin API -- config, create_weights, TP weight
loaders (sharding index math incl. packed dims and kv-head replication), and the world_size-2
gloo path of Row/ColumnParallelLinear.  The oracle stands in for the CUDA kernels ONLY as the
checker inside these tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from layer_util import awq_cfg, build_layers, dense, layer_dense, make_checkpoint, to_torch

from sglang_awq_b200.srt.layers import linear as L
from sglang_awq_b200.srt.layers.parameter import GroupQuantScaleParameter, PackedvLLMParameter
from sglang_awq_b200.srt.layers.quantization import (AWQConfig, AWQLinearMethod, UnquantizedLinearMethod,
                                                     get_quantization_config)


def test_config_and_registry():
    cls = get_quantization_config("awq")
    assert cls is AWQConfig
    with pytest.raises(ValueError):
        get_quantization_config("gptq_marlin")
    cfg = cls.from_config({"bits": 4, "group_size": 128, "zero_point": True, "modules_to_not_convert": ["lm_head"]})
    assert (cfg.weight_bits, cfg.group_size, cfg.zero_point, cfg.pack_factor) == (4, 128, True, 8)
    assert cfg.get_name() == "awq" and cfg.get_supported_act_dtypes() == [torch.float16]
    assert cfg.get_config_filenames() == ["quant_config.json", "quantize_config.json"]
    assert cfg.get_scaled_act_names() == []
    with pytest.raises(ValueError):
        AWQConfig(8, 128, True)
    with pytest.raises(ValueError):
        cls.from_config({"bits": 4})
    lin = L.ReplicatedLinear(256, 128, bias=False, quant_config=cfg, params_dtype=torch.float16, prefix="model.x")
    assert type(lin.quant_method).__name__ == "AWQLinearMethod"
    skipped = L.ReplicatedLinear(256, 128, bias=False, quant_config=cfg, params_dtype=torch.float16, prefix="lm_head")
    assert isinstance(skipped.quant_method, UnquantizedLinearMethod) and skipped.weight.shape == (128, 256)
    assert cfg.get_quant_method(torch.nn.ReLU(), "x") is None
    assert "AWQLinearMethod" in L.WEIGHT_LOADER_V2_SUPPORTED


def test_create_weights_shapes_and_errors():
    cfg = awq_cfg()
    lin = L.ColumnParallelLinear(4096, 11008, bias=True, quant_config=cfg, params_dtype=torch.float16, tp_rank=0,
                                 tp_size=2)
    assert isinstance(lin.qweight, PackedvLLMParameter) and isinstance(lin.scales, GroupQuantScaleParameter)
    assert lin.qweight.shape == (4096, 688) and lin.qweight.dtype == torch.int32
    assert lin.qzeros.shape == (32, 688) and lin.scales.shape == (32, 5504) and lin.bias.shape == (5504,)
    assert (lin.qweight.input_dim, lin.qweight.output_dim, lin.qweight.packed_dim, lin.qweight.packed_factor) == (0, 1, 1, 8)
    assert lin.qweight.weight_loader == lin.weight_loader_v2
    with pytest.raises(ValueError, match="input size is not aligned"):
        L.RowParallelLinear(4096 + 64, 128, quant_config=cfg, params_dtype=torch.float16, tp_rank=0, tp_size=1)
    with pytest.raises(ValueError, match="output size is not aligned"):
        L.ColumnParallelLinear(256, 132, quant_config=cfg, params_dtype=torch.float16, tp_rank=0, tp_size=1)
    # Llama-3-70B TP=8 per-rank shapes
    qkv = L.QKVParallelLinear(8192, 128, 64, 8, bias=False, quant_config=cfg, params_dtype=torch.float16, tp_rank=0,
                              tp_size=8)
    assert qkv.qweight.shape == (8192, 160)
    down = L.RowParallelLinear(28672, 8192, bias=False, quant_config=cfg, params_dtype=torch.float16, tp_rank=7,
                               tp_size=8)
    assert down.qweight.shape == (3584, 1024) and down.scales.shape == (28, 8192)


@pytest.mark.parametrize("tp", [1, 2, 4, 8])
def test_tp_loaders_shard_exactly(oracle, tp):
    """concat_N(column shards) == unsharded and stack_K(row shards) == unsharded, bit-exact, with
    kv heads replicated when tp > kv_heads (8 q heads, 2 kv heads)."""
    dims = dict(hidden=1024, heads=8, kv_heads=2, head_dim=128, inter=2048)  # K/8 stays a multiple of 128
    ck = make_checkpoint(oracle, **dims)
    full = {k: dense(oracle, m) for k, m in ck.items()}
    per_rank = [build_layers(ck, awq_cfg(), r, tp, **dims) for r in range(tp)]
    q = np.concatenate([layer_dense(oracle, pr["qkv"])[:, :pr["qkv"].q_proj_shard_size] for pr in per_rank], axis=1)
    assert np.array_equal(q, full["q"])
    for r, pr in enumerate(per_rank):
        l = pr["qkv"]
        d = layer_dense(oracle, l)
        qs, kvs = l.q_proj_shard_size, l.kv_proj_shard_size
        kv_index = r // l.num_kv_head_replicas
        assert np.array_equal(d[:, qs:qs + kvs], full["k"][:, kv_index * kvs:(kv_index + 1) * kvs])
        assert np.array_equal(d[:, qs + kvs:], full["v"][:, kv_index * kvs:(kv_index + 1) * kvs])
    if tp > 2:
        assert per_rank[0]["qkv"].num_kv_heads == 1 and per_rank[0]["qkv"].num_kv_head_replicas == tp // 2
    inter_r = 2048 // tp
    gate = np.concatenate([layer_dense(oracle, pr["gate_up"])[:, :inter_r] for pr in per_rank], axis=1)
    up = np.concatenate([layer_dense(oracle, pr["gate_up"])[:, inter_r:] for pr in per_rank], axis=1)
    assert np.array_equal(gate, full["gate"]) and np.array_equal(up, full["up"])
    for name in ("o", "down"):
        stacked = np.concatenate([layer_dense(oracle, pr[name]) for pr in per_rank], axis=0)
        assert np.array_equal(stacked, full[name])
    # and the reference's alignment error when a K shard is not a whole number of groups
    if tp == 8:
        with pytest.raises(ValueError, match="input size is not aligned"):
            build_layers(make_checkpoint(oracle), awq_cfg(), 0, 8)


def test_fused_checkpoint_tensors_are_split(oracle):
    """Checkpoints that store qkv / gate_up already fused are split by packed offsets."""
    ck = make_checkpoint(oracle)
    cfg = awq_cfg()
    for tp_rank in (0, 1):
        ref = build_layers(ck, cfg, tp_rank, 2)
        gu = L.MergedColumnParallelLinear(512, [1024, 1024], bias=False, quant_config=cfg, params_dtype=torch.float16,
                                          tp_rank=tp_rank, tp_size=2)
        qkv = L.QKVParallelLinear(512, 64, 8, 2, bias=False, quant_config=cfg, params_dtype=torch.float16,
                                  tp_rank=tp_rank, tp_size=2)
        for pname in ("qweight", "qzeros", "scales"):
            fused = torch.cat([to_torch(ck["gate"])[pname], to_torch(ck["up"])[pname]], dim=1)
            p = getattr(gu, pname)
            p.weight_loader(p, fused)
            assert torch.equal(p.data, getattr(ref["gate_up"], pname).data)
            fused = torch.cat([to_torch(ck[n])[pname] for n in ("q", "k", "v")], dim=1)
            p = getattr(qkv, pname)
            p.weight_loader(p, fused)
            assert torch.equal(p.data, getattr(ref["qkv"], pname).data)


def test_bias_loaders():
    cfg = awq_cfg()
    qkv = L.QKVParallelLinear(512, 64, 8, 2, bias=True, quant_config=cfg, params_dtype=torch.float16, tp_rank=1,
                              tp_size=2)
    bq, bk, bv = torch.arange(512.).half(), torch.arange(128.).half() + 1000, torch.arange(128.).half() + 2000
    for b, sid in ((bq, "q"), (bk, "k"), (bv, "v")):
        qkv.bias.weight_loader(qkv.bias, b, sid)
    assert torch.equal(qkv.bias.data, torch.cat([bq[256:], bk[64:], bv[64:]]))
    row = L.RowParallelLinear(1024, 512, bias=True, quant_config=cfg, params_dtype=torch.float16, tp_rank=1, tp_size=2)
    row.bias.weight_loader(row.bias, torch.ones(512).half())
    assert row.bias.data.eq(1).all()


def test_loader_shape_mismatch_raises(oracle):
    cfg = awq_cfg()
    lin = L.ColumnParallelLinear(256, 256, bias=False, quant_config=cfg, params_dtype=torch.float16, tp_rank=0,
                                 tp_size=2)
    with pytest.raises((AssertionError, ValueError, RuntimeError)):
        lin.qweight.weight_loader(lin.qweight, torch.zeros(128, 32, dtype=torch.int32))
    with pytest.raises((AssertionError, ValueError, RuntimeError)):
        lin.qweight.weight_loader(lin.qweight, torch.zeros(256, 8, dtype=torch.int32))  # too few columns for rank


# ---------------------------------------------------------------------------------------------
# world_size = 2 over gloo: the N > 1 code path (all-reduce / all-gather) on CPU
# ---------------------------------------------------------------------------------------------

def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tp_worker(rank, world, port, ret):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from layer_util import awq_cfg, build_layers, dense, make_checkpoint
    from oracle import awq_oracle as oracle
    from sglang_awq_b200.srt import distributed as D
    from sglang_awq_b200.srt.layers.quantization import awq as awq_mod

    D.init_distributed_environment(world_size=world, rank=rank, distributed_init_method=f"tcp://127.0.0.1:{port}",
                                   backend="gloo")
    D.initialize_model_parallel(world)
    assert D.get_tensor_model_parallel_world_size() == world and D.get_tensor_model_parallel_rank() == rank

    # test-only checker: same class NAME (the layers select weight_loader_v2 by it), apply = oracle
    class AWQLinearMethod(awq_mod.AWQLinearMethod):
        def process_weights_after_loading(self, layer):
            return

        def apply(self, layer, x, bias=None):
            bits = layer.scales.data.view(torch.int16).numpy().view(np.uint16)
            w = oracle.bits_to_f32(oracle.dequantize(layer.qweight.data.numpy(), bits, layer.qzeros.data.numpy(), 0), 0)
            y = x.float() @ torch.from_numpy(w)
            if bias is not None:
                y = y + bias.float()
            return y

    class Cfg(awq_mod.AWQConfig):
        def get_quant_method(self, layer, prefix):
            return AWQLinearMethod(self)

    cfg = Cfg(4, 128, True)
    ck = make_checkpoint(oracle)
    layers = build_layers(ck, cfg, None, None)  # tp rank/size come from the process group
    assert layers["down"].tp_size == world and layers["down"].tp_rank == rank
    g = torch.Generator().manual_seed(5)
    x = torch.randn(3, 512, generator=g)
    h = torch.randn(3, 1024, generator=g)
    # column parallel + gather_output
    layers["gate_up"].gather_output = True
    gu, _ = layers["gate_up"](x)
    # row parallel: input already split
    hr = h[:, rank * 512:(rank + 1) * 512].contiguous()
    dn, _ = layers["down"](hr)
    # row parallel: input_is_parallel=False splits internally; bias added once (rank 0)
    layers["down"].input_is_parallel = False
    layers["down"].bias = torch.nn.Parameter(torch.full((512,), 0.5))
    dn2, _ = layers["down"](h)
    if rank == 0:
        full = {k: torch.from_numpy(dense(oracle, m)) for k, m in ck.items()}
        ref_gu = torch.cat([(x @ full["gate"])[:, :512], (x @ full["up"])[:, :512],
                            (x @ full["gate"])[:, 512:], (x @ full["up"])[:, 512:]], dim=1)
        ref_dn = h @ full["down"]
        ret["gu"] = float((gu - ref_gu).abs().max())
        ret["dn"] = float((dn - ref_dn).abs().max() / ref_dn.abs().max())
        ret["dn2"] = float((dn2 - (ref_dn + 0.5)).abs().max() / ref_dn.abs().max())
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_world_size_2_gloo():
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_tp_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret["gu"] == 0.0          # all-gather of column shards: exact
    assert ret["dn"] < 1e-5          # all-reduce of row partials: summation order only
    assert ret["dn2"] < 1e-5


def test_awq_marlin_front_end_config_and_routing():
    """reference awq.py:182-349: the config an AWQ checkpoint resolves to by default, its override
    rule, its support checks (marlin_utils.py:115-233) and its fall-back to AWQLinearMethod"""
    from sglang_awq_b200.srt.layers.quantization import AWQMarlinConfig, AWQMarlinLinearMethod
    from sglang_awq_b200.srt.layers.quantization import marlin_utils as mu

    hf = {"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True, "modules_to_not_convert": ["lm_head"]}
    assert get_quantization_config("awq_marlin") is AWQMarlinConfig
    cfg = AWQMarlinConfig.from_config(hf)
    assert cfg.get_name() == "awq_marlin" and cfg.pack_factor == 8 and cfg.full_config is hf
    assert cfg.get_supported_act_dtypes() == [torch.half, torch.bfloat16]
    assert cfg.get_config_filenames() == ["quantize_config.json"] and not cfg.lm_head_quantized
    # support checks: zero point required, group sizes {-1, 32, 64, 128}, 4 bits
    for bad in ({**hf, "zero_point": False}, {**hf, "group_size": 256}, {**hf, "bits": 8}, {**hf, "bits": 3}):
        with pytest.raises(ValueError):
            AWQMarlinConfig.from_config(bad)
    assert mu.check_marlin_supported(4, 64, has_zp=True) and not mu.check_marlin_supported(4, 64, has_zp=False)
    # override rule needs a CUDA device to say yes (reference: `if not _is_cuda: return False`)
    want = "awq_marlin" if torch.cuda.is_available() else None
    assert AWQMarlinConfig.override_quantization_method(hf, None) == want
    assert AWQMarlinConfig.override_quantization_method(hf, "awq") is None
    assert AWQMarlinConfig.override_quantization_method({**hf, "quant_method": "gptq"}, None) is None
    assert AWQMarlinConfig.is_awq_marlin_compatible({"quant_method": "awq", "bits": 4}) is False
    # layer routing
    lin = L.ColumnParallelLinear(4096, 11008, bias=False, quant_config=cfg, params_dtype=torch.float16, tp_rank=1,
                                 tp_size=2, prefix="model.layers.0.mlp.up_proj")
    assert type(lin.quant_method) is AWQMarlinLinearMethod and isinstance(lin.quant_method, AWQLinearMethod)
    assert lin.qweight.shape == (4096, 688) and lin.qzeros.shape == (32, 688) and lin.scales.shape == (32, 5504)
    assert (lin.input_size_per_partition, lin.output_size_per_partition, lin.num_groups) == (4096, 5504, 32)
    assert lin.qweight.weight_loader == lin.weight_loader_v2 and "AWQMarlinLinearMethod" in L.WEIGHT_LOADER_V2_SUPPORTED
    skipped = L.ReplicatedLinear(256, 128, bias=False, quant_config=cfg, params_dtype=torch.float16, prefix="lm_head")
    assert isinstance(skipped.quant_method, UnquantizedLinearMethod)
    # N not divisible by 64: "not supported by AWQMarlin", falls back to the plain AWQ method
    odd = L.ReplicatedLinear(256, 72, bias=False, quant_config=cfg, params_dtype=torch.float16, prefix="model.odd")
    assert type(odd.quant_method) is AWQLinearMethod
    # shape errors surface from create_weights with the reference's wording
    with pytest.raises(ValueError, match="min_thread_k"):
        mu.verify_marlin_supports_shape(128, 192, 4096, 128)
    with pytest.raises(ValueError, match="min_thread_n"):
        mu.verify_marlin_supports_shape(96, 256, 4096, 128)
    with pytest.raises(ValueError, match="divisible by group_size"):
        mu.verify_marlin_supports_shape(128, 384, 4096, 256)
    # group_size -1 means one group over the whole input (awq.py:478-482)
    whole = AWQMarlinLinearMethod(AWQMarlinConfig.from_config({**hf, "group_size": -1}))
    mod = torch.nn.Module()
    whole.create_weights(mod, 384, [128], 384, 128, torch.float16, weight_loader=None)
    assert mod.num_groups == 1 and mod.scales.shape == (1, 128)
    assert cfg.get_quant_method(torch.nn.ReLU(), "x") is None


def test_awq_marlin_loads_like_awq():
    """same checkpoint through both front-ends and a TP=2 shard: identical parameter tensors"""
    from sglang_awq_b200.srt.layers.quantization import AWQMarlinConfig
    from oracle import awq_oracle as oracle

    dims = dict(hidden=1024, heads=8, kv_heads=8, head_dim=128, inter=2048)
    ck = make_checkpoint(oracle, **dims)
    mcfg = AWQMarlinConfig.from_config({"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True})
    for rank, size in ((0, 1), (1, 2)):
        a = build_layers(ck, awq_cfg(), rank, size, **dims)
        m = build_layers(ck, mcfg, rank, size, **dims)
        for name in a:
            assert type(m[name].quant_method).__name__ == "AWQMarlinLinearMethod"
            for pn in ("qweight", "qzeros", "scales"):
                assert torch.equal(getattr(a[name], pn).data, getattr(m[name], pn).data), (name, pn, rank, size)


def test_awq_moe_create_weights_and_loader_cpu():
    """AWQMoEMethod.create_weights mirrors the reference's parameter names / shapes (awq.py:669-765) and the expert
    loader shards w1 / w3 along the output and w2 along the input (is_transposed AWQ tensors)"""
    from sglang_awq_b200.srt.layers.moe import FusedMoE
    from sglang_awq_b200.srt.layers.quantization import AWQConfig, AWQMarlinConfig, AWQMoEMethod

    E, hidden, inter, tp = 4, 256, 512, 2
    for cfg in (AWQConfig.from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True}),):
        for rank in range(tp):
            m = FusedMoE(E, hidden, inter, 2, torch.float16, cfg, tp_rank=rank, tp_size=tp)
            assert isinstance(m.quant_method, AWQMoEMethod)
            shapes = {n: tuple(p.shape) for n, p in m.named_parameters()}
            assert shapes == {"w13_qweight": (E, hidden, 2 * inter // tp // 8), "w2_qweight": (E, inter // tp, hidden // 8),
                              "w13_scales": (E, hidden // 128, 2 * inter // tp), "w2_scales": (E, inter // tp // 128, hidden),
                              "w13_qzeros": (E, hidden // 128, 2 * inter // tp // 8), "w2_qzeros": (E, inter // tp // 128, hidden // 8)}
            assert m.w13_qweight.is_transposed and m.w13_qweight.quant_method == "group"
            w1 = torch.arange(hidden * inter // 8, dtype=torch.int32).reshape(hidden, inter // 8)
            w3 = -w1
            w2 = torch.arange(inter * hidden // 8, dtype=torch.int32).reshape(inter, hidden // 8)
            m.weight_loader(m.w13_qweight, w1, "w1", "w1", 1)
            m.weight_loader(m.w13_qweight, w3, "w3", "w3", 1)
            m.weight_loader(m.w2_qweight, w2, "w2", "w2", 3)
            half = inter // tp // 8
            assert torch.equal(m.w13_qweight.data[1, :, :half], w1[:, rank * half:(rank + 1) * half])
            assert torch.equal(m.w13_qweight.data[1, :, half:], w3[:, rank * half:(rank + 1) * half])
            assert torch.equal(m.w2_qweight.data[3], w2[rank * (inter // tp):(rank + 1) * (inter // tp)])
    with pytest.raises(ValueError):
        FusedMoE(E, hidden, 192, 2, torch.float16, AWQConfig.from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True}),
                 tp_rank=0, tp_size=1)
