"""-m gpu: AWQLinearMethod.create_weights / process_weights_after_loading / apply and the
Column/RowParallel layers on the real kernels (reference has no layer-level test; SURVEY.md §4
lists the gaps this closes)."""
import numpy as np
import pytest
import torch

from gpu_util import rel_fro
from layer_util import awq_cfg, build_layers, dense, make_checkpoint, make_matrix, to_torch

from sglang_awq_b200.srt.layers import linear as L
from sglang_awq_b200.srt.layers.quantization import AWQLinearMethod

pytestmark = pytest.mark.gpu
DIMS = dict(hidden=1024, heads=8, kv_heads=2, head_dim=128, inter=2048)


def _finish(layers):
    for l in layers.values():
        l.cuda()
        l.quant_method.process_weights_after_loading(l)
    return layers


def test_apply_matches_oracle_single_rank(cabi, oracle):
    ck = make_checkpoint(oracle, **DIMS)
    layers = _finish(build_layers(ck, awq_cfg(), 0, 1, **DIMS))
    full = {k: torch.from_numpy(dense(oracle, m)).cuda() for k, m in ck.items()}
    g = torch.Generator(device="cuda").manual_seed(1)
    for M in (1, 7, 64, 130):
        x = torch.randn(2, M, 1024, generator=g, device="cuda").half()
        y, bias = layers["qkv"](x)
        assert bias is None and y.shape == (2, M, 1024 + 2 * 256)
        ref = x.float() @ torch.cat([full["q"], full["k"], full["v"]], dim=1)
        assert rel_fro(y, ref) <= 1e-3
        h = torch.randn(M, 2048, generator=g, device="cuda").half()
        y, _ = layers["down"](h)
        assert rel_fro(y, h.float() @ full["down"]) <= 1e-3
    assert layers["qkv"].awq_layout == "tiled" and layers["qkv"].qweight.shape == (8 * 12, 2144)


@pytest.mark.parametrize("tp", [2, 4, 8])
def test_tp_equivalence_on_one_gpu(cabi, oracle, tp):
    """Every rank's layer is built and run on this GPU: concat of column-parallel outputs and the
    SUM of row-parallel partials (what the all-reduce computes) equal the unsharded result."""
    ck = make_checkpoint(oracle, **DIMS)
    full = {k: torch.from_numpy(dense(oracle, m)).cuda() for k, m in ck.items()}
    ranks = [_finish(build_layers(ck, awq_cfg(), r, tp, **DIMS)) for r in range(tp)]
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(5, 1024, generator=g, device="cuda").half()
    inter_r = 2048 // tp
    outs = [rk["gate_up"](x)[0] for rk in ranks]
    gate = torch.cat([o[:, :inter_r] for o in outs], dim=1)
    up = torch.cat([o[:, inter_r:] for o in outs], dim=1)
    one = _finish(build_layers(ck, awq_cfg(), 0, 1, **DIMS))
    ref = one["gate_up"](x)[0]
    # column shards see the same x: bit-identical to the unsharded layer whenever the measured router
    # sends both widths to the same decode kernel (same k order); within the GEMV bound otherwise
    L = cabi.lib()
    same_kernel = L.awq_b200_gemv_route(5, 1024, 2 * inter_r, 0) == L.awq_b200_gemv_route(5, 1024, 4096, 0)
    for got, want in ((gate, ref[:, :2048]), (up, ref[:, 2048:])):
        if same_kernel:
            assert torch.equal(got.view(torch.int16), want.view(torch.int16))
        else:
            assert rel_fro(got, want.float()) <= 1e-3
    h = torch.randn(5, 2048, generator=g, device="cuda").half()
    partial = sum(rk["down"](h[:, r * inter_r:(r + 1) * inter_r].contiguous(), skip_all_reduce=True)[0].float()
                  for r, rk in enumerate(ranks))
    assert rel_fro(partial, h.float() @ full["down"]) <= 1e-3


def test_bias_fused_and_row_parallel_bias_rank0_only(cabi, oracle):
    m = make_matrix(oracle, 512, 384, 128, seed=3)
    cfg = awq_cfg()
    W = torch.from_numpy(dense(oracle, m)).cuda()
    t = to_torch(m)
    outs = []
    for rank in (0, 1):
        lin = L.RowParallelLinear(512, 384, bias=True, quant_config=cfg, params_dtype=torch.float16, tp_rank=rank,
                                  tp_size=2)
        for p in ("qweight", "qzeros", "scales"):
            getattr(lin, p).weight_loader(getattr(lin, p), t[p])
        lin.bias.data.fill_(0.25)
        lin.cuda()
        lin.quant_method.process_weights_after_loading(lin)
        x = torch.ones(3, 256, device="cuda").half()
        outs.append(lin(x, skip_all_reduce=True)[0].float())
    ref = torch.ones(3, 512, device="cuda") @ W + 0.25
    assert rel_fro(outs[0] + outs[1], ref) <= 1e-3  # bias counted once


def test_dummy_loader_ordering_and_idempotence(cabi, oracle):
    """DummyModelLoader runs process_weights_after_loading BEFORE filling float params with
    uniform noise in place (reference loader.py:1211-1218): the repacked `scales` must stay a
    float Parameter that tolerates that, and a second post-process call must be a no-op."""
    lin = L.ColumnParallelLinear(512, 256, bias=False, quant_config=awq_cfg(), params_dtype=torch.float16, tp_rank=0,
                                 tp_size=1).cuda()
    lin.qweight.data.random_(-2**31, 2**31 - 1)
    lin.qzeros.data.random_(-2**31, 2**31 - 1)
    lin.scales.data.fill_(1.0)
    lin.quant_method.process_weights_after_loading(lin)
    before = lin.qweight.data_ptr()
    lin.quant_method.process_weights_after_loading(lin)
    assert lin.qweight.data_ptr() == before and lin.awq_layout == "tiled"
    assert lin.scales.dtype == torch.float16 and lin.scales.is_floating_point()
    assert sorted(n for n, _ in lin.named_parameters()) == ["qweight", "qzeros", "scales"]
    for _, p in lin.named_parameters():  # what initialize_dummy_weights does (weight_utils.py:1108-1137)
        if torch.is_floating_point(p):
            p.data.uniform_(-1e-3, 1e-3)
    y, _ = lin(torch.randn(4, 512, device="cuda").half())
    assert torch.isfinite(y).all() and y.abs().max() < 1.0
    W = AWQLinearMethod.dequantize(lin)
    assert W.shape == (512, 256) and W.abs().max() <= 15e-3 + 1e-6


@pytest.mark.parametrize("G", [32, 64, 256, -1])
@pytest.mark.parametrize("M", [9, 200])
def test_other_group_sizes_take_the_fused_path(cabi, oracle, G, M):
    """the reference op's other group sizes (awq_triton.py:10: -1, 32, 64): 32 / 64 -> records with 4 / 2 metadata
    sections (tcgen05 kernel), multiples of 128 (incl. -1 = K) -> plain records with the group's metadata repeated"""
    K, N = 512, 256
    g_eff = K if G == -1 else G
    m = make_matrix(oracle, K, N, g_eff, seed=4)
    lin = L.ReplicatedLinear(K, N, bias=True, quant_config=awq_cfg(G), params_dtype=torch.float16)
    t = to_torch(m)
    for p in ("qweight", "qzeros", "scales"):
        getattr(lin, p).weight_loader(getattr(lin, p), t[p])
    lin.bias.data.zero_()
    lin.cuda()
    lin.quant_method.process_weights_after_loading(lin)
    assert lin.awq_layout == "tiled" and lin.qweight.shape[1] == {32: 2432, 64: 2240}.get(g_eff, 2144)
    W = torch.from_numpy(dense(oracle, m)).cuda()
    assert torch.equal(AWQLinearMethod.dequantize(lin).float(), W), "tiled dequant must give the reference's bits"
    x = torch.randn(M, K, device="cuda").half()
    y, _ = lin(x)
    assert rel_fro(y, x.float() @ W) <= 1e-3
    # the float views reach the bytes the kernels read (dummy loader)
    lin.scales.data.mul_(2)
    y2, _ = lin(x)
    assert rel_fro(y2, 2 * (x.float() @ W)) <= 1e-3


def test_unsupported_group_size_keeps_the_reference_path(cabi, oracle):
    m = make_matrix(oracle, 384, 128, 96, seed=4)  # 96 divides 384 but is neither 32 / 64 nor a multiple of 128
    lin = L.ReplicatedLinear(384, 128, bias=False, quant_config=awq_cfg(96), params_dtype=torch.float16)
    t = to_torch(m)
    for p in ("qweight", "qzeros", "scales"):
        getattr(lin, p).weight_loader(getattr(lin, p), t[p])
    lin.cuda()
    lin.quant_method.process_weights_after_loading(lin)
    assert lin.awq_layout == "awq" and type(lin.qweight) is torch.nn.Parameter
    x = torch.randn(9, 384, device="cuda").half()
    y, _ = lin(x)
    assert rel_fro(y, x.float() @ torch.from_numpy(dense(oracle, m)).cuda()) <= 1e-3


def test_cpu_tensor_is_rejected(cabi, oracle):
    lin = L.ReplicatedLinear(256, 128, bias=False, quant_config=awq_cfg(), params_dtype=torch.float16)
    lin.quant_method.process_weights_after_loading(lin)  # CPU-resident: stays in reference layout
    with pytest.raises((NotImplementedError, RuntimeError)):
        lin(torch.randn(1, 256).half())  # no CPU fallback


def test_awq_marlin_front_end_runs_the_tile_kernels(cabi, oracle):
    """`awq_marlin` (what the reference resolves AWQ checkpoints to, awq.py:267-289) through the
    layer API: same results as `awq`, bit for bit (same repack, same kernels), within 1e-3 of the
    dense reference, fp16 and bf16 activations (its supported act dtypes)"""
    from sglang_awq_b200.srt.layers.quantization import AWQMarlinConfig

    ck = make_checkpoint(oracle, **DIMS)
    mcfg = AWQMarlinConfig.from_config({"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True})
    a = _finish(build_layers(ck, awq_cfg(), 0, 1, **DIMS))
    m = _finish(build_layers(ck, mcfg, 0, 1, **DIMS))
    full = {k: torch.from_numpy(dense(oracle, mat)).cuda() for k, mat in ck.items()}
    g = torch.Generator(device="cuda").manual_seed(2)
    for M in (1, 16, 200):
        x = torch.randn(M, 1024, generator=g, device="cuda").half()
        ya, _ = a["gate_up"](x)
        ym, _ = m["gate_up"](x)
        assert type(m["gate_up"].quant_method).__name__ == "AWQMarlinLinearMethod" and m["gate_up"].awq_layout == "tiled"
        assert torch.equal(ya.view(torch.int16), ym.view(torch.int16))
        assert rel_fro(ym, x.float() @ torch.cat([full["gate"], full["up"]], dim=1)) <= 1e-3


def test_tiled_layer_state_dict_round_trip_and_dtype_guard(cabi, oracle):
    """`scales` / `qzeros` of a tiled layer are strided VIEWS into `qweight`'s records: a state_dict
    must load back in place into a tiled layer of the same shape (keeping the aliasing) and still
    dequantise bit-exactly; `retile_views` restores the aliasing after a de-aliasing copy; and an
    activation dtype other than the one the scales were repacked with is refused, as the
    reference's matmul would (awq.py:447)."""
    import copy

    m = make_matrix(oracle, 512, 384, 128, seed=9)
    t = to_torch(m)

    def make(fill):
        lin = L.ColumnParallelLinear(512, 384, bias=False, quant_config=awq_cfg(), params_dtype=torch.float16, tp_rank=0,
                                     tp_size=1)
        if fill:
            for p in ("qweight", "qzeros", "scales"):
                getattr(lin, p).weight_loader(getattr(lin, p), t[p])
        else:
            lin.qweight.data.zero_(), lin.qzeros.data.zero_(), lin.scales.data.zero_()
        lin.cuda()
        lin.quant_method.process_weights_after_loading(lin)
        return lin

    a, b = make(True), make(False)
    ref_bits = oracle.dequantize(m["qweight"], m["scales"], m["qzeros"], 0)
    wa = AWQLinearMethod.dequantize(a)
    assert np.array_equal(wa.view(torch.int16).cpu().numpy().view(np.uint16), ref_bits)
    assert not AWQLinearMethod.dequantize(b).any()
    b.load_state_dict(a.state_dict())
    assert b.scales.data_ptr() == b.qweight.data_ptr() + 8192  # still a view into the records
    assert torch.equal(AWQLinearMethod.dequantize(b).view(torch.int16), wa.view(torch.int16))
    x = torch.randn(3, 512, device="cuda").half()
    assert torch.equal(b(x)[0].view(torch.int16), a(x)[0].view(torch.int16))
    c = copy.deepcopy(a)
    AWQLinearMethod.retile_views(c)
    assert c.scales.data_ptr() == c.qweight.data_ptr() + 8192
    assert torch.equal(c(x)[0].view(torch.int16), a(x)[0].view(torch.int16))
    with pytest.raises(RuntimeError, match="dtype"):
        a(x.to(torch.bfloat16))
