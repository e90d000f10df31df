"""-m gpu: the fused epilogues either side of the AWQ linears, through the C ABI.

  gate_up_proj + SiluAndMul   reference python/sglang/srt/models/llama.py:94-106 (+ activation.py SiluAndMul)
  residual add + RMSNorm      reference python/sglang/srt/layers/layernorm.py:104-126, :178-222

Bar: relative Frobenius error <= 1e-3 against the oracle (double-precision activation / normalisation over
the fp32-accumulated linear on the reference-dequantised weights); the repack is checked bit for bit; the
residual output of the norm is one rounding of an exact sum and is checked bit for bit against the golden
vectors made by the reference's own forward_native."""
import glob
import os

import numpy as np
import pytest
import torch

from gpu_util import TDT, big_inputs, dev_inputs, from_bits, rel_fro, to_bits

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("K,inter", [(128, 8), (256, 200), (512, 192), (1024, 1024)])
@pytest.mark.parametrize("dt", [0, 1])
def test_repack_gate_up_is_the_plain_repack_of_permuted_columns(cabi, oracle, K, inter, dt):
    N = 2 * inter
    d = dev_inputs(oracle, K, N, 128, dt, seed=inter + dt)
    perm = oracle.gate_up_interleave_perm(inter)
    qw, sc, qz = oracle.permute_columns(d["np"]["qweight"], d["np"]["scales"], d["np"]["qzeros"], perm)
    want = cabi.repack(torch.from_numpy(qw).cuda(), from_bits(sc, dt), torch.from_numpy(qz).cuda())
    got = cabi.repack_gate_up(d["qweight"], d["scales"], d["qzeros"])
    assert torch.equal(got, want)
    # and against the oracle's byte-level spec of the tile layout
    assert np.array_equal(got.cpu().numpy().view(np.uint8).reshape(-1, 8576), oracle.repack(qw, sc, qz))


def _silu_ref(oracle, d, dt, cabi):
    """oracle: T-rounded fp32-accumulated linear over the bit-exact dequantised weights, then silu(g) * u in double"""
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
    gu = (d["x"].float() @ w).to(TDT[dt])
    _, y = oracle.silu_mul(to_bits(gu), dt)
    return torch.from_numpy(y).cuda()


# impl: 0 = routed, 3 = persistent kernel, 5 = tcgen05 decode kernel; M > 32 -> tcgen05 GEMM
@pytest.mark.parametrize("M,impl", [(1, 0), (1, 3), (3, 3), (8, 3), (8, 5), (9, 0), (16, 3), (16, 5), (17, 5), (32, 5),
                                    (33, 0), (64, 0), (200, 0), (600, 0)])
@pytest.mark.parametrize("K,inter", [(512, 192), (1024, 200), (2048, 1024)])
def test_linear_silu_mul_vs_oracle(cabi, oracle, M, impl, K, inter):
    N = 2 * inter
    d = dev_inputs(oracle, K, N, 128, 0, seed=M + K, M=M, signed_scales=True)
    d["x"] = d["x"] * 0.25
    tiles = cabi.repack_gate_up(d["qweight"], d["scales"], d["qzeros"])
    L = cabi.lib()
    L.awq_b200_debug_set_gemv_impl(impl, 0, -1)
    try:
        y = cabi.linear_silu_mul(d["x"], tiles, N)
        y2 = cabi.linear_silu_mul(d["x"], tiles, N)
    finally:
        L.awq_b200_debug_set_gemv_impl(0, 0, -1)
    assert y.shape == (M, inter) and y.dtype == torch.float16
    assert torch.equal(y.view(torch.int16), y2.view(torch.int16))
    assert rel_fro(y, _silu_ref(oracle, d, 0, cabi)) <= 1e-3


@pytest.mark.parametrize("M", [2, 16, 100])
def test_linear_silu_mul_bf16(cabi, oracle, M):
    K, inter = 1024, 512
    d = dev_inputs(oracle, K, 2 * inter, 128, 1, seed=M, M=M)
    d["x"] = d["x"] * 0.25
    tiles = cabi.repack_gate_up(d["qweight"], d["scales"], d["qzeros"])
    y = cabi.linear_silu_mul(d["x"], tiles, 2 * inter)
    ref = _silu_ref(oracle, d, 1, cabi)
    assert y.dtype == torch.bfloat16 and rel_fro(y, ref) <= 6e-3 and rel_fro(y, ref.to(torch.bfloat16)) <= 4e-3


# Llama-3-8B / 70B / Qwen2.5-32B gate_up (full and per-rank), every kernel family that serves them
@pytest.mark.parametrize("K,inter,M", [(4096, 14336, 1), (4096, 14336, 16), (4096, 14336, 512), (8192, 28672, 4),
                                       (8192, 28672, 16), (8192, 3584, 1), (8192, 3584, 32), (5120, 27648, 8),
                                       (5120, 6912, 16), (5120, 13824, 96)])
def test_linear_silu_mul_model_shapes(cabi, oracle, K, inter, M):
    N = 2 * inter
    d = big_inputs(K, N, 128, torch.float16, M=M, seed=K + inter + M)
    d["scales"] = (d["scales"].float() * (0.6 / (K ** 0.5) / 0.015)).half()  # gate_up outputs of O(1): silu is not linear
    tiles = cabi.repack_gate_up(d["qweight"], d["scales"], d["qzeros"])
    y = cabi.linear_silu_mul(d["x"], tiles, N)
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
    gu = (d["x"].float() @ w).half().double()
    g, u = gu[:, :inter], gu[:, inter:]
    assert rel_fro(y, g / (1 + torch.exp(-g)) * u) <= 1e-3
    # the fused launch against the unfused sequence on the SAME kernels: plain repack + linear, then the activation
    plain = cabi.linear(d["x"], cabi.repack(d["qweight"], d["scales"], d["qzeros"]), N).double()
    g, u = plain[:, :inter], plain[:, inter:]
    assert rel_fro(y, g / (1 + torch.exp(-g)) * u) <= 1e-3


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "epilogue_rmsnorm_*.npz"))), ids=os.path.basename)
def test_residual_rmsnorm_vs_reference_golden(cabi, oracle, path):
    g = np.load(path)
    dt = int(g["dtype"])
    x, res, w = from_bits(g["x"], dt), from_bits(g["residual"], dt), from_bits(g["weight"], dt)
    out = cabi.residual_rmsnorm(x, res, w, float(g["eps"]), out=torch.empty_like(x))
    assert np.array_equal(to_bits(res), g["residual_out"]), "residual <- T(x + residual) must be bit-exact"
    assert rel_fro(out, from_bits(g["out"], dt)) <= (2e-4 if dt == 0 else 2e-3)
    _, _, y64 = oracle.residual_rmsnorm(g["x"], g["residual"], g["weight"], float(g["eps"]), dt)
    assert rel_fro(out, torch.from_numpy(y64).cuda()) <= (1e-3 if dt == 0 else 4e-3)


@pytest.mark.parametrize("rows,hidden", [(1, 1024), (1, 8192), (16, 8192), (3, 5120), (64, 4096), (7, 3072), (5, 1000),
                                         (2, 16384), (300, 8192)])
@pytest.mark.parametrize("dt", [0, 1])
def test_residual_rmsnorm_shapes_in_place(cabi, oracle, rows, hidden, dt):
    """cluster-per-row kernel (hidden % 1024 == 0, <= 8192) and block-per-row kernel (the rest); out aliases x as
    the reference's fused_add_rmsnorm does"""
    g = torch.Generator(device="cuda").manual_seed(rows * hidden + dt)
    x = (torch.randn(rows, hidden, generator=g, device="cuda") * 0.8).to(TDT[dt])
    res = (torch.randn(rows, hidden, generator=g, device="cuda") * 1.1).to(TDT[dt])
    w = (torch.rand(hidden, generator=g, device="cuda") + 0.5).to(TDT[dt])
    xb, rb, wb = to_bits(x), to_bits(res), to_bits(w)
    cabi.residual_rmsnorm(x, res, w, 1e-5)
    _, want_res, y64 = oracle.residual_rmsnorm(xb, rb, wb, 1e-5, dt)
    assert np.array_equal(to_bits(res), want_res)
    assert rel_fro(x, torch.from_numpy(y64).cuda()) <= (1e-3 if dt == 0 else 4e-3)


def test_fused_add_rmsnorm_op_and_layer(cabi, oracle):
    """the torch op keeps the reference schema (sgl-kernel/csrc/common_extension.cc:69-70) and semantics (in place)"""
    from sglang_awq_b200 import sgl_kernel
    from sglang_awq_b200.srt.layers.layernorm import RMSNorm

    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(4, 4096, generator=g, device="cuda").half()
    res = torch.randn(4, 4096, generator=g, device="cuda").half()
    norm = RMSNorm(4096, eps=1e-6, params_dtype=torch.float16).cuda()
    norm.weight.data.copy_(torch.rand(4096, generator=g, device="cuda") + 0.5)
    want, want_res = norm.forward_native(x.clone(), res.clone())
    x1, r1 = x.clone(), res.clone()
    torch.ops.sgl_kernel.fused_add_rmsnorm.default(x1, r1, norm.weight.data, 1e-6, True)
    assert torch.equal(r1, want_res) and rel_fro(x1, want) <= 2e-4
    x2, r2 = norm(x.clone(), res.clone())
    assert torch.equal(x2, x1) and torch.equal(r2, r1)
    y = norm(x.clone())
    assert rel_fro(y, norm.forward_native(x.clone())) <= 2e-4
    with pytest.raises(RuntimeError):
        sgl_kernel.fused_add_rmsnorm(x1, r1[:, :128], norm.weight.data)


def test_epilogue_argument_checks(cabi):
    L = cabi.lib()
    x = torch.zeros(2, 256, dtype=torch.float16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.awq_b200_residual_rmsnorm(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), 2, 250, 1e-6, 0, st) != 0
    assert L.awq_b200_residual_rmsnorm(x.data_ptr(), x.data_ptr(), None, x.data_ptr(), 2, 256, 1e-6, 0, st) != 0
    assert L.awq_b200_silu_and_mul(x.data_ptr(), x.data_ptr(), 2, 100, 0, st) != 0
    assert L.awq_b200_residual_rmsnorm(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), 0, 256, 1e-6, 0, st) == 0
    t = torch.zeros(2144 * 4, dtype=torch.int32, device="cuda")
    ws = cabi.workspace(x.device)
    assert L.awq_b200_linear_silu_mul(x.data_ptr(), t.data_ptr(), x.data_ptr(), 2, 256, 24, 0, ws.data_ptr(), ws.numel(), st) != 0
    assert L.awq_b200_repack_gate_up(t.data_ptr(), x.data_ptr(), t.data_ptr(), t.data_ptr(), 128, 24, 0, st) != 0


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "epilogue_silu_mul_*.npz"))), ids=os.path.basename)
def test_silu_and_mul_op_vs_reference_golden(cabi, oracle, path):
    """the unfused activation op (reference schema silu_and_mul(Tensor! out, Tensor input)) against the oracle and the
    bits of the reference's forward_native (which rounds silu(g) once more)"""
    from sglang_awq_b200 import sgl_kernel

    g = np.load(path)
    dt = int(g["dtype"])
    x = from_bits(g["x"], dt)
    out = sgl_kernel.silu_and_mul(x)
    _, y64 = oracle.silu_mul(g["x"], dt)
    assert rel_fro(out, torch.from_numpy(y64).cuda()) <= (1e-3 if dt == 0 else 4e-3)
    assert rel_fro(out, from_bits(g["out"], dt)) <= (1e-3 if dt == 0 else 6e-3)
    out2 = torch.empty_like(out)
    torch.ops.sgl_kernel.silu_and_mul.default(out2, x)
    assert torch.equal(out, out2) and torch.equal(cabi.silu_and_mul(x), out)


@pytest.mark.parametrize("rows,hidden", [(1, 8192), (5, 5120), (3, 1000)])
def test_rmsnorm_without_residual(cabi, oracle, rows, hidden):
    from sglang_awq_b200 import sgl_kernel

    g = torch.Generator(device="cuda").manual_seed(hidden)
    x = torch.randn(rows, hidden, generator=g, device="cuda").half()
    w = (torch.rand(hidden, generator=g, device="cuda") + 0.5).half()
    out = sgl_kernel.rmsnorm(x, w, 1e-6)
    zero = np.zeros((rows, hidden), dtype=np.uint16)
    _, _, y64 = oracle.residual_rmsnorm(to_bits(x), zero, to_bits(w), 1e-6, 0)
    assert rel_fro(out, torch.from_numpy(y64).cuda()) <= 1e-3
