"""-m gpu: gemv_sub_kernel, the decode kernel of group-64 / group-32 checkpoints behind awq_b200_linear_g
(sglang_awq_b200/csrc/awq_gemv_sub.cu), against
  * the fp32 matmul over the bit-exact dequantised weights (the reference's AWQLinearMethod.apply semantics,
    python/sglang/srt/layers/quantization/awq.py:434-451, at the group size the checkpoint names), and
  * the tcgen05 kernel on the same records (the route those batches took before; diagnostic knob).
Tolerance: relative Frobenius error <= 1e-3 (BASELINE.json north_star), as in the other decode-kernel tests."""
import pytest
import torch

from gpu_util import big_inputs, dev_inputs, rel_fro

pytestmark = pytest.mark.gpu

MAX_M = {64: 4, 32: 4}   # token columns: 4 / 2; group 32 at M = 3, 4 is the per-weight variant
# KT = 8 (one record per warp), 9 (ragged split), 32 (even stream, team of 8), 40 (ragged), 64; ragged N; one slab
SHAPES = [(1024, 8), (1024, 384), (1152, 264), (4096, 4096), (5120, 1000), (8192, 2048), (3584, 8192)]


@pytest.fixture
def sub_decode(cabi):
    L = cabi.lib()
    yield lambda on: L.awq_b200_debug_set_sub_decode(int(on))
    L.awq_b200_debug_set_sub_decode(1)


def _signed(d, K, N, G, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    sign = (torch.rand(K // G, N, generator=g, device="cuda") > 0.25).to(torch.float16) * 2 - 1
    d["scales"] = (d["scales"] * sign).contiguous()
    return g


@pytest.mark.parametrize("G", [64, 32])
@pytest.mark.parametrize("K,N", SHAPES)
def test_sub_group_decode_kernel(cabi, sub_decode, G, K, N):
    for M in range(1, MAX_M[G] + 1):
        d = big_inputs(K, N, G, torch.float16, M=M, seed=K + N + M + G)
        g = _signed(d, K, N, G, seed=K + M)
        tiles = cabi.repack_g(d["qweight"], d["scales"], d["qzeros"])
        bias = torch.randn(N, generator=g, device="cuda").to(torch.float16)
        w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()   # bit-exact with the reference op
        ref = d["x"].float() @ w + bias.float()
        sub_decode(True)
        y = cabi.linear_g(d["x"], tiles, N, bias=bias)
        assert y.dtype == torch.float16 and y.shape == (M, N)
        err = rel_fro(y, ref)
        assert err <= 1e-3, f"G={G} K={K} N={N} M={M}: {err:.2e}"
        y2 = cabi.linear_g(d["x"], tiles, N, bias=bias)
        assert torch.equal(y.view(torch.int16), y2.view(torch.int16)), "the decode kernel must be deterministic"
        sub_decode(False)
        y_tc = cabi.linear_g(d["x"], tiles, N, bias=bias)
        assert rel_fro(y, y_tc.float()) <= 1e-3, f"G={G} K={K} N={N} M={M} vs the tcgen05 kernel"
        sub_decode(True)
        y_nb = cabi.linear_g(d["x"], tiles, N)
        assert rel_fro(y_nb, ref - bias.float()) <= 1e-3


@pytest.mark.parametrize("G", [64, 32])
def test_sub_group_decode_vs_oracle(cabi, oracle, G):
    # the CPU oracle (oracle/awq_oracle.py, pinned to the reference's golden vectors) on the same seeded inputs
    K, N = 1024, 640
    for M in range(1, MAX_M[G] + 1):
        d = dev_inputs(oracle, K, N, G, 0, seed=G + M, M=M, bias=True, signed_scales=True)
        tiles = cabi.repack_g(d["qweight"], d["scales"], d["qzeros"])
        ref = torch.from_numpy(oracle.linear_f32(d["np"]["x"], d["np"]["qweight"], d["np"]["scales"], d["np"]["qzeros"], 0,
                                                 d["np"]["bias"])).cuda()
        y = cabi.linear_g(d["x"], tiles, N, bias=d["bias"])
        assert rel_fro(y, ref) <= 1e-3, f"G={G} M={M}"


@pytest.mark.parametrize("G", [64, 32])
def test_sub_group_decode_separates_the_groups(cabi, G):
    # one sub-group with a huge scale and zero point 15, its neighbours tiny: a kernel that applied a record-wide
    # scale (or leaked x across sub-groups) misses by orders of magnitude
    K, N, M = 1024, 256, MAX_M[G]
    d = big_inputs(K, N, G, torch.float16, M=M, seed=99 + G)
    rows = torch.arange(K // G, device="cuda")
    big = (rows % (128 // G)) == 1
    d["scales"] = torch.where(big[:, None], torch.full_like(d["scales"], 2.0), torch.full_like(d["scales"], 1e-3)).contiguous()
    d["qzeros"] = torch.where(big[:, None], torch.full_like(d["qzeros"], -1), d["qzeros"]).contiguous()   # all nibbles 15
    tiles = cabi.repack_g(d["qweight"], d["scales"], d["qzeros"])
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
    ref = d["x"].float() @ w
    assert rel_fro(cabi.linear_g(d["x"], tiles, N), ref) <= 1e-3


def test_sub_group_decode_full_size(cabi):
    # Llama-3-70B gate_up at group 64 (BASELINE.json full size): linearity in x is the size-independent property
    K, N, G, M = 8192, 57344, 64, 4
    d = big_inputs(K, N, G, torch.float16, M=M, seed=5)
    tiles = cabi.repack_g(d["qweight"], d["scales"], d["qzeros"])
    x = d["x"]
    y = cabi.linear_g(x, tiles, N).float()
    x2 = torch.cat([x[1:2] + x[0:1], x[0:1] * 2, x[2:3] - x[3:4], x[3:4]]).contiguous()
    y2 = cabi.linear_g(x2, tiles, N).float()
    want = torch.cat([y[1:2] + y[0:1], y[0:1] * 2, y[2:3] - y[3:4], y[3:4]])
    assert rel_fro(y2, want) <= 2e-3
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
    assert rel_fro(y, x.float() @ w) <= 1e-3


# ---- per-weight variant (gemv_sub_pw_kernel): bf16 up to M = 8, fp16 M = 5 .. 8, and everything when forced ----
@pytest.mark.parametrize("G", [64, 32])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("K,N", [(1024, 8), (1152, 264), (4096, 4096), (5120, 1000), (3584, 8192)])
def test_sub_group_decode_per_weight_kernel(cabi, sub_decode, G, dtype, K, N):
    for M in range(1, 9):
        d = big_inputs(K, N, G, dtype, M=M, seed=K + N + M + G)
        g = torch.Generator(device="cuda").manual_seed(K + M)
        sign = (torch.rand(K // G, N, generator=g, device="cuda") > 0.25).to(dtype) * 2 - 1
        d["scales"] = (d["scales"] * sign).contiguous()
        tiles = cabi.repack_g(d["qweight"], d["scales"], d["qzeros"])
        bias = torch.randn(N, generator=g, device="cuda").to(dtype)
        w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
        ref = d["x"].float() @ w + bias.float()
        sub_decode(2)      # the per-weight kernel for every batch it can take
        y = cabi.linear_g(d["x"], tiles, N, bias=bias)
        y2 = cabi.linear_g(d["x"], tiles, N, bias=bias)
        sub_decode(0)      # the tcgen05 kernel
        y_tc = cabi.linear_g(d["x"], tiles, N, bias=bias)
        sub_decode(1)      # default routing
        y_def = cabi.linear_g(d["x"], tiles, N, bias=bias)
        assert torch.equal(y.view(torch.int16), y2.view(torch.int16)), "deterministic"
        if dtype == torch.float16:
            tol = 1e-3
            assert rel_fro(y, ref) <= tol, f"G={G} K={K} N={N} M={M}: {rel_fro(y, ref):.2e}"
        else:                       # bf16 outputs carry 8 significant bits: the bar of tests/test_gpu_decode_kernels.py
            tol = 4e-3
            assert rel_fro(y, ref.to(torch.bfloat16)) <= 1e-3, f"G={G} K={K} N={N} M={M}"
            assert rel_fro(y, ref) <= tol
        assert rel_fro(y, y_tc.float()) <= tol and rel_fro(y_def, ref) <= tol
