"""-m gpu: every decode kernel behind awq_b200_gemv, forced one by one through the diagnostic
knob, against the fp32 matmul over the bit-exact dequantised weights (the reference's
AWQLinearMethod.apply semantics, python/sglang/srt/layers/quantization/awq.py:434-451).

impl 1 = one-slab-per-CTA, 2 = legacy stream-K ring, 3 = persistent team kernel,
impl 4 = cluster / DSMEM split-K kernel, 5 = tcgen05 decode kernel (fp16), 0 = the shape router."""
import pytest
import torch

from gpu_util import big_inputs, rel_fro

pytestmark = pytest.mark.gpu

IMPLS = {1: "direct", 2: "legacy", 3: "persist", 4: "cluster", 5: "tcgen05", 0: "router"}
# (K, N): ragged N, a single tile, K/128 not divisible by 8 teams x depth (generic stream loop),
# row-parallel TP shards of Llama-3-70B (1024 x 8192, 3584 x 8192), Qwen2.5-32B (5120 x 7168)
SHAPES = [(128, 8), (512, 72), (1024, 384), (1152, 256), (1024, 8192), (3584, 8192), (5120, 7168), (4096, 4096)]


@pytest.fixture
def force(cabi):
    L = cabi.lib()
    yield lambda impl, split=0, variant=-1: L.awq_b200_debug_set_gemv_impl(impl, split, variant)
    L.awq_b200_debug_set_gemv_impl(0, 0, -1)


def _case(cabi, K, N, M, dtype, seed):
    d = big_inputs(K, N, 128, dtype, M=M, seed=seed)
    g = torch.Generator(device="cuda").manual_seed(seed)
    # signed scales: the sign of (w - z) * s must survive the group-sum formulation
    sign = (torch.rand(K // 128, N, generator=g, device="cuda") > 0.25).to(dtype) * 2 - 1
    d["scales"] = (d["scales"] * sign).contiguous()
    tiles = cabi.repack(d["qweight"], d["scales"], d["qzeros"])
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
    bias = torch.randn(N, generator=g, device="cuda").to(dtype)
    ref = d["x"].float() @ w + bias.float()
    return d, tiles, bias, ref


@pytest.mark.parametrize("impl", sorted(IMPLS))
@pytest.mark.parametrize("K,N", SHAPES)
def test_every_decode_kernel_fp16(cabi, force, impl, K, N):
    for M in (1, 2, 3, 4, 5, 8, 9, 13, 16) + ((17, 24, 32) if impl == 5 else ()):
        d, tiles, bias, ref = _case(cabi, K, N, M, torch.float16, seed=K + N + M)
        force(impl)
        y = cabi.gemv(d["x"], tiles, N, bias=bias)
        assert y.dtype == torch.float16 and y.shape == (M, N)
        err = rel_fro(y, ref)
        assert err <= 1e-3, f"{IMPLS[impl]} K={K} N={N} M={M}: {err:.2e}"
        y2 = cabi.gemv(d["x"], tiles, N, bias=bias)
        assert torch.equal(y.view(torch.int16), y2.view(torch.int16)), "decode kernels must be deterministic"


@pytest.mark.parametrize("impl", sorted(IMPLS))
def test_every_decode_kernel_bf16(cabi, force, impl):
    for K, N in ((1024, 384), (4096, 4096), (3584, 8192)):
        for M in (1, 4, 8, 16):
            d, tiles, bias, ref = _case(cabi, K, N, M, torch.bfloat16, seed=K + M)
            force(impl)
            y = cabi.gemv(d["x"], tiles, N, bias=bias)
            assert y.dtype == torch.bfloat16
            assert rel_fro(y, ref.to(torch.bfloat16)) <= 1e-3
            assert rel_fro(y, ref) <= 4e-3


@pytest.mark.parametrize("impl,split", [(4, 1), (4, 2), (4, 4), (4, 8), (3, 8), (3, 16), (3, 32), (5, 1), (5, 3), (5, 37), (5, 148)])
def test_forced_splits(cabi, force, impl, split):
    """cluster sizes / team sizes the heuristics would not pick for this shape"""
    K, N = 4096, 2048
    for M in (1, 7, 16):
        d, tiles, bias, ref = _case(cabi, K, N, M, torch.float16, seed=split + M)
        force(impl, split)
        assert rel_fro(cabi.gemv(d["x"], tiles, N, bias=bias), ref) <= 1e-3


@pytest.mark.parametrize("impl", [1, 3, 4])
@pytest.mark.parametrize("variant", [0, 1, 3])
def test_every_instantiation(cabi, force, impl, variant):
    K, N = 2048, 1536
    for M in (2, 12):
        d, tiles, bias, ref = _case(cabi, K, N, M, torch.float16, seed=variant + M)
        force(impl, 0, variant)
        assert rel_fro(cabi.gemv(d["x"], tiles, N, bias=bias), ref) <= 1e-3


def test_tcgen05_decode_long_streams(cabi, force):
    """every instantiation of the tcgen05 decode kernel per M range (convert groups x records in flight
    x operand buffers per group) over streams long enough to wrap the operand buffers, the activation
    stages and the accumulator ring many times, with split n-tiles (2 CTAs: 100+ tiles per CTA) and
    without"""
    for K, N in ((2048, 1536), (4096, 4096), (1152, 264), (8192, 2048), (128, 136)):
        for M in (1, 3, 8, 12, 16, 17, 29, 32):
            d, tiles, bias, ref = _case(cabi, K, N, M, torch.float16, seed=M + K)
            for variant in (0, 1, 2, 3):
                for ctas in (0, 2, 5):
                    force(5, ctas, variant)
                    y = cabi.gemv(d["x"], tiles, N, bias=bias)
                    assert rel_fro(y, ref) <= 1e-3, f"K={K} N={N} M={M} ctas={ctas} variant={variant}"
                    assert torch.equal(y.view(torch.int16), cabi.gemv(d["x"], tiles, N, bias=bias).view(torch.int16))


def test_tcgen05_decode_is_exact_on_integer_data(cabi, force):
    """With integer-valued x, unit scales and sums below 2^24 every product and partial sum is exact
    in fp32: the masked-nibble operands (w * 2^-24, 16 w * 2^-24 folded back by scale-input-d) and the
    -z * Sum(x) correction must reproduce the integer result bit for bit."""
    K, N, M = 1024, 512, 16
    g = torch.Generator(device="cuda").manual_seed(3)
    qw = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    qz = torch.randint(-(2**31), 2**31, (K // 128, N // 8), generator=g, device="cuda", dtype=torch.int64).to(torch.int32)
    sc = torch.ones(K // 128, N, device="cuda", dtype=torch.float16)
    x = torch.randint(-1, 2, (M, K), generator=g, device="cuda").half()
    tiles = cabi.repack(qw, sc, qz)
    ref = x.double() @ cabi.dequantize(qw, sc, qz).double()
    assert ref.abs().max() < 2048  # representable in fp16
    for variant in (0, 1, 2, 3):
        force(5, 0, variant)
        y = cabi.gemv(x, tiles, N)
        assert torch.equal(y.double(), ref), f"variant {variant}"


def test_extreme_values_survive_the_group_sum_form(cabi, force):
    """The fp16 fast path computes sum(w x) - z sum(x) with w as fp16 subnormals: large and tiny
    activations, all-zero weights (w == z) and zero scales must still meet the bound."""
    K, N, M = 1024, 256, 4
    d, tiles, bias, ref = _case(cabi, K, N, M, torch.float16, seed=5)
    for scale in (2.0 ** -10, 1.0, 2.0 ** 6):
        x = (d["x"].float() * scale).half()
        w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
        r = x.float() @ w
        for impl in (1, 3, 4, 5):
            force(impl)
            assert rel_fro(cabi.gemv(x, tiles, N), r) <= 1e-3
    # w == z everywhere -> exactly zero output whatever x is
    qz = d["qzeros"].clone()
    qw = qz.repeat_interleave(128, dim=0).contiguous()
    t0 = cabi.repack(qw, d["scales"], qz)
    for impl in (1, 3, 4, 5):
        force(impl)
        y = cabi.gemv(d["x"], t0, N).float()
        assert y.abs().max().item() <= 2e-3 * d["x"].float().abs().sum(dim=1).max().item() * d["scales"].float().abs().max().item()


def test_router_follows_the_measured_table(cabi):
    """awq_b200_gemv_route picks, for shapes of the measured matrix, the kernel measured fastest
    (profiles/r02_kernel_matrix.json): the tcgen05 kernel for 9..32-token batches of the big
    projections, the persistent kernel for 70B down_proj at batch 1-2 (the hand-tuned router sent
    those to the cluster kernel: 42 us instead of 27), and only kernels that can take the problem."""
    import json
    import os
    L = cabi.lib()
    rows = json.load(open(os.path.join(os.path.dirname(__file__), "..", "profiles", "r02_kernel_matrix.json")))["rows"]
    ids = {"direct": 1, "legacy": 2, "persist": 3, "cluster": 4, "tc": 5}
    assert L.awq_b200_gemv_route(1, 28672, 8192, 0) == 3
    assert L.awq_b200_gemv_route(2, 28672, 8192, 0) == 3
    assert L.awq_b200_gemv_route(16, 8192, 57344, 0) == 5
    assert L.awq_b200_gemv_route(32, 8192, 57344, 0) == 5
    for r in rows:
        got = L.awq_b200_gemv_route(r["M"], r["K"], r["N"], 0)
        assert got == ids[r["best"]], (r["shape"], r["M"], got, r["best"])
        # bf16 never reaches the fp16-only tensor-core kernel, and always reaches something that fits
        got_bf16 = L.awq_b200_gemv_route(r["M"], r["K"], r["N"], 1)
        assert got_bf16 != 5 and L.awq_b200_debug_gemv_impl_fits(got_bf16, r["M"], r["K"], r["N"], 1) == 1
    # unmeasured shapes and in-between batches still land on a kernel that fits
    for M, K, N in ((3, 5120, 13824), (7, 1536, 1536), (11, 16384, 2048), (27, 7168, 4608), (40, 4096, 4096)):
        got = L.awq_b200_gemv_route(M, K, N, 0)
        assert got in (1, 2, 3, 4, 5) and L.awq_b200_debug_gemv_impl_fits(got, M, K, N, 0) == 1
