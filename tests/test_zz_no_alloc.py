"""-m gpu: awq_b200_gemv_static, the allocation-free decode entry point (static device scratch inside the library;
no counterpart in the reference, include/awq_b200.h), against the existing entry point awq_b200_gemv on the same
inputs.  Collected last on purpose (file name).

The call is made through the C ABI with every buffer (x, tiles, bias, y) allocated BEFORE it; between the two
calls in a row nothing is allocated or cleared by the test, so the second call runs on whatever the first left in
the shared scratch."""
import pytest
import torch

from gpu_util import big_inputs, rel_fro

pytestmark = pytest.mark.gpu

DT_F16 = 0
# ragged N / single tile, a split-heavy row-parallel shard, Llama-3-8B qkv-sized, Llama-3-70B gate_up (TP=1)
SHAPES = [(128, 8), (1152, 264), (3584, 8192), (4096, 6144), (8192, 57344)]


def _err(cabi):
    return cabi.lib().awq_b200_last_error().decode()


def _problem(cabi, K, N, M, seed):
    d = big_inputs(K, N, 128, torch.float16, M=M, seed=seed)
    tiles = cabi.repack(d["qweight"], d["scales"], d["qzeros"])
    g = torch.Generator(device="cuda").manual_seed(seed + 1)
    bias = torch.randn(N, generator=g, device="cuda").to(torch.float16)
    return d["x"], tiles, bias, d


def _static(cabi, x, tiles, bias, y, N):
    M, K = x.shape
    st = torch.cuda.current_stream().cuda_stream
    rc = cabi.lib().awq_b200_gemv_static(x.data_ptr(), tiles.data_ptr(), bias.data_ptr() if bias is not None else None,
                                        y.data_ptr(), M, K, N, DT_F16, st)
    return rc


@pytest.mark.parametrize("K,N", SHAPES)
@pytest.mark.parametrize("M", [1, 4, 16, 32])
def test_static_scratch_matches_existing_entry_point_twice_in_a_row(cabi, K, N, M):
    x, tiles, bias, d = _problem(cabi, K, N, M, seed=K + N + M)
    ref = cabi.gemv(x, tiles, N, bias=bias)                # existing entry point, shape router, caller workspace
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"]).float()
    exact = x.float() @ w + bias.float()
    y1 = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
    y2 = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    assert _static(cabi, x, tiles, bias, y1, N) == 0, _err(cabi)
    assert _static(cabi, x, tiles, bias, y2, N) == 0, _err(cabi)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for y in (y1, y2):
        assert not torch.isnan(y).any()
        # the tolerance of the decode-kernel parity tests (tests/test_gpu_decode_kernels.py)
        assert rel_fro(y, ref.float()) <= 1e-3, f"K={K} N={N} M={M} vs awq_b200_gemv"
        assert rel_fro(y, exact) <= 1e-3, f"K={K} N={N} M={M} vs fp32 matmul"
    assert torch.equal(y1.view(torch.int16), y2.view(torch.int16)), "second call on the used scratch differs"


def test_static_scratch_is_bit_identical_to_the_same_kernel_with_a_caller_workspace(cabi):
    L = cabi.lib()
    K, N = 3584, 8192
    for M in (2, 16, 24):
        x, tiles, bias, _ = _problem(cabi, K, N, M, seed=7 + M)
        try:
            L.awq_b200_debug_set_gemv_impl(5, 0, -1)       # awq_b200_gemv forced to the tcgen05 decode kernel
            ref = cabi.gemv(x, tiles, N, bias=bias)
        finally:
            L.awq_b200_debug_set_gemv_impl(0, 0, -1)
        y = torch.empty((M, N), dtype=torch.float16, device="cuda")
        for _ in range(2):
            y.fill_(float("nan"))
            assert _static(cabi, x, tiles, bias, y, N) == 0, _err(cabi)
            assert torch.equal(y.view(torch.int16), ref.view(torch.int16))


def test_static_scratch_alternating_shapes_and_no_bias(cabi):
    # one shape after another on the shared array: leftovers of a larger problem must not leak into a smaller one
    cases = [(8192, 57344, 16), (128, 8, 1), (3584, 8192, 32), (1152, 264, 3), (8192, 57344, 1)]
    probs = [(_problem(cabi, K, N, M, seed=11 * i + 3), N) for i, (K, N, M) in enumerate(cases)]
    refs = [cabi.gemv(x, tiles, N) for (x, tiles, _, _), N in probs]
    outs = [torch.empty_like(r) for r in refs]
    for _ in range(2):
        for ((x, tiles, _, _), N), y in zip(probs, outs):
            assert _static(cabi, x, tiles, None, y, N) == 0, _err(cabi)
    torch.cuda.synchronize()
    for y, r in zip(outs, refs):
        assert rel_fro(y, r.float()) <= 1e-3


def test_static_scratch_rejects_what_it_cannot_hold(cabi):
    L = cabi.lib()
    x = torch.zeros((32, 128), dtype=torch.float16, device="cuda")
    y = torch.zeros((32, 8), dtype=torch.float16, device="cuda")
    t = torch.zeros(16, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    # N beyond what 16 MiB of scratch holds at M = 32 / M = 16 (checked before any pointer is touched)
    assert L.awq_b200_gemv_static(x.data_ptr(), t.data_ptr(), None, y.data_ptr(), 32, 128, 131072, DT_F16, st) != 0
    assert "static maximum" in _err(cabi)
    assert L.awq_b200_gemv_static(x.data_ptr(), t.data_ptr(), None, y.data_ptr(), 16, 128, 262144, DT_F16, st) != 0
    assert "static maximum" in _err(cabi)
    # the kernel's own limits
    assert L.awq_b200_gemv_static(x.data_ptr(), t.data_ptr(), None, y.data_ptr(), 33, 128, 8, DT_F16, st) != 0
    assert "M" in _err(cabi)
    assert L.awq_b200_gemv_static(x.data_ptr(), t.data_ptr(), None, y.data_ptr(), 1, 128, 8, 1, st) != 0
    assert "fp16" in _err(cabi)
    assert L.awq_b200_gemv_static(None, t.data_ptr(), None, y.data_ptr(), 1, 128, 8, DT_F16, st) != 0
    torch.cuda.synchronize()
