"""-m gpu: ours against the REFERENCE'S OWN CUDA kernel on the same device.

oracle/_ref/ref_awq.so is sgl-kernel/csrc/gemm/awq_kernel.cu compiled unmodified for sm_100a by
oracle/build_ref.py (in the build container, where /root/reference exists; the binary travels to
the GPU box).  Bit-exact for the dequant op on the reference's full test grid
(sgl-kernel/tests/test_awq_dequant.py:68-73) at full sizes, and the fused linear against the
reference's apply sequence (awq.py:434-451: awq_dequantize -> torch.matmul -> add_)."""
import pytest
import torch

from gpu_util import big_inputs, rel_fro

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_dequant():
    from oracle import build_ref

    fn = build_ref.load()
    if fn is None:
        pytest.skip("oracle/_ref/ref_awq.so not built (python -m oracle.build_ref, needs /root/reference)")
    return fn


# the reference's own grid, untrimmed: rows x packed cols, group size = rows (as there) and 128
@pytest.mark.parametrize("rows", [3584, 18944, 128, 256, 512, 1024, 1536])
@pytest.mark.parametrize("cols", [448, 576, 4736, 16, 32, 64, 128, 72])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_dequant_bit_exact_vs_reference_cuda_kernel(cabi, ref_dequant, rows, cols, dtype):
    for G in (rows, 128):
        d = big_inputs(rows, cols * 8, G, dtype, seed=rows + cols)
        g = torch.Generator(device="cuda").manual_seed(cols)
        sign = (torch.rand(d["scales"].shape, generator=g, device="cuda") > 0.3).to(dtype) * 2 - 1
        tiny = (torch.rand(d["scales"].shape, generator=g, device="cuda") > 0.9).to(dtype) * 2.0 ** -20  # subnormal products
        scales = (d["scales"] * sign * torch.where(tiny > 0, tiny, torch.ones_like(tiny))).contiguous()
        want = ref_dequant(d["qweight"], scales, d["qzeros"])
        got = cabi.dequantize(d["qweight"], scales, d["qzeros"])
        assert got.dtype == want.dtype and got.shape == want.shape
        assert torch.equal(got.view(torch.int16), want.view(torch.int16)), (rows, cols, G, dtype)


@pytest.mark.parametrize("K,N", [(4096, 11008), (4096, 6144), (14336, 4096), (8192, 7168)])
@pytest.mark.parametrize("M", [1, 4, 16, 32, 64, 512])
def test_linear_vs_reference_apply_sequence(cabi, ref_dequant, K, N, M):
    """fused kernel vs the reference's three launches (fp16 cuBLAS GEMM over ITS dequantised weights);
    both are compared with the fp32 matmul over those weights, and with each other"""
    d = big_inputs(K, N, 128, torch.float16, M=M, seed=K + N + M)
    bias = torch.randn(N, device="cuda").half()
    w = ref_dequant(d["qweight"], d["scales"], d["qzeros"])
    ref_out = torch.matmul(d["x"], w)
    ref_out.add_(bias)
    exact = d["x"].float() @ w.float() + bias.float()
    tiles = cabi.repack(d["qweight"], d["scales"], d["qzeros"])
    ours = cabi.linear(d["x"], tiles, N, bias=bias)
    assert rel_fro(ours, exact) <= 1e-3
    assert rel_fro(ours, ref_out.float()) <= 1.5e-3  # two fp16 roundings apart
    assert rel_fro(ours, exact) <= rel_fro(ref_out, exact) * 1.5 + 1e-5, "not less accurate than the reference path"


@pytest.fixture(scope="module")
def marlin():
    from oracle import marlin_ref

    if not marlin_ref.load():
        pytest.skip("oracle/_ref/ref_marlin.so not built (python -m oracle.marlin_ref, needs /root/reference, 16 min)")
    return marlin_ref


@pytest.mark.parametrize("K,N", [(512, 256), (4096, 11008), (4096, 6144), (8192, 7168)])
@pytest.mark.parametrize("M", [1, 16, 64, 512])
def test_linear_vs_reference_marlin(cabi, marlin, K, N, M):
    """ours against the reference's AWQ-Marlin path on the same device (what an AWQ checkpoint resolves to by default,
    awq.py:267-289): awq_marlin_repack + the python permutes (restated in oracle/marlin_ref.py) + gptq_marlin_gemm, all
    compiled from the reference's sources.  Marlin's own pin is mean|d|/mean|ref| < 0.04
    (sgl-kernel/tests/test_marlin_gemm.py:119-127); both paths are held to the 1e-3 Frobenius bar here."""
    d = big_inputs(K, N, 128, torch.float16, M=M, seed=K + N + M)
    w = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"])
    exact = d["x"].float() @ w.float()
    ref_out = marlin.linear(d["x"], marlin.prepare(d["qweight"], d["scales"], d["qzeros"]))
    ours = cabi.linear(d["x"], cabi.repack(d["qweight"], d["scales"], d["qzeros"]), N)
    assert rel_fro(ref_out.float(), exact) <= 1e-3, "the restated Marlin permutes do not reproduce the reference path"
    assert rel_fro(ours, exact) <= 1e-3
    assert rel_fro(ours, ref_out.float()) <= 1.5e-3
