"""-m gpu: kernel 1 (awq_b200_dequantize) and kernel 4 (repack) through the C ABI, bit-exact
against the reference golden vectors and the CPU oracle."""
import glob
import os

import numpy as np
import pytest
import torch

from gpu_util import TDT, big_inputs, dev_inputs, from_bits, to_bits

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "dequant_*.npz")))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[8:-4] for p in GOLDEN])
def test_dequant_matches_reference_golden(cabi, path):
    d = np.load(path)
    dt = int(d["dtype"])
    out = cabi.dequantize(torch.from_numpy(d["qweight"]).cuda(), from_bits(d["scales"], dt),
                          torch.from_numpy(d["qzeros"]).cuda())
    assert np.array_equal(to_bits(out), d["out"])


# the reference's own test grid (sgl-kernel/tests/test_awq_dequant.py:68-73), trimmed to what
# the CPU oracle finishes in seconds; group_size = K as there, plus G = 128 and G = 32
@pytest.mark.parametrize("rows", [128, 256, 512, 1024, 1536, 3584])
@pytest.mark.parametrize("cols", [16, 32, 64, 72, 128, 448, 576])
@pytest.mark.parametrize("dt", [0, 1])
def test_dequant_reference_grid(cabi, oracle, rows, cols, dt):
    for G in (rows, 128, 32):
        d = dev_inputs(oracle, rows, cols * 8, G, dt, seed=rows + cols, signed_scales=True)
        out = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"])
        ref = oracle.dequantize(d["np"]["qweight"], d["np"]["scales"], d["np"]["qzeros"], dt)
        assert np.array_equal(to_bits(out), ref), (rows, cols, G, dt)


@pytest.mark.parametrize("K,N", [(4096, 11008), (18944, 4736 * 8), (8192, 57344)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_dequant_full_size_row_slices(cabi, oracle, K, N, dtype):
    """Full-size shapes: the oracle checks random 128-row slices bit-exactly."""
    dt = 0 if dtype == torch.float16 else 1
    d = big_inputs(K, N, 128, dtype, seed=K + N)
    out = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"])
    assert out.shape == (K, N) and out.dtype == dtype
    rng = np.random.default_rng(0)
    for g in list(rng.integers(0, K // 128, size=3)) + [0, K // 128 - 1]:
        sl = slice(128 * g, 128 * (g + 1))
        ref = oracle.dequantize(d["qweight"][sl].cpu().numpy(), to_bits(d["scales"][g:g + 1]),
                                d["qzeros"][g:g + 1].cpu().numpy(), dt)
        assert np.array_equal(to_bits(out[sl]), ref)


def test_dequant_edge_cases(cabi, oracle):
    # empty
    q = torch.zeros(0, 16, dtype=torch.int32, device="cuda")
    s = torch.zeros(1, 128, dtype=torch.float16, device="cuda")
    assert cabi.dequantize(q, s, torch.zeros(1, 16, dtype=torch.int32, device="cuda")).shape == (0, 128)
    # w == z with negative scale gives -0, NaN/Inf scales propagate
    K, N = 128, 64
    qw = torch.full((K, N // 8), 0x33333333, dtype=torch.int32, device="cuda")
    qz = torch.full((1, N // 8), 0x33333333, dtype=torch.int32, device="cuda")
    sc = torch.full((1, N), -0.5, dtype=torch.float16, device="cuda")
    sc[0, 1] = float("inf")
    sc[0, 2] = float("nan")
    out = cabi.dequantize(qw, sc, qz)
    ref = oracle.dequantize(qw.cpu().numpy(), to_bits(sc), qz.cpu().numpy(), 0)
    got = to_bits(out)
    nan = np.isnan(ref.view(np.float16))
    assert np.array_equal(got[~nan], ref[~nan]) and np.isnan(got.view(np.float16)[nan]).all()
    assert got[0, 0] == 0x8000  # -0
    # rows not a multiple of the 64-row block, odd group size
    d = dev_inputs(oracle, 200, 72, 40, 0, seed=9)
    out = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"])
    assert np.array_equal(to_bits(out), oracle.dequantize(d["np"]["qweight"], d["np"]["scales"], d["np"]["qzeros"], 0))


def test_dequant_argument_errors(cabi):
    L = cabi.lib()
    q = torch.zeros(128, 16, dtype=torch.int32, device="cuda")
    s = torch.ones(1, 128, dtype=torch.float16, device="cuda")
    o = torch.empty(128, 128, dtype=torch.float16, device="cuda")
    assert L.awq_b200_dequantize(q.data_ptr(), s.data_ptr(), q.data_ptr(), o.data_ptr(), 128, 100, 1, 0, None) != 0
    assert L.awq_b200_dequantize(q.data_ptr(), s.data_ptr(), q.data_ptr(), o.data_ptr(), 128, 128, 3, 0, None) != 0
    assert L.awq_b200_dequantize(q.data_ptr(), s.data_ptr(), q.data_ptr(), o.data_ptr(), 128, 128, 1, 7, None) != 0
    assert L.awq_b200_dequantize(None, s.data_ptr(), q.data_ptr(), o.data_ptr(), 128, 128, 1, 0, None) != 0


def test_torch_op_matches_c_abi(cabi, oracle):
    from sglang_awq_b200 import sgl_kernel

    d = dev_inputs(oracle, 512, 1024, 128, 0, seed=5)
    a = sgl_kernel.awq_dequantize(d["qweight"], d["scales"], d["qzeros"])
    b = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"])
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    with pytest.raises(RuntimeError):
        sgl_kernel.awq_dequantize(d["qweight"].t(), d["scales"], d["qzeros"])  # non-contiguous
    with pytest.raises(RuntimeError):
        sgl_kernel.awq_dequantize(d["qweight"].float(), d["scales"], d["qzeros"])  # wrong dtype


@pytest.mark.parametrize("K,N", [(128, 128), (256, 200), (512, 576), (1024, 1024)])
@pytest.mark.parametrize("dt", [0, 1])
def test_repack_bit_exact_vs_oracle(cabi, oracle, K, N, dt):
    d = dev_inputs(oracle, K, N, 128, dt, seed=K * 7 + N, signed_scales=True)
    tiles = cabi.repack(d["qweight"], d["scales"], d["qzeros"])
    ref = oracle.repack(d["np"]["qweight"], d["np"]["scales"], d["np"]["qzeros"])
    assert tiles.shape == (ref.shape[0], 2144) and tiles.dtype == torch.int32
    assert np.array_equal(tiles.cpu().numpy().view(np.uint8), ref)


@pytest.mark.parametrize("K,N", [(256, 200), (4096, 11008), (14336, 4096)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_tiled_dequant_equals_reference_layout_dequant(cabi, K, N, dtype):
    """repack -> dequantize_tiled must give the same bits as dequantize on the originals
    (full-size property check: the repack loses nothing)."""
    d = big_inputs(K, N, 128, dtype, seed=3)
    tiles = cabi.repack(d["qweight"], d["scales"], d["qzeros"])
    a = cabi.dequantize_tiled(tiles, K, N, dtype)
    b = cabi.dequantize(d["qweight"], d["scales"], d["qzeros"])
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))


def test_repack_rejects_other_group_sizes(cabi):
    q = torch.zeros(192, 16, dtype=torch.int32, device="cuda")
    s = torch.ones(3, 128, dtype=torch.float16, device="cuda")
    with pytest.raises(RuntimeError):
        cabi.repack(q, s, torch.zeros(3, 16, dtype=torch.int32, device="cuda"))
