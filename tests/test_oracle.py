"""CPU: the oracle (C and numpy statements) against golden vectors produced by the REFERENCE's
own dequant functions (tests/golden/make_golden.py), plus self-consistency of the fixtures."""
import glob
import os

import numpy as np
import pytest

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "dequant_*.npz")))


def test_golden_present():
    assert len(GOLDEN) >= 8


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[8:-4] for p in GOLDEN])
def test_oracle_matches_reference_golden(oracle, path):
    d = np.load(path)
    dt = int(d["dtype"])
    got_c = oracle.dequantize(d["qweight"], d["scales"], d["qzeros"], dt)
    got_np = oracle.dequantize_np(d["qweight"], d["scales"], d["qzeros"], dt)
    assert np.array_equal(got_c, d["out"]), "C oracle differs from the reference"
    assert np.array_equal(got_np, d["out"]), "numpy oracle differs from the reference"


def _bits_to_torch(bits, dt):
    import torch

    t = torch.from_numpy(np.ascontiguousarray(bits).view(np.int16))
    return t.view(torch.float16 if dt == 0 else torch.bfloat16)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[8:-4] for p in GOLDEN])
def test_torch_port_matches_reference_golden(path):
    """oracle/torch_port.py is what bench.py times as the reference's CPU path (cpu_baseline and
    --impl reference): its dequant must reproduce the reference's bits on every golden vector."""
    import torch

    from oracle import torch_port

    d = np.load(path)
    dt = int(d["dtype"])
    got = torch_port.awq_dequantize_cpu(torch.from_numpy(d["qweight"]), _bits_to_torch(d["scales"], dt),
                                        torch.from_numpy(d["qzeros"]))
    got_bits = got.contiguous().view(torch.int16).numpy().view(np.uint16)
    assert got_bits.shape == d["out"].shape
    assert np.array_equal(got_bits, d["out"]), "torch port differs from the reference"


@pytest.mark.parametrize("dt", [0, 1])
@pytest.mark.parametrize("upcast", [False, True])
def test_torch_port_linear_matches_oracle(oracle, dt, upcast):
    """awq_linear_cpu (dequant -> matmul -> add_, awq.py:434-451) against the oracle's fp32(+) linear."""
    import torch

    from oracle import torch_port

    inp = oracle.make_inputs(384, 136, 128, dt, seed=11, M=7, bias=True, signed_scales=True)
    y = torch_port.awq_linear_cpu(_bits_to_torch(inp["x"], dt), torch.from_numpy(inp["qweight"]),
                                  _bits_to_torch(inp["scales"], dt), torch.from_numpy(inp["qzeros"]),
                                  _bits_to_torch(inp["bias"], dt),
                                  matmul_dtype=torch.float32 if upcast else None)
    ref = oracle.linear_f32(inp["x"], inp["qweight"], inp["scales"], inp["qzeros"], dt, inp["bias"])
    assert y.shape == ref.shape
    # the port rounds its result (and the bias add) to the 16-bit type, the oracle does not
    tol = 2e-3 if dt == 0 else 1.5e-2
    assert oracle.rel_fro(y.float().numpy(), ref) < tol


def test_nibble_order_kat(oracle):
    kat = np.load(os.path.join(os.path.dirname(__file__), "golden", "nibble_order_kat.npz"))["nibble_to_column"]
    assert list(kat) == list(oracle.AWQ_ORDER) == [0, 2, 4, 6, 1, 3, 5, 7]
    assert [list(oracle.AWQ_ORDER).index(j) for j in range(8)] == list(oracle.AWQ_REV)


def test_pack_unpack_roundtrip(oracle):
    rng = np.random.default_rng(0)
    vals = rng.integers(0, 16, size=(64, 72), dtype=np.uint8)
    packed = oracle.pack(vals)
    assert packed.shape == (64, 9)
    assert np.array_equal(oracle.unpack(packed), vals)
    assert np.array_equal(oracle.unpack_np(packed), vals.astype(np.int32))


@pytest.mark.parametrize("dt", [0, 1])
def test_half_conversions_exhaustive(oracle, dt):
    """RN conversion helpers of the C oracle agree with numpy on every 16-bit pattern and on
    products that land in subnormals / ties."""
    L = oracle.lib()
    bits = np.arange(65536, dtype=np.uint16)
    f = oracle.bits_to_f32(bits, dt)
    back = oracle.f32_to_bits(f, dt)
    finite = np.isfinite(f)
    assert np.array_equal(back[finite], bits[finite])
    if dt == 0:
        for b in [0x0001, 0x03FF, 0x0400, 0x7BFF, 0x8001, 0xFBFF, 0x3C00, 0x0000, 0x8000]:
            assert L.awq_oracle_float_to_half(L.awq_oracle_half_to_float(b)) == b
        # ties and subnormal rounding
        rng = np.random.default_rng(1)
        v = (rng.standard_normal(20000).astype(np.float32) * np.float32(1e-6)).astype(np.float32)
        got = np.array([L.awq_oracle_float_to_half(float(x)) for x in v], dtype=np.uint16)
        assert np.array_equal(got, oracle.f32_to_bits(v, 0))
    else:
        rng = np.random.default_rng(2)
        v = rng.standard_normal(20000).astype(np.float32)
        got = np.array([L.awq_oracle_float_to_bf16(float(x)) for x in v], dtype=np.uint16)
        assert np.array_equal(got, oracle.f32_to_bits(v, 1))


@pytest.mark.parametrize("dt", [0, 1])
def test_linear_oracle_matches_numpy(oracle, dt):
    inp = oracle.make_inputs(256, 64, 128, dt, seed=3, M=5, bias=True, signed_scales=True)
    y = oracle.linear_f32(inp["x"], inp["qweight"], inp["scales"], inp["qzeros"], dt, inp["bias"])
    w = oracle.bits_to_f32(oracle.dequantize(inp["qweight"], inp["scales"], inp["qzeros"], dt), dt).astype(np.float64)
    ref = oracle.bits_to_f32(inp["x"], dt).astype(np.float64) @ w + oracle.bits_to_f32(inp["bias"], dt)
    assert oracle.rel_fro(y, ref) < 1e-6


def test_repack_layout_roundtrip(oracle):
    """The tile layout is a permutation: decoding it with the spec'd index math gives back the
    logical weights / zeros / scales (including an N that is not a multiple of 128)."""
    K, N = 256, 200
    inp = oracle.make_inputs(K, N, 128, 0, seed=4)
    tiles = oracle.repack(inp["qweight"], inp["scales"], inp["qzeros"])
    W = oracle.unpack(inp["qweight"])          # [K, N]
    Z = oracle.unpack(inp["qzeros"])           # [K/128, N]
    NT, KT = (N + 127) // 128, K // 128
    assert tiles.shape == (NT * KT, 8576)
    for nt in range(NT):
        for kt in range(KT):
            rec = tiles[nt * KT + kt]
            wt = rec[:8192].view(np.uint32)
            for r in range(128):
                n = 128 * nt + r
                meta = rec[8192 + 48 * (r >> 4):8192 + 48 * (r >> 4) + 48]
                j = r & 15
                sv = meta[:32].view(np.uint16)[((j & 7) << 1) | (j >> 3)]
                zn = meta[32 + (((j & 7) << 1) | (j >> 3))]
                if n >= N:
                    assert sv == 0 and zn == 0 and not wt[r * 16:(r + 1) * 16].any()
                    continue
                assert sv == inp["scales"][kt, n]
                assert zn == Z[kt, n]
                for c in range(4):
                    for sp in range(4):
                        word = int(wt[r * 16 + 4 * (c ^ ((r >> 1) & 3)) + sp])
                        for p in range(8):
                            k = 128 * kt + 8 * (4 * c + sp) + oracle.AWQ_ORDER[p]
                            assert (word >> (4 * p)) & 0xF == W[k, n]


# ---- epilogues either side of the linears: numpy oracle pinned to the reference's own forward_native ----
import glob  # noqa: E402
import os  # noqa: E402

_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(_GOLD, "epilogue_rmsnorm_*.npz"))), ids=os.path.basename)
def test_residual_rmsnorm_oracle_matches_reference_golden(oracle, path):
    """RMSNorm.forward_native (layernorm.py:178-222) executed by tests/golden/make_golden_epilogues.py: the residual
    output is one rounding of an exact fp32 sum (bit-exact); the normalised output differs from torch's fp32
    evaluation only by the order of the mean's additions (<= 1 ulp of T)."""
    g = np.load(path)
    dt = int(g["dtype"])
    out, res, _ = oracle.residual_rmsnorm(g["x"], g["residual"], g["weight"], float(g["eps"]), dt)
    assert np.array_equal(res, g["residual_out"])
    a, b = oracle.bits_to_f32(out, dt), oracle.bits_to_f32(g["out"], dt)
    assert oracle.rel_fro(a, b) < (2e-4 if dt == 0 else 2e-3)
    assert np.mean(out != g["out"]) < 0.02  # a handful of ties broken the other way, no systematic difference


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(_GOLD, "epilogue_silu_mul_*.npz"))), ids=os.path.basename)
def test_silu_mul_oracle_matches_reference_golden(oracle, path):
    """SiluAndMul.forward_native (activation.py:68-70) rounds silu(g) to T before the product; the oracle (like the
    CUDA op) rounds once: within one rounding of T of each other."""
    g = np.load(path)
    dt = int(g["dtype"])
    out, _ = oracle.silu_mul(g["x"], dt)
    assert oracle.rel_fro(oracle.bits_to_f32(out, dt), oracle.bits_to_f32(g["out"], dt)) < (6e-4 if dt == 0 else 5e-3)


def test_gate_up_interleave_perm_is_a_permutation(oracle):
    for inter in (8, 64, 200, 1024):
        p = oracle.gate_up_interleave_perm(inter)
        assert sorted(p.tolist()) == list(range(2 * inter))
        assert p[0] == 0 and p[8] == inter and (inter < 16 or p[16] == 8)
    inp = oracle.make_inputs(128, 64, 128, 0, seed=9)
    qw, sc, qz = oracle.permute_columns(inp["qweight"], inp["scales"], inp["qzeros"], np.arange(64))
    assert np.array_equal(qw, inp["qweight"]) and np.array_equal(qz, inp["qzeros"]) and np.array_equal(sc, inp["scales"])
