"""Python face of the CPU oracle (oracle/awq_oracle.c) + a numpy restatement of the same math.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (sglang_awq_b200/) never imports it.

Reference behaviour restated (paths relative to /root/reference):
  * dequant arithmetic ..... sgl-kernel/csrc/gemm/awq_kernel.cu:126-184
  * nibble order ........... python/sglang/srt/layers/quantization/awq_triton.py:342-368
  * linear op .............. python/sglang/srt/layers/quantization/awq.py:434-451
Pinned against the reference's own functions by tests/golden/make_golden.py -> tests/golden/*.npz.

All 16-bit float tensors cross this boundary as numpy uint16 bit patterns, so no test depends
on how numpy/torch round.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libawq_oracle.so")

F16, BF16 = 0, 1
AWQ_ORDER = (0, 2, 4, 6, 1, 3, 5, 7)  # nibble p holds logical column ORDER[p]
AWQ_REV = (0, 4, 1, 5, 2, 6, 3, 7)  # logical column j lives at nibble REV[j]

_lib = None


def build(force: bool = False) -> str:
    """Compile oracle/awq_oracle.c with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "awq_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        p = ctypes.c_void_p
        i = ctypes.c_int
        L.awq_oracle_dequantize.argtypes = [p, p, p, p, i, i, i, i]
        L.awq_oracle_dequantize.restype = None
        L.awq_oracle_linear_f32.argtypes = [p, p, p, p, p, p, i, i, i, i, i]
        L.awq_oracle_linear_f32.restype = None
        L.awq_oracle_pack.argtypes = [p, p, i, i]
        L.awq_oracle_pack.restype = None
        L.awq_oracle_unpack.argtypes = [p, p, i, i]
        L.awq_oracle_unpack.restype = None
        L.awq_oracle_repack.argtypes = [p, p, p, p, i, i]
        L.awq_oracle_repack.restype = None
        L.awq_oracle_float_to_half.argtypes = [ctypes.c_float]
        L.awq_oracle_float_to_half.restype = ctypes.c_uint16
        L.awq_oracle_half_to_float.argtypes = [ctypes.c_uint16]
        L.awq_oracle_half_to_float.restype = ctypes.c_float
        L.awq_oracle_float_to_bf16.argtypes = [ctypes.c_float]
        L.awq_oracle_float_to_bf16.restype = ctypes.c_uint16
        _lib = L
    return _lib


def _c(a: np.ndarray, dtype) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


# ---------------------------------------------------------------------------------------
# C oracle entry points
# ---------------------------------------------------------------------------------------


def dequantize(qweight: np.ndarray, scales_bits: np.ndarray, qzeros: np.ndarray, dtype: int) -> np.ndarray:
    """uint16 bit patterns of the reference's dequantised [K, N] matrix."""
    qweight = _c(qweight, np.int32)
    qzeros = _c(qzeros, np.int32)
    scales_bits = _c(scales_bits, np.uint16)
    K, NP = qweight.shape
    N = NP * 8
    assert scales_bits.shape[1] == N and qzeros.shape == (scales_bits.shape[0], NP)
    out = np.empty((K, N), dtype=np.uint16)
    lib().awq_oracle_dequantize(_ptr(qweight), _ptr(scales_bits), _ptr(qzeros), _ptr(out), K, N,
                                scales_bits.shape[0], dtype)
    return out


def linear_f32(x_bits, qweight, scales_bits, qzeros, dtype: int, bias_bits=None) -> np.ndarray:
    """float32 [M, N] = x @ dequant(...) (+bias), double accumulation."""
    x_bits = _c(x_bits, np.uint16)
    qweight = _c(qweight, np.int32)
    qzeros = _c(qzeros, np.int32)
    scales_bits = _c(scales_bits, np.uint16)
    M, K = x_bits.shape
    N = qweight.shape[1] * 8
    assert qweight.shape[0] == K
    if bias_bits is not None:
        bias_bits = _c(bias_bits, np.uint16)
    y = np.empty((M, N), dtype=np.float32)
    lib().awq_oracle_linear_f32(_ptr(x_bits), _ptr(qweight), _ptr(scales_bits), _ptr(qzeros),
                                _ptr(bias_bits), _ptr(y), M, K, N, scales_bits.shape[0], dtype)
    return y


def pack(vals: np.ndarray) -> np.ndarray:
    vals = _c(vals, np.uint8)
    K, N = vals.shape
    out = np.empty((K, N // 8), dtype=np.int32)
    lib().awq_oracle_pack(_ptr(vals), _ptr(out), K, N)
    return out


def unpack(packed: np.ndarray) -> np.ndarray:
    packed = _c(packed, np.int32)
    K, NP = packed.shape
    out = np.empty((K, NP * 8), dtype=np.uint8)
    lib().awq_oracle_unpack(_ptr(packed), _ptr(out), K, NP * 8)
    return out


TILE_BYTES = 8576


def repack(qweight, scales_bits, qzeros) -> np.ndarray:
    """The B200 tile layout (spec in awq_oracle.c): uint8 [T, 8576] records,
    T = ceil(N/128) * K/128, each = 8192 B weights | 8 x 48 B per-16-row metadata (scales, zeros)."""
    qweight = _c(qweight, np.int32)
    qzeros = _c(qzeros, np.int32)
    scales_bits = _c(scales_bits, np.uint16)
    K, NP = qweight.shape
    N = NP * 8
    assert K % 128 == 0 and scales_bits.shape[0] == K // 128
    T = ((N + 127) // 128) * (K // 128)
    tiles = np.empty((T, TILE_BYTES), dtype=np.uint8)
    lib().awq_oracle_repack(_ptr(qweight), _ptr(scales_bits), _ptr(qzeros), _ptr(tiles), K, N)
    return tiles


# ---------------------------------------------------------------------------------------
# numpy restatement (second, independent statement of the same arithmetic; tests check
# C == numpy == golden)
# ---------------------------------------------------------------------------------------


def bits_to_f32(bits: np.ndarray, dtype: int) -> np.ndarray:
    bits = np.asarray(bits, dtype=np.uint16)
    if dtype == F16:
        return bits.view(np.float16).astype(np.float32)
    return (bits.astype(np.uint32) << 16).view(np.float32)


def f32_to_bits(v: np.ndarray, dtype: int) -> np.ndarray:
    v = np.asarray(v, dtype=np.float32)
    if dtype == F16:
        with np.errstate(over="ignore"):
            return v.astype(np.float16).view(np.uint16)
    u = v.view(np.uint32)
    rounded = ((u + (0x7FFF + ((u >> 16) & 1))) >> 16).astype(np.uint16)
    nan = (u & 0x7FFFFFFF) > 0x7F800000
    return np.where(nan, ((u >> 16) | 0x40).astype(np.uint16), rounded)


def unpack_np(packed: np.ndarray) -> np.ndarray:
    p = np.asarray(packed, dtype=np.int32).view(np.uint32)
    cols = [(p >> np.uint32(4 * AWQ_REV[j])) & np.uint32(0xF) for j in range(8)]
    return np.stack(cols, axis=-1).reshape(p.shape[0], -1).astype(np.int32)


def dequantize_np(qweight, scales_bits, qzeros, dtype: int) -> np.ndarray:
    w = unpack_np(qweight)
    z = unpack_np(qzeros)
    K = w.shape[0]
    G = K // z.shape[0]
    s = bits_to_f32(scales_bits, dtype)
    d = (w - np.repeat(z, G, axis=0)).astype(np.float32)
    return f32_to_bits(d * np.repeat(s, G, axis=0), dtype)


def rel_fro(a: np.ndarray, ref: np.ndarray) -> float:
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    den = np.linalg.norm(ref)
    return float(np.linalg.norm(a - ref) / (den if den > 0 else 1.0))


# ---------------------------------------------------------------------------------------
# epilogues either side of the linears; pinned to tests/golden/epilogue_*.npz
# ---------------------------------------------------------------------------------------


def residual_rmsnorm(x_bits, residual_bits, weight_bits, eps: float, dtype: int):
    """RMSNorm.forward_native with a residual, python/sglang/srt/layers/layernorm.py:178-222:
    r = x + residual in fp32; residual_out = T(r); out = T(r * rsqrt(mean(r^2) + eps) * weight).
    Returns (out_bits, residual_out_bits, out_f64): the last one unrounded, for tolerance checks."""
    r = bits_to_f32(x_bits, dtype) + bits_to_f32(residual_bits, dtype)
    var = np.mean(r.astype(np.float64) ** 2, axis=-1, keepdims=True)
    y = r.astype(np.float64) / np.sqrt(var + np.float64(np.float32(eps))) * bits_to_f32(weight_bits, dtype).astype(np.float64)
    return f32_to_bits(y.astype(np.float32), dtype), f32_to_bits(r, dtype), y


def silu_mul(gate_up_bits, dtype: int):
    """SiluAndMul on [.., 2d] (gate columns first): silu(g) * u, evaluated in double from the T-rounded
    inputs (the CUDA op, sgl_kernel.silu_and_mul, evaluates in fp32 and rounds once; forward_native,
    python/sglang/srt/layers/activation.py:68-70, rounds silu(g) to T before the product).
    Returns (out_bits, out_f64)."""
    v = bits_to_f32(gate_up_bits, dtype).astype(np.float64)
    d = v.shape[-1] // 2
    g, u = v[..., :d], v[..., d:]
    y = g / (1.0 + np.exp(-g)) * u
    return f32_to_bits(y.astype(np.float32), dtype), y


def gate_up_interleave_perm(inter: int) -> np.ndarray:
    """perm[n'] = logical column (gate j -> j, up j -> inter + j) of PHYSICAL row n' of tiles made by
    awq_b200_repack_gate_up: slab s (16 rows) = gate rows 8s..8s+7 then up rows 8s..8s+7."""
    n = np.arange(2 * inter)
    s, p = n // 16, n % 16
    return np.where(p < 8, 8 * s + p, inter + 8 * s + (p - 8))


def permute_columns(qweight, scales_bits, qzeros, perm):
    """AWQ tensors whose logical column n' is column perm[n'] of the inputs (unpack, gather, pack)."""
    w = unpack_np(qweight)[:, perm]
    z = unpack_np(qzeros)[:, perm]
    return pack_np(w), np.ascontiguousarray(np.asarray(scales_bits)[:, perm]), pack_np(z)


def pack_np(vals: np.ndarray) -> np.ndarray:
    v = np.asarray(vals, dtype=np.uint32).reshape(vals.shape[0], -1, 8)
    out = np.zeros(v.shape[:2], dtype=np.uint32)
    for j in range(8):
        out |= (v[:, :, j] & np.uint32(0xF)) << np.uint32(4 * AWQ_REV[j])
    return out.view(np.int32)


# ---------------------------------------------------------------------------------------
# seeded synthetic inputs
# ---------------------------------------------------------------------------------------


def make_inputs(K: int, N: int, G: int, dtype: int, seed: int = 1234, M: int = 0, bias: bool = False,
                signed_scales: bool = False):
    """Full-range int32 words (sign bit is data), AWQ-magnitude scales, randn activations."""
    rng = np.random.default_rng(seed)
    qweight = rng.integers(-(2**31), 2**31, size=(K, N // 8), dtype=np.int64).astype(np.int32)
    qzeros = rng.integers(-(2**31), 2**31, size=(K // G, N // 8), dtype=np.int64).astype(np.int32)
    s = (rng.random((K // G, N), dtype=np.float32) * 0.02 + 0.005).astype(np.float32)
    if signed_scales:
        s *= rng.choice(np.array([-1.0, 1.0], dtype=np.float32), size=s.shape)
    out = {"qweight": qweight, "qzeros": qzeros, "scales": f32_to_bits(s, dtype)}
    if M:
        out["x"] = f32_to_bits(rng.standard_normal((M, K), dtype=np.float32), dtype)
    if bias:
        out["bias"] = f32_to_bits(rng.standard_normal(N, dtype=np.float32) * 0.1, dtype)
    return out
