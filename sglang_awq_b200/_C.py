"""ctypes binding of the C ABI (include/awq_b200.h -> lib/libawq_b200.so).

This is the same boundary a non-Python host would bind.  It takes torch
tensors only as a convenient owner of device memory: every call below passes raw device
pointers, sizes and the current CUDA stream handle.  There is no fallback: if the shared
library is missing, or the device is not sm_100, calls raise.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Optional, Tuple

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("AWQ_B200_LIB") or os.path.join(_PKG, "lib", "libawq_b200.so")  # override: dev builds only
HEADER_PATH = os.path.join(os.path.dirname(_PKG), "include", "awq_b200.h")

F16, BF16 = 0, 1
_lib: Optional[ctypes.CDLL] = None

_vp, _i, _i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64

_SIGNATURES = {
    "awq_b200_abi_version": (_i, []),
    "awq_b200_last_error": (ctypes.c_char_p, []),
    "awq_b200_check_device": (_i, []),
    "awq_b200_dequantize": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp]),
    "awq_b200_num_tiles": (_i64, [_i64, _i64]),
    "awq_b200_repack": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i, _vp]),
    "awq_b200_repack_gate_up": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i, _vp]),
    "awq_b200_dequantize_tiled": (_i, [_vp, _vp, _i64, _i64, _i, _vp]),
    "awq_b200_tile_bytes": (_i64, [_i64]),
    "awq_b200_repack_g": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp]),
    "awq_b200_dequantize_tiled_g": (_i, [_vp, _vp, _i64, _i64, _i64, _i, _vp]),
    "awq_b200_linear_g": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i, _vp, _i64, _vp]),
    "awq_b200_workspace_bytes": (_i64, [_i64, _i64]),
    "awq_b200_gemv": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp, _i64, _vp]),
    "awq_b200_gemm": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp, _i64, _vp]),
    "awq_b200_gemv_static": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp]),
    "awq_b200_linear": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp, _i64, _vp]),
    "awq_b200_linear_silu_mul": (_i, [_vp, _vp, _vp, _i64, _i64, _i64, _i, _vp, _i64, _vp]),
    "awq_b200_linear_route": (_i, [_i64, _i64, _i64]),
    "awq_b200_gemv_route": (_i, [_i64, _i64, _i64, _i]),
    "awq_b200_allreduce_buffer_bytes": (_i64, [_i64]),
    "awq_b200_debug_set_allreduce_packets": (None, [_i]),
    "awq_b200_allreduce": (_i, [_vp, _vp, _i64, _i, _vp, _i64, _i, _i, _vp]),
    "awq_b200_residual_rmsnorm": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, ctypes.c_float, _i, _vp]),
    "awq_b200_allreduce_residual_rmsnorm": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, ctypes.c_float, _i, _vp, _i64, _i, _i, _vp]),
    "awq_b200_moe_linear": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i, _i, _vp]),
    "awq_b200_moe_sum": (_i, [_vp, _vp, _i64, _i64, _i64, _i, _vp]),
    "awq_b200_silu_and_mul": (_i, [_vp, _vp, _i64, _i64, _i, _vp]),
    "awq_b200_linear_allreduce": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i, _vp, _i64, _vp, _i64, _i, _i, _vp]),
    "awq_b200_debug_set_linear_allreduce_fusion": (None, [_i]),
    "awq_b200_set_pdl": (None, [_i]),
    "awq_b200_set_l2_prefetch_mb": (None, [_i]),
    "awq_b200_set_gemv_ctas_per_sm": (None, [_i]),
    "awq_b200_set_gemv_shared_x_min_m": (None, [_i]),
    "awq_b200_set_linear_gemv_max_m": (None, [_i]),
    "awq_b200_set_gemv_direct_max_m": (None, [_i]),
    "awq_b200_debug_set_gemv_impl": (None, [_i, _i, _i]),
    "awq_b200_debug_set_dequant_variant": (None, [_i]),
    "awq_b200_debug_set_gemm_plan": (None, [_i, _i]),
    "awq_b200_debug_gemv_impl_fits": (_i, [_i, _i64, _i64, _i64, _i]),
    "awq_b200_debug_set_tc_trace": (None, [_vp]),
    "awq_b200_debug_set_sub_decode": (None, [_i]),
    "awq_b200_debug_set_moe_chunk": (None, [_i]),
    "awq_b200_debug_set_moe_blocks_per_sm": (None, [_i]),
}


def declared_symbols() -> list:
    """Every function name include/awq_b200.h declares (used by the symbol-export test)."""
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(awq_b200_[a-z0-9_]+)\s*\(", text)))


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -m sglang_awq_b200.build` (nvcc, sm_100a). "
                "There is no CPU or PyTorch fallback for the AWQ kernels.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            if os.environ.get("AWQ_B200_LIB") and not hasattr(L, name):
                continue  # a development build of an older source (A/B timing): the symbols it lacks stay unbound
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.awq_b200_abi_version() != 1:
            raise RuntimeError("libawq_b200.so ABI version mismatch")
        _lib = L
    return _lib


def _check(rc: int):
    if rc != 0:
        raise RuntimeError(lib().awq_b200_last_error().decode())


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float16:
        return F16
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"expected float16/bfloat16, got {t.dtype}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


_workspaces = {}
_retired_workspaces = []  # never freed: a captured CUDA graph may still hold the pointer


TILE_WORDS = 2144  # 8576-byte record as int32


def workspace(device: torch.device, max_m: int = 64, max_n: int = 65536) -> torch.Tensor:
    """Zero-initialised split-K scratch per (device, stream).  Sized on first use for every decode
    shape and prefill up to M = 8192; a buffer that was handed out is never freed (graphs keep the
    pointer) and growth during stream capture raises."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _stream())
    need = lib().awq_b200_workspace_bytes(max_m, max_n)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < need:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("awq_b200: the split-K workspace must grow during CUDA graph capture; "
                               "run this shape once eagerly on the capture stream first")
        need = max(need, lib().awq_b200_workspace_bytes(64, 65536), lib().awq_b200_workspace_bytes(8192, 65536))
        if ws is not None:
            _retired_workspaces.append(ws)
        ws = torch.zeros(need, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    K, NP = qweight.shape
    out = torch.empty((K, NP * 8), dtype=scales.dtype, device=scales.device)
    with torch.cuda.device(qweight.device):
        _check(lib().awq_b200_dequantize(_p(qweight), _p(scales), _p(qzeros), _p(out), K, NP * 8, scales.shape[0],
                                         _dt(scales), _stream()))
    return out


def repack(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """-> tiles int32 [T, 2144]: T records of 8576 bytes (weights | scales | zeros)."""
    K, NP = qweight.shape
    N = NP * 8
    T = lib().awq_b200_num_tiles(K, N)
    if T < 0:
        raise RuntimeError("awq_b200_repack: K must be a multiple of 128")
    tiles = torch.empty((T, TILE_WORDS), dtype=torch.int32, device=qweight.device)
    with torch.cuda.device(qweight.device):
        _check(lib().awq_b200_repack(_p(qweight), _p(scales), _p(qzeros), _p(tiles), K, N, _dt(scales), _stream()))
    return tiles


def repack_gate_up(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """merged gate_up weight -> tiles with gate / up rows interleaved by 8 (for linear_silu_mul)."""
    K, NP = qweight.shape
    N = NP * 8
    T = lib().awq_b200_num_tiles(K, N)
    if T < 0:
        raise RuntimeError("awq_b200_repack_gate_up: K must be a multiple of 128")
    tiles = torch.empty((T, TILE_WORDS), dtype=torch.int32, device=qweight.device)
    with torch.cuda.device(qweight.device):
        _check(lib().awq_b200_repack_gate_up(_p(qweight), _p(scales), _p(qzeros), _p(tiles), K, N, _dt(scales), _stream()))
    return tiles


def dequantize_tiled(tiles: torch.Tensor, K: int, N: int, dtype: torch.dtype) -> torch.Tensor:
    out = torch.empty((K, N), dtype=dtype, device=tiles.device)
    with torch.cuda.device(tiles.device):
        _check(lib().awq_b200_dequantize_tiled(_p(tiles), _p(out), K, N, _dt(out), _stream()))
    return out


def _mm(fn_name: str, x, tiles, N: int, bias=None, out=None) -> torch.Tensor:
    M, K = x.shape
    y = out if out is not None else torch.empty((M, N), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, max_m=max(16, M))
        fn = getattr(lib(), fn_name)
        _check(fn(_p(x), _p(tiles), _p(bias), _p(y), M, K, N, _dt(x), _p(ws), ws.numel(), _stream()))
    return y


def gemv(x, tiles, N: int, bias=None, out=None) -> torch.Tensor:
    return _mm("awq_b200_gemv", x, tiles, N, bias, out)


def gemv_static(x, tiles, N: int, bias=None, out=None) -> torch.Tensor:
    """awq_b200_gemv_static: the tcgen05 decode kernel on the library's static device scratch.  `out` must be
    given by callers that may not allocate; calls share the scratch, keep them on one stream."""
    M, K = x.shape
    y = out if out is not None else torch.empty((M, N), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        _check(lib().awq_b200_gemv_static(_p(x), _p(tiles), _p(bias), _p(y), M, K, N, _dt(x), _stream()))
    return y


def gemm(x, tiles, N: int, bias=None, out=None) -> torch.Tensor:
    return _mm("awq_b200_gemm", x, tiles, N, bias, out)


def linear(x, tiles, N: int, bias=None, out=None) -> torch.Tensor:
    return _mm("awq_b200_linear", x, tiles, N, bias, out)


def linear_silu_mul(x, tiles, N: int, out=None) -> torch.Tensor:
    """y[M, N/2] = silu(x @ W_gate) * (x @ W_up); tiles from repack_gate_up."""
    M, K = x.shape
    y = out if out is not None else torch.empty((M, N // 2), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, max_m=max(16, M))
        _check(lib().awq_b200_linear_silu_mul(_p(x), _p(tiles), _p(y), M, K, N, _dt(x), _p(ws), ws.numel(), _stream()))
    return y


def silu_and_mul(x, out=None) -> torch.Tensor:
    """[rows, 2 d] -> [rows, d] = silu(x[:, :d]) * x[:, d:]"""
    rows, d2 = x.shape
    out = torch.empty((rows, d2 // 2), dtype=x.dtype, device=x.device) if out is None else out
    with torch.cuda.device(x.device):
        _check(lib().awq_b200_silu_and_mul(_p(x), _p(out), rows, d2 // 2, _dt(x), _stream()))
    return out


def residual_rmsnorm(x, residual, weight, eps: float = 1e-6, out=None) -> torch.Tensor:
    """residual <- x + residual (in place; None: plain RMSNorm); returns rmsnorm(residual) * weight (into `out`, default x)."""
    rows, hidden = x.shape
    out = x if out is None else out
    with torch.cuda.device(x.device):
        _check(lib().awq_b200_residual_rmsnorm(_p(x), _p(residual), _p(weight), _p(out), rows, hidden, float(eps), _dt(x),
                                               _stream()))
    return out


def moe_linear(x, tiles, expert_ids, N: int, x_div: int = 1, row_scale=None, silu: bool = False) -> torch.Tensor:
    """one GEMV per (token, k) pair against the pair's expert; tiles int32 [E, T, 2144]; -> [pairs, N] ([pairs, N/2] with silu)"""
    pairs, K = expert_ids.numel(), x.shape[-1]
    y = torch.empty((pairs, N // 2 if silu else N), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        _check(lib().awq_b200_moe_linear(_p(x), _p(tiles), _p(expert_ids), _p(row_scale), _p(y), pairs, x_div, K, N, tiles.shape[0],
                                         1 if silu else 0, _dt(x), _stream()))
    return y


def moe_sum(o, topk: int) -> torch.Tensor:
    pairs, hidden = o.shape
    out = torch.empty((pairs // topk, hidden), dtype=o.dtype, device=o.device)
    with torch.cuda.device(o.device):
        _check(lib().awq_b200_moe_sum(_p(o), _p(out), pairs // topk, topk, hidden, _dt(o), _stream()))
    return out


def repack_g(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """any group size of the reference op (K / scales rows: 32, 64 or a multiple of 128) -> tiles int32 [T, record words]"""
    K, NP = qweight.shape
    N, G = NP * 8, K // scales.shape[0]
    rec = lib().awq_b200_tile_bytes(G)
    T = lib().awq_b200_num_tiles(K, N)
    if rec < 0 or T < 0:
        raise RuntimeError(f"awq_b200_repack_g: unsupported group size {G} / K {K}")
    tiles = torch.empty((T, rec // 4), dtype=torch.int32, device=qweight.device)
    with torch.cuda.device(qweight.device):
        _check(lib().awq_b200_repack_g(_p(qweight), _p(scales), _p(qzeros), _p(tiles), K, N, G, _dt(scales), _stream()))
    return tiles


def _group_of(tiles: torch.Tensor) -> int:
    return {2144: 128, 2240: 64, 2432: 32}[tiles.shape[-1]]


def dequantize_tiled_g(tiles: torch.Tensor, K: int, N: int, dtype: torch.dtype) -> torch.Tensor:
    out = torch.empty((K, N), dtype=dtype, device=tiles.device)
    with torch.cuda.device(tiles.device):
        _check(lib().awq_b200_dequantize_tiled_g(_p(tiles), _p(out), K, N, _group_of(tiles), _dt(out), _stream()))
    return out


def linear_g(x, tiles, N: int, bias=None) -> torch.Tensor:
    M, K = x.shape
    y = torch.empty((M, N), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        ws = workspace(x.device, max_m=max(16, M))
        _check(lib().awq_b200_linear_g(_p(x), _p(tiles), _p(bias), _p(y), M, K, N, _group_of(tiles), _dt(x), _p(ws), ws.numel(),
                                       _stream()))
    return y
