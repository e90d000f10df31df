"""Python wrappers over the registered torch ops (reference: sgl-kernel/python/sgl_kernel/gemm.py:8-11)."""
from typing import Optional

import torch


def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """[K, N/8] int32, [K/G, N] f16|bf16, [K/G, N/8] int32 -> [K, N] in scales.dtype.

    Same signature and result (bit-for-bit) as the reference op, gemm.py:8-11."""
    return torch.ops.sgl_kernel.awq_dequantize.default(qweight, scales, qzeros)


def awq_b200_repack(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """One-time load-time repack into the B200 tile layout (group size 128 only).

    Returns `tiles`, int32 [T, 2144]: T = ceil(N/128) * K/128 records of 8576 bytes =
    8192 B packed weights (128 rows x 64 B) followed by 8 metadata blocks of 48 B, one per 16
    rows: 16 scales (32 B) then 16 zero-point bytes (include/awq_b200.h)."""
    return torch.ops.sgl_kernel.awq_b200_repack.default(qweight, scales, qzeros)


def awq_b200_dequantize_tiled(tiles: torch.Tensor, in_features: int, out_features: int,
                              dtype: torch.dtype = torch.float16) -> torch.Tensor:
    """Dense [K, N] matrix from the tile layout; same bits as awq_dequantize on the originals."""
    return torch.ops.sgl_kernel.awq_b200_dequantize_tiled.default(tiles, in_features, out_features, dtype)


def awq_b200_linear(x: torch.Tensor, tiles: torch.Tensor, bias: Optional[torch.Tensor],
                    out_features: int) -> torch.Tensor:
    """y[..., N] = x[..., K] @ W (+ bias) straight from the packed tiles (fused dequant)."""
    return torch.ops.sgl_kernel.awq_b200_linear.default(x, tiles, bias, out_features)


# Meta ("fake") implementations for torch.compile / tracing.  The reference registers a fake
# for awq_dequantize with SIX arguments although the op takes three
# (python/sglang/srt/layers/quantization/awq.py:947-960 vs common_extension.cc:126); these match
# the real schemas.
@torch.library.register_fake("sgl_kernel::awq_dequantize")
def _(qweight, scales, qzeros):
    return qweight.new_empty((qweight.shape[0], qweight.shape[1] * 8), dtype=scales.dtype)


@torch.library.register_fake("sgl_kernel::awq_b200_repack")
def _(qweight, scales, qzeros):
    K, N = qweight.shape[0], qweight.shape[1] * 8
    T = ((N + 127) // 128) * (K // 128)
    return qweight.new_empty((T, 2144))


@torch.library.register_fake("sgl_kernel::awq_b200_dequantize_tiled")
def _(tiles, in_features, out_features, dtype):
    return tiles.new_empty((in_features, out_features), dtype=dtype)


@torch.library.register_fake("sgl_kernel::awq_b200_linear")
def _(x, tiles, bias, out_features):
    return x.new_empty(x.shape[:-1] + (out_features,))
