"""Python wrappers over the registered torch ops (reference: sgl-kernel/python/sgl_kernel/gemm.py:8-11)."""
from typing import Optional, Tuple

import torch


def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """[K, N/8] int32, [K/G, N] f16|bf16, [K/G, N/8] int32 -> [K, N] in scales.dtype.

    Same signature and result (bit-for-bit) as the reference op, gemm.py:8-11."""
    return torch.ops.sgl_kernel.awq_dequantize.default(qweight, scales, qzeros)


def awq_b200_repack(qweight: torch.Tensor, scales: torch.Tensor,
                    qzeros: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """One-time load-time repack into the B200 tile layout (group size 128 only)."""
    return torch.ops.sgl_kernel.awq_b200_repack.default(qweight, scales, qzeros)


def awq_b200_dequantize_tiled(qweight_t: torch.Tensor, scales_t: torch.Tensor, qzeros_t: torch.Tensor,
                              in_features: int, out_features: int) -> torch.Tensor:
    """Dense [K, N] matrix from the tile layout; same bits as awq_dequantize on the originals."""
    return torch.ops.sgl_kernel.awq_b200_dequantize_tiled.default(qweight_t, scales_t, qzeros_t, in_features,
                                                                  out_features)


def awq_b200_linear(x: torch.Tensor, qweight_t: torch.Tensor, scales_t: torch.Tensor, qzeros_t: torch.Tensor,
                    bias: Optional[torch.Tensor], out_features: int) -> torch.Tensor:
    """y[..., N] = x[..., K] @ W (+ bias) straight from the packed tiles (fused dequant)."""
    return torch.ops.sgl_kernel.awq_b200_linear.default(x, qweight_t, scales_t, qzeros_t, bias, out_features)


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
    return (qweight.new_empty((T, 2048)), scales.new_empty((T, 128)),
            qweight.new_empty((T, 64), dtype=torch.uint8))


@torch.library.register_fake("sgl_kernel::awq_b200_dequantize_tiled")
def _(qweight_t, scales_t, qzeros_t, in_features, out_features):
    return scales_t.new_empty((in_features, out_features))


@torch.library.register_fake("sgl_kernel::awq_b200_linear")
def _(x, qweight_t, scales_t, qzeros_t, bias, out_features):
    return x.new_empty(x.shape[:-1] + (out_features,))
