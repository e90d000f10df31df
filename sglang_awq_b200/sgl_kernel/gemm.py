"""Python wrappers over the registered torch ops (reference: sgl-kernel/python/sgl_kernel/gemm.py:8-11)."""
from typing import Optional

import torch


def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """[K, N/8] int32, [K/G, N] f16|bf16, [K/G, N/8] int32 -> [K, N] in scales.dtype.

    Same signature and result (bit-for-bit) as the reference op, gemm.py:8-11."""
    return torch.ops.sgl_kernel.awq_dequantize.default(qweight, scales, qzeros)


def awq_b200_repack(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """One-time load-time repack into the B200 tile layout.  Group size = K / scales.shape[0] as in the reference op:
    128 and every other multiple of 128 that divides K (incl. K itself, the reference's -1) give the 8576-byte records
    described below; 64 / 32 give 8960- / 9728-byte records (2 / 4 metadata sections) that the tcgen05 kernel reads.

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


def awq_b200_repack_gate_up(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """Repack of a MERGED gate_up weight ([K, 2 * inter / 8], gate columns first) with gate / up rows
    interleaved by 8 inside every 16-row slab; only for awq_b200_linear_silu_mul."""
    return torch.ops.sgl_kernel.awq_b200_repack_gate_up.default(qweight, scales, qzeros)


def awq_b200_linear_silu_mul(x: torch.Tensor, tiles: torch.Tensor, out_features: int) -> torch.Tensor:
    """[..., out_features / 2] = silu(x @ W_gate) * (x @ W_up): gate_up_proj + SiluAndMul in one launch."""
    return torch.ops.sgl_kernel.awq_b200_linear_silu_mul.default(x, tiles, out_features)


def fused_add_rmsnorm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6,
                      enable_pdl: Optional[bool] = None) -> None:
    """In place: residual += input; input = residual / RMS(residual) * weight.  Same signature as the
    reference wrapper (sgl-kernel/python/sgl_kernel/elementwise.py:49-83)."""
    torch.ops.sgl_kernel.fused_add_rmsnorm.default(input, residual, weight, eps, True if enable_pdl is None else enable_pdl)


def awq_b200_moe_linear(x: torch.Tensor, tiles: torch.Tensor, expert_ids: torch.Tensor, row_scale: Optional[torch.Tensor],
                        x_div: int, out_features: int, silu: bool) -> torch.Tensor:
    """[pairs, out_features (/ 2 with silu)]: one GEMV per (token, k) pair against expert `expert_ids[pair]` of the
    stacked tile arrays `tiles` [E, T, 2144]; row pair // x_div of x feeds pair."""
    return torch.ops.sgl_kernel.awq_b200_moe_linear.default(x, tiles, expert_ids, row_scale, x_div, out_features, silu)


def awq_b200_moe_sum(o: torch.Tensor, topk: int) -> torch.Tensor:
    """[tokens, hidden] = sum over the topk rows of each token (fp32 accumulation, one rounding)"""
    return torch.ops.sgl_kernel.awq_b200_moe_sum.default(o, topk)


def silu_and_mul(input: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """reference wrapper sgl-kernel/python/sgl_kernel/elementwise.py:172-184"""
    if input.shape[-1] * input.dtype.itemsize % 16 != 0:
        raise ValueError("The pointers must be multiple of 16 bytes.")
    if out is None:
        out = torch.empty(input.shape[:-1] + (input.shape[-1] // 2,), device=input.device, dtype=input.dtype)
    torch.ops.sgl_kernel.silu_and_mul.default(out, input)
    return out


def rmsnorm(input: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6, out: Optional[torch.Tensor] = None,
            enable_pdl: Optional[bool] = None) -> torch.Tensor:
    """reference wrapper sgl-kernel/python/sgl_kernel/elementwise.py (rmsnorm): out = input / RMS(input) * weight"""
    if out is None:
        out = torch.empty_like(input)
    torch.ops.sgl_kernel.rmsnorm.default(out, input, weight, eps, True if enable_pdl is None else enable_pdl)
    return out


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
    G = K // scales.shape[0]
    return qweight.new_empty((T, {32: 2432, 64: 2240}.get(G, 2144)))


@torch.library.register_fake("sgl_kernel::awq_b200_dequantize_tiled")
def _(tiles, in_features, out_features, dtype):
    return tiles.new_empty((in_features, out_features), dtype=dtype)


@torch.library.register_fake("sgl_kernel::awq_b200_linear")
def _(x, tiles, bias, out_features):
    return x.new_empty(x.shape[:-1] + (out_features,))


@torch.library.register_fake("sgl_kernel::awq_b200_repack_gate_up")
def _(qweight, scales, qzeros):
    K, N = qweight.shape[0], qweight.shape[1] * 8
    return qweight.new_empty((((N + 127) // 128) * (K // 128), 2144))


@torch.library.register_fake("sgl_kernel::awq_b200_linear_silu_mul")
def _(x, tiles, out_features):
    return x.new_empty(x.shape[:-1] + (out_features // 2,))


@torch.library.register_fake("sgl_kernel::fused_add_rmsnorm")
def _(input, residual, weight, eps, enable_pdl):
    return None


@torch.library.register_fake("sgl_kernel::awq_b200_moe_linear")
def _(x, tiles, expert_ids, row_scale, x_div, out_features, silu):
    return x.new_empty((expert_ids.numel(), out_features // 2 if silu else out_features))


@torch.library.register_fake("sgl_kernel::awq_b200_moe_sum")
def _(o, topk):
    return o.new_empty((o.shape[0] // topk, o.shape[1]))


@torch.library.register_fake("sgl_kernel::silu_and_mul")
def _(out, input):
    return None


@torch.library.register_fake("sgl_kernel::rmsnorm")
def _(output, input, weight, eps, enable_pdl):
    return None
