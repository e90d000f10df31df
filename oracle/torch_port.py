"""Torch-CPU restatement of the reference's CPU path for the AWQ linear op.

TEST INFRASTRUCTURE / TIMED BASELINE ONLY (bench.py cpu_baseline and --impl reference, tests).
The sgl-kernel CPU build has no AWQ/int4 kernel (sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:326-504),
so what the reference can run on host cores is its pure-torch dequant
(python/sglang/srt/layers/quantization/awq_triton.py:342-368, awq_dequantize_decomposition)
followed by torch.matmul and the bias add_ of AWQLinearMethod.apply (awq.py:434-451).  This
file restates that with the same torch ops (shift/mask per nibble in AWQ order, cast, subtract
the per-group zero, multiply by the per-group scale, matmul); tests/test_oracle.py pins it
bit-for-bit to the golden vectors generated from the reference itself.
"""
from typing import Optional

import torch

_NIBBLE_OF_COLUMN = (0, 4, 1, 5, 2, 6, 3, 7)  # logical column j of a word lives at nibble REV[j]


def awq_dequantize_cpu(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """[K, N/8] int32, [K/G, N] f16|bf16, [K/G, N/8] int32 -> [K, N] in scales.dtype."""
    K, groups = qweight.shape[0], qzeros.shape[0]
    w_cols, z_cols = [], []
    for nib in _NIBBLE_OF_COLUMN:
        w_cols.append((qweight >> (4 * nib)) & 0xF)
        z_cols.append((qzeros >> (4 * nib)) & 0xF)
    w = torch.stack(w_cols, dim=-1).reshape(K, -1).to(scales.dtype)
    z = torch.stack(z_cols, dim=-1).reshape(groups, -1).to(scales.dtype)
    w = w.reshape(groups, K // groups, -1)
    return ((w - z.unsqueeze(1)) * scales.unsqueeze(1)).reshape(K, -1)


def awq_linear_cpu(x: torch.Tensor, qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor,
                   bias: Optional[torch.Tensor] = None, matmul_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """AWQLinearMethod.apply on CPU tensors.  matmul_dtype=torch.float32 upcasts for the GEMM
    (CPU half GEMMs may be unoptimised; BASELINE.md times both)."""
    w = awq_dequantize_cpu(qweight, scales, qzeros)
    x2 = x.reshape(-1, x.shape[-1])
    if matmul_dtype is not None:
        out = torch.matmul(x2.to(matmul_dtype), w.to(matmul_dtype)).to(x.dtype)
    else:
        out = torch.matmul(x2, w)
    if bias is not None:
        out.add_(bias)
    return out.reshape(x.shape[:-1] + (w.shape[1],))
