"""Mirror of the slice of the reference's `sgl_kernel` package that the AWQ path uses.

Reference: sgl-kernel/python/sgl_kernel/__init__.py:5,42-43 (loads common_ops, re-exports
`awq_dequantize`), sgl-kernel/python/sgl_kernel/load_utils.py:48-188 (picks an sm90/ or sm100/
build).  Here there is ONE sm_100a library and no architecture dispatch: importing this package
loads `common_ops.so`, which registers the `sgl_kernel::*` torch ops, and fails loudly if the
library has not been built.
"""
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_OPS_PATH = os.path.join(_HERE, "common_ops.so")


def _load_ops():
    if not os.path.exists(_OPS_PATH):
        raise ImportError(
            f"{_OPS_PATH} is missing: run `python -m sglang_awq_b200.build`. The AWQ ops have no "
            "Python/CPU fallback.")
    torch.ops.load_library(_OPS_PATH)


_load_ops()

from .gemm import (  # noqa: E402
    awq_b200_dequantize_tiled,
    awq_b200_linear,
    awq_b200_linear_silu_mul,
    awq_b200_moe_linear,
    awq_b200_moe_sum,
    awq_b200_repack,
    awq_b200_repack_gate_up,
    awq_dequantize,
    fused_add_rmsnorm,
    rmsnorm,
    silu_and_mul,
)

__all__ = ["awq_dequantize", "awq_b200_repack", "awq_b200_linear", "awq_b200_dequantize_tiled", "awq_b200_repack_gate_up",
           "awq_b200_linear_silu_mul", "fused_add_rmsnorm", "awq_b200_moe_linear", "awq_b200_moe_sum", "silu_and_mul", "rmsnorm"]
