"""Quantization method registry (reference: python/sglang/srt/layers/quantization/__init__.py:49-94).
Only the entries on the AWQ hot path exist here ("awq", and "awq_marlin", which the reference
resolves AWQ checkpoints to by default; both run on the same B200 kernels)."""
from typing import Dict, Type

from .awq import AWQConfig, AWQLinearMethod, AWQMarlinConfig, AWQMarlinLinearMethod, AWQMoEMethod
from .base_config import LinearMethodBase, QuantizationConfig, QuantizeMethodBase
from .unquant import UnquantizedLinearMethod

BASE_QUANTIZATION_METHODS: Dict[str, Type[QuantizationConfig]] = {"awq": AWQConfig, "awq_marlin": AWQMarlinConfig}
QUANTIZATION_METHODS = {**BASE_QUANTIZATION_METHODS}


def get_quantization_config(quantization: str) -> Type[QuantizationConfig]:
    if quantization not in QUANTIZATION_METHODS:
        raise ValueError(f"Invalid quantization method: {quantization}. "
                         f"Available methods: {list(QUANTIZATION_METHODS.keys())}")
    return QUANTIZATION_METHODS[quantization]


__all__ = ["AWQConfig", "AWQLinearMethod", "AWQMarlinConfig", "AWQMarlinLinearMethod", "AWQMoEMethod", "LinearMethodBase", "QuantizationConfig", "QuantizeMethodBase",
           "UnquantizedLinearMethod", "get_quantization_config", "QUANTIZATION_METHODS"]
