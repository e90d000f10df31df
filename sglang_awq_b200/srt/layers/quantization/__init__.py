"""Quantization method registry (reference: python/sglang/srt/layers/quantization/__init__.py:49-94).
Only the entry on the AWQ hot path exists here."""
from typing import Dict, Type

from .awq import AWQConfig, AWQLinearMethod
from .base_config import LinearMethodBase, QuantizationConfig, QuantizeMethodBase
from .unquant import UnquantizedLinearMethod

BASE_QUANTIZATION_METHODS: Dict[str, Type[QuantizationConfig]] = {"awq": AWQConfig}
QUANTIZATION_METHODS = {**BASE_QUANTIZATION_METHODS}


def get_quantization_config(quantization: str) -> Type[QuantizationConfig]:
    if quantization not in QUANTIZATION_METHODS:
        raise ValueError(f"Invalid quantization method: {quantization}. "
                         f"Available methods: {list(QUANTIZATION_METHODS.keys())}")
    return QUANTIZATION_METHODS[quantization]


__all__ = ["AWQConfig", "AWQLinearMethod", "LinearMethodBase", "QuantizationConfig", "QuantizeMethodBase",
           "UnquantizedLinearMethod", "get_quantization_config", "QUANTIZATION_METHODS"]
