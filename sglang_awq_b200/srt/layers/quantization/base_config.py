"""Quantization plugin ABCs -- the interface AWQLinearMethod plugs into.

Same contract as the reference's python/sglang/srt/layers/quantization/base_config.py
(QuantizeMethodBase :16-40, LinearMethodBase :42-81, QuantizationConfig :117-229); the MoE /
ModelOpt specific members are out of scope.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict, List, Optional

import torch
from torch import nn


class QuantizeMethodBase(ABC):
    @abstractmethod
    def create_weights(self, layer: nn.Module, *weight_args, **extra_weight_attrs):
        """Register the (quantized) parameters on `layer`."""
        raise NotImplementedError

    @abstractmethod
    def apply(self, layer: nn.Module, *args, **kwargs) -> torch.Tensor:
        """Run the layer's computation with the parameters created by create_weights."""
        raise NotImplementedError

    def process_weights_after_loading(self, layer: nn.Module) -> None:
        """Hook run once after the checkpoint has been copied in (loader.py:618-630)."""
        return


class LinearMethodBase(QuantizeMethodBase):
    @abstractmethod
    def create_weights(self, layer: nn.Module, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        """`output_partition_sizes`: per-rank widths of the logical matrices fused in this layer
        (e.g. [q, k, v]); `extra_weight_attrs["weight_loader"]` is the layer's loader."""
        raise NotImplementedError

    @abstractmethod
    def apply(self, layer: nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError


class QuantizationConfig(ABC):
    def __init__(self):
        super().__init__()
        self.packed_modules_mapping: Dict[str, List[str]] = dict()

    @abstractmethod
    def get_name(self) -> str:
        raise NotImplementedError

    @abstractmethod
    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def get_min_capability(cls) -> int:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def get_config_filenames() -> List[str]:
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def from_config(cls, config: Dict[str, Any]) -> "QuantizationConfig":
        raise NotImplementedError

    @classmethod
    def override_quantization_method(cls, hf_quant_cfg, user_quant) -> Optional[str]:
        return None

    @staticmethod
    def get_from_keys(config: Dict[str, Any], keys: List[str]) -> Any:
        for key in keys:
            if key in config:
                return config[key]
        raise ValueError(f"Cannot find any of {keys} in the model's quantization config.")

    @staticmethod
    def get_from_keys_or(config: Dict[str, Any], keys: List[str], default: Any) -> Any:
        try:
            return QuantizationConfig.get_from_keys(config, keys)
        except ValueError:
            return default

    @abstractmethod
    def get_quant_method(self, layer: nn.Module, prefix: str) -> Optional[QuantizeMethodBase]:
        raise NotImplementedError

    @abstractmethod
    def get_scaled_act_names(self) -> List[str]:
        raise NotImplementedError
