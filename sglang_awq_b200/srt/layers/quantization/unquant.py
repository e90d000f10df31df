"""Unquantized linear method (reference: python/sglang/srt/layers/quantization/unquant.py:94-150):
what AWQConfig.get_quant_method returns for `modules_to_not_convert` layers."""
from typing import List, Optional

import torch
import torch.nn.functional as F
from torch.nn import Parameter

from .base_config import LinearMethodBase
from .utils import set_weight_attrs


class UnquantizedLinearMethod(LinearMethodBase):
    def create_weights(self, layer, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        weight = Parameter(torch.empty(sum(output_partition_sizes), input_size_per_partition, dtype=params_dtype),
                           requires_grad=False)
        set_weight_attrs(weight, {"input_dim": 1, "output_dim": 0})
        layer.register_parameter("weight", weight)
        set_weight_attrs(weight, extra_weight_attrs)

    def apply(self, layer, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        return F.linear(x, layer.weight, bias)
