"""FusedMoE: the slice of the reference's mixture-of-experts layer that the AWQ method touches
(python/sglang/srt/layers/moe/fused_moe_triton/layer.py `FusedMoE`: parameter creation through the quant method,
the per-expert weight loader with w1 / w3 -> w13 halves and TP sharding, forward = quant_method.apply + all-reduce).
Routing (gate linear, top-k, renormalisation), expert parallelism and the dispatcher classes are outside the path
(SURVEY.md §8): `forward` takes the router's top-k output."""
from __future__ import annotations

from typing import Optional

import torch

from ..distributed import get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size, tensor_model_parallel_all_reduce
from .quantization.base_config import QuantizationConfig


class FusedMoE(torch.nn.Module):
    def __init__(self, num_experts: int, hidden_size: int, intermediate_size: int, top_k: int,
                 params_dtype: Optional[torch.dtype] = None, quant_config: Optional[QuantizationConfig] = None, prefix: str = "",
                 reduce_results: bool = True, tp_rank: Optional[int] = None, tp_size: Optional[int] = None):
        super().__init__()
        self.tp_size = get_tensor_model_parallel_world_size() if tp_size is None else tp_size
        self.tp_rank = get_tensor_model_parallel_rank() if tp_rank is None else tp_rank
        assert intermediate_size % self.tp_size == 0
        self.num_experts, self.hidden_size, self.top_k = num_experts, hidden_size, top_k
        self.intermediate_size_per_partition = intermediate_size // self.tp_size
        self.reduce_results = reduce_results
        self.params_dtype = params_dtype or torch.get_default_dtype()
        assert quant_config is not None, "only quantized (AWQ) experts are on this path"
        self.quant_method = quant_config.get_quant_method(self, prefix)
        assert self.quant_method is not None
        self.quant_method.create_weights(layer=self, num_experts=num_experts, hidden_size=hidden_size,
                                         intermediate_size_per_partition=self.intermediate_size_per_partition,
                                         params_dtype=self.params_dtype, weight_loader=self.weight_loader)

    def weight_loader(self, param: torch.nn.Parameter, loaded_weight: torch.Tensor, weight_name: str, shard_id: str,
                      expert_id: int) -> None:
        """checkpoint tensor of ONE expert's w1 (gate) / w3 (up) / w2 (down) -> this rank's slice of the stacked
        parameter.  AWQ tensors are stored transposed ([in, out / 8] etc., `is_transposed`): w1 / w3 shard their LAST
        dim (the output), w2 its FIRST (the input), in units of the parameter's own packing (reference layer.py
        `_load_w13` / `_load_w2` with is_transposed=True)."""
        assert shard_id in ("w1", "w2", "w3")
        dst = param.data[expert_id]
        if shard_id == "w2":
            rows = dst.shape[0]
            dst.copy_(loaded_weight.narrow(0, self.tp_rank * rows, rows))
            return
        half = dst.shape[1] // 2
        src = loaded_weight.narrow(1, self.tp_rank * half, half)
        dst.narrow(1, 0 if shard_id == "w1" else half, half).copy_(src)

    def forward(self, hidden_states: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor) -> torch.Tensor:
        out = self.quant_method.apply(self, hidden_states, topk_weights, topk_ids)
        if self.reduce_results and self.tp_size > 1:
            out = tensor_model_parallel_all_reduce(out)
        return out
