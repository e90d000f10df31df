"""Tensor-parallel linear layers that host the AWQ linear method.

Same classes, constructor arguments, attributes and forward contracts as the reference's
python/sglang/srt/layers/linear.py: LinearBase :135-172, ReplicatedLinear :175-267,
ColumnParallelLinear :270-458, MergedColumnParallelLinear :461-775, QKVParallelLinear :778-1209,
RowParallelLinear :1212-1422.  Column-parallel layers split N and need no communication;
row-parallel layers split K on group boundaries and close with ONE all-reduce (sum) over the TP
group.  Only the loader paths the AWQ checkpoint format exercises are implemented
(weight_loader_v2 on the TP-aware parameter classes, plus the plain loader for biases and
unquantized weights); GGUF / bitsandbytes / block-fp8 / Marlin special cases are out of scope.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
from torch.nn.parameter import Parameter

from ..distributed import (
    divide,
    get_tensor_model_parallel_rank,
    get_tensor_model_parallel_world_size,
    split_tensor_along_last_dim,
    tensor_model_parallel_all_gather,
    tensor_model_parallel_all_reduce,
)
from .parameter import BasevLLMParameter, PackedvLLMParameter, RowvLLMParameter, _ColumnvLLMParameter
from .quantization.base_config import QuantizationConfig, QuantizeMethodBase
from .quantization.unquant import UnquantizedLinearMethod
from .quantization.utils import set_weight_attrs

# the layers choose weight_loader_v2 by the quant method's class NAME (reference linear.py:51-69)
WEIGHT_LOADER_V2_SUPPORTED = ["AWQLinearMethod", "AWQMarlinLinearMethod"]


def _tp(tp_rank: Optional[int], tp_size: Optional[int]) -> Tuple[int, int]:
    if tp_rank is None:
        tp_rank = get_tensor_model_parallel_rank()
    if tp_size is None:
        tp_size = get_tensor_model_parallel_world_size()
    return tp_rank, tp_size


class LinearBase(torch.nn.Module):
    """Holds sizes and picks the quant method (reference :135-172)."""

    def __init__(self, input_size: int, output_size: int, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None, quant_config: Optional[QuantizationConfig] = None,
                 prefix: str = ""):
        super().__init__()
        self.input_size = input_size
        self.output_size = output_size
        self.skip_bias_add = skip_bias_add
        self.params_dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        self.quant_config = quant_config
        if quant_config is None:
            self.quant_method: Optional[QuantizeMethodBase] = UnquantizedLinearMethod()
        else:
            self.quant_method = quant_config.get_quant_method(self, prefix=prefix)

    def _loader_for_method(self):
        v2 = self.quant_method.__class__.__name__ in WEIGHT_LOADER_V2_SUPPORTED
        return self.weight_loader_v2 if v2 else self.weight_loader

    def _make_bias(self, size: int, params_dtype, bias: bool):
        if bias:
            self.bias = Parameter(torch.empty(size, dtype=params_dtype))
            set_weight_attrs(self.bias, {"output_dim": 0, "weight_loader": self.weight_loader})
        else:
            self.register_parameter("bias", None)

    def forward(self, x: torch.Tensor):
        raise NotImplementedError


class ReplicatedLinear(LinearBase):
    """Every rank holds the whole matrix (reference :175-267)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None, quant_config: Optional[QuantizationConfig] = None,
                 prefix: str = ""):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix=prefix)
        assert self.quant_method is not None
        self.quant_method.create_weights(self, self.input_size, [self.output_size], self.input_size,
                                         self.output_size, self.params_dtype, weight_loader=self.weight_loader)
        self._make_bias(self.output_size, self.params_dtype, bias)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        assert param.size() == loaded_weight.size()
        param.data.copy_(loaded_weight)

    def forward(self, x: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        bias = self.bias if not self.skip_bias_add else None
        output = self.quant_method.apply(self, x, bias)
        return output, (self.bias if self.skip_bias_add else None)

    def extra_repr(self) -> str:
        return f"in_features={self.input_size}, output_features={self.output_size}, bias={self.bias is not None}"


class ColumnParallelLinear(LinearBase):
    """Y = X A + b with A split along its output dimension: A = [A_1 .. A_p] (reference :270-458)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, gather_output: bool = False,
                 skip_bias_add: bool = False, params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None, output_sizes: Optional[List[int]] = None,
                 prefix: str = "", tp_rank: Optional[int] = None, tp_size: Optional[int] = None,
                 use_presharded_weights: bool = False):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix)
        self.gather_output = gather_output
        self.use_presharded_weights = use_presharded_weights
        self.tp_rank, self.tp_size = _tp(tp_rank, tp_size)
        assert self.quant_method is not None
        self.output_size_per_partition = divide(self.output_size, self.tp_size)
        self.output_partition_sizes = [self.output_size_per_partition]
        if hasattr(self, "output_sizes"):  # QKV / merged: width of every fused logical matrix on this rank
            self.output_partition_sizes = [divide(s, self.tp_size) for s in self.output_sizes]
        self.quant_method.create_weights(
            layer=self, input_size_per_partition=self.input_size,
            output_partition_sizes=self.output_partition_sizes, input_size=self.input_size,
            output_size=self.output_size, params_dtype=self.params_dtype, weight_loader=self._loader_for_method())
        self._make_bias(self.output_size_per_partition, params_dtype, bias)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        output_dim = getattr(param, "output_dim", None)
        param_data = param.data
        if output_dim is not None and not self.use_presharded_weights:
            shard = param_data.shape[output_dim]
            loaded_weight = loaded_weight.narrow(output_dim, self.tp_rank * shard, shard)
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        assert param_data.shape == loaded_weight.shape
        param_data.copy_(loaded_weight)

    def weight_loader_v2(self, param: Parameter, loaded_weight: torch.Tensor):
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        param.load_column_parallel_weight(loaded_weight, tp_rank=self.tp_rank,
                                          use_presharded_weights=self.use_presharded_weights)

    def forward(self, input_):
        bias = self.bias if not self.skip_bias_add else None
        output_parallel = self.quant_method.apply(self, input_, bias)
        output = tensor_model_parallel_all_gather(output_parallel) if self.gather_output else output_parallel
        return output, (self.bias if self.skip_bias_add else None)

    def extra_repr(self) -> str:
        return (f"in_features={self.input_size}, output_features={self.output_size_per_partition}, "
                f"bias={self.bias is not None}, tp_size={self.tp_size}, gather_output={self.gather_output}")


class MergedColumnParallelLinear(ColumnParallelLinear):
    """Several column-parallel matrices fused along N, e.g. gate_up_proj (reference :461-775).
    Each logical matrix is sharded separately, so rank r holds [gate_r | up_r]."""

    def __init__(self, input_size: int, output_sizes: List[int], bias: bool = True, gather_output: bool = False,
                 skip_bias_add: bool = False, params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None, prefix: str = "",
                 tp_rank: Optional[int] = None, tp_size: Optional[int] = None, use_presharded_weights: bool = False):
        self.output_sizes = output_sizes
        tp_rank, tp_size = _tp(tp_rank, tp_size)
        assert all(s % tp_size == 0 for s in output_sizes)
        super().__init__(input_size=input_size, output_size=sum(output_sizes), bias=bias,
                         gather_output=gather_output, skip_bias_add=skip_bias_add, params_dtype=params_dtype,
                         quant_config=quant_config, prefix=prefix, tp_rank=tp_rank, tp_size=tp_size,
                         use_presharded_weights=use_presharded_weights)
        self.prefix = prefix

    def forward_silu_mul(self, input_):
        """gate_up_proj + SiluAndMul as the MLP uses them (reference python/sglang/srt/models/llama.py:94-106:
        `gate_up, _ = self.gate_up_proj(x); x = self.act_fn(gate_up)`): returns [..., intermediate / tp].  One fused
        launch when `self.awq_fuse_silu_mul` was set before process_weights_after_loading."""
        assert len(self.output_sizes) == 2 and self.bias is None and not self.gather_output
        return self.quant_method.apply_silu_mul(self, input_)

    def _rank_slice(self, shard_id: int) -> Tuple[int, int]:
        return sum(self.output_sizes[:shard_id]) // self.tp_size, self.output_sizes[shard_id] // self.tp_size

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor, loaded_shard_id: Optional[int] = None):
        """Plain tensors (bias, unquantized weight)."""
        output_dim = getattr(param, "output_dim", None)
        if loaded_shard_id is None:  # checkpoint tensor is already fused: split and recurse
            if output_dim is None:
                assert param.data.shape == loaded_weight.shape
                param.data.copy_(loaded_weight)
                return
            offset = 0
            for i, size in enumerate(self.output_sizes):
                self.weight_loader(param, loaded_weight.narrow(output_dim, offset, size), i)
                offset += size
            return
        assert loaded_shard_id < len(self.output_sizes)
        param_data = param.data
        if output_dim is not None:
            shard_offset, shard_size = self._rank_slice(loaded_shard_id)
            param_data = param_data.narrow(output_dim, shard_offset, shard_size)
            if not self.use_presharded_weights:
                loaded_weight = loaded_weight.narrow(output_dim, self.tp_rank * shard_size, shard_size)
        assert param_data.shape == loaded_weight.shape
        param_data.copy_(loaded_weight)

    def _load_fused_module_from_checkpoint(self, param: BasevLLMParameter, loaded_weight: torch.Tensor):
        offset = 0
        for shard_id, size in enumerate(self.output_sizes):
            shard_size, shard_offset = size, offset
            if isinstance(param, PackedvLLMParameter) and param.packed_dim == param.output_dim:
                shard_size, shard_offset = param.adjust_shard_indexes_for_packing(shard_size=size,
                                                                                 shard_offset=offset)
            self.weight_loader_v2(param, loaded_weight.narrow(param.output_dim, shard_offset, shard_size), shard_id)
            offset += size

    def weight_loader_v2(self, param: BasevLLMParameter, loaded_weight: torch.Tensor,
                         loaded_shard_id: Optional[int] = None):
        if loaded_shard_id is None:
            if type(param) in (RowvLLMParameter, BasevLLMParameter):
                param.load_merged_column_weight(loaded_weight=loaded_weight)
                return
            self._load_fused_module_from_checkpoint(param, loaded_weight)
            return
        assert loaded_shard_id < len(self.output_sizes)
        shard_offset, shard_size = self._rank_slice(loaded_shard_id)
        param.load_merged_column_weight(loaded_weight=loaded_weight, shard_id=loaded_shard_id,
                                        shard_offset=shard_offset, shard_size=shard_size,
                                        use_presharded_weights=self.use_presharded_weights,
                                        tp_rank=self.tp_rank, tp_size=self.tp_size)


class QKVParallelLinear(ColumnParallelLinear):
    """Fused q/k/v projection, parallelised over heads; when there are fewer kv heads than ranks
    each kv head is replicated on tp_size / kv_heads ranks (reference :778-1209, head math :831-848)."""

    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int,
                 total_num_kv_heads: Optional[int] = None, bias: bool = True, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None, quant_config: Optional[QuantizationConfig] = None,
                 prefix: str = "", tp_rank: Optional[int] = None, tp_size: Optional[int] = None,
                 load_presharded_attn: bool = False):
        self.hidden_size = hidden_size
        self.head_size = head_size
        self.total_num_heads = total_num_heads
        self.total_num_kv_heads = total_num_kv_heads if total_num_kv_heads is not None else total_num_heads
        tp_rank, tp_size = _tp(tp_rank, tp_size)
        self.num_heads = divide(self.total_num_heads, tp_size)
        if tp_size >= self.total_num_kv_heads:
            self.num_kv_heads = 1
            self.num_kv_head_replicas = divide(tp_size, self.total_num_kv_heads)
        else:
            self.num_kv_heads = divide(self.total_num_kv_heads, tp_size)
            self.num_kv_head_replicas = 1
        self.q_proj_shard_size = self.num_heads * self.head_size
        self.kv_proj_shard_size = self.num_kv_heads * self.head_size
        output_size = (self.num_heads + 2 * self.num_kv_heads) * tp_size * self.head_size
        self.output_sizes = [
            self.num_heads * self.head_size * tp_size,     # q_proj
            self.num_kv_heads * self.head_size * tp_size,  # k_proj
            self.num_kv_heads * self.head_size * tp_size,  # v_proj
        ]
        super().__init__(input_size=hidden_size, output_size=output_size, bias=bias, gather_output=False,
                         skip_bias_add=skip_bias_add, params_dtype=params_dtype, quant_config=quant_config,
                         prefix=prefix, tp_rank=tp_rank, tp_size=tp_size,
                         use_presharded_weights=load_presharded_attn)

    def _get_shard_offset_mapping(self, loaded_shard_id: str):
        return {
            "q": 0,
            "k": self.num_heads * self.head_size,
            "v": (self.num_heads + self.num_kv_heads) * self.head_size,
            "total": (self.num_heads + 2 * self.num_kv_heads) * self.head_size,
        }.get(loaded_shard_id)

    def _get_shard_size_mapping(self, loaded_shard_id: str):
        return {
            "q": self.num_heads * self.head_size,
            "k": self.num_kv_heads * self.head_size,
            "v": self.num_kv_heads * self.head_size,
        }.get(loaded_shard_id)

    def _checkpoint_slices(self):
        q = self.total_num_heads * self.head_size
        kv = self.total_num_kv_heads * self.head_size
        return [("q", 0, q), ("k", q, kv), ("v", q + kv, kv)]

    def _load_fused_module_from_checkpoint(self, param: BasevLLMParameter, loaded_weight: torch.Tensor):
        for shard_id, shard_offset, shard_size in self._checkpoint_slices():
            if isinstance(param, PackedvLLMParameter) and param.packed_dim == param.output_dim:
                shard_size, shard_offset = param.adjust_shard_indexes_for_packing(shard_size=shard_size,
                                                                                 shard_offset=shard_offset)
            self.weight_loader_v2(param, loaded_weight.narrow(param.output_dim, shard_offset, shard_size), shard_id)

    def weight_loader_v2(self, param: BasevLLMParameter, loaded_weight: torch.Tensor,
                         loaded_shard_id: Optional[str] = None):
        if loaded_shard_id is None:
            if type(param) in (RowvLLMParameter, BasevLLMParameter):
                param.load_qkv_weight(loaded_weight=loaded_weight)
                return
            self._load_fused_module_from_checkpoint(param, loaded_weight)
            return
        assert loaded_shard_id in ["q", "k", "v"]
        param.load_qkv_weight(loaded_weight=loaded_weight, num_heads=self.num_kv_head_replicas,
                              shard_id=loaded_shard_id, shard_offset=self._get_shard_offset_mapping(loaded_shard_id),
                              shard_size=self._get_shard_size_mapping(loaded_shard_id), tp_rank=self.tp_rank,
                              use_presharded_weights=self.use_presharded_weights)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor, loaded_shard_id: Optional[str] = None):
        """Plain tensors (bias, unquantized weight)."""
        output_dim = getattr(param, "output_dim", None)
        if loaded_shard_id is None:
            if output_dim is None:
                assert param.data.shape == loaded_weight.shape
                param.data.copy_(loaded_weight)
                return
            for shard_id, off, size in self._checkpoint_slices():
                self.weight_loader(param, loaded_weight.narrow(output_dim, off, size), shard_id)
            return
        assert loaded_shard_id in ["q", "k", "v"]
        param_data = param.data
        if output_dim is not None:
            shard_offset = self._get_shard_offset_mapping(loaded_shard_id)
            shard_size = self._get_shard_size_mapping(loaded_shard_id)
            param_data = param_data.narrow(output_dim, shard_offset, shard_size)
            index = self.tp_rank if loaded_shard_id == "q" else self.tp_rank // self.num_kv_head_replicas
            if not self.use_presharded_weights:
                loaded_weight = loaded_weight.narrow(output_dim, index * shard_size, shard_size)
        assert param_data.shape == loaded_weight.shape
        param_data.copy_(loaded_weight)


class RowParallelLinear(LinearBase):
    """Y = X A + b with A split along its input dimension and X along its last dimension; the
    per-rank partial products are summed with one all-reduce (reference :1212-1422)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, input_is_parallel: bool = True,
                 skip_bias_add: bool = False, params_dtype: Optional[torch.dtype] = None,
                 reduce_results: bool = True, quant_config: Optional[QuantizationConfig] = None, prefix: str = "",
                 tp_rank: Optional[int] = None, tp_size: Optional[int] = None, use_presharded_weights: bool = False):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix)
        self.input_is_parallel = input_is_parallel
        self.reduce_results = reduce_results
        self.tp_rank, self.tp_size = _tp(tp_rank, tp_size)
        self.input_size_per_partition = divide(input_size, self.tp_size)
        assert self.quant_method is not None
        self.use_presharded_weights = use_presharded_weights
        self.quant_method.create_weights(
            layer=self, input_size_per_partition=self.input_size_per_partition,
            output_partition_sizes=[self.output_size], input_size=self.input_size, output_size=self.output_size,
            params_dtype=self.params_dtype, weight_loader=self._loader_for_method())
        self._make_bias(self.output_size, params_dtype, bias)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        input_dim = getattr(param, "input_dim", None)
        param_data = param.data
        if input_dim is not None and not self.use_presharded_weights:
            shard = param_data.shape[input_dim]
            loaded_weight = loaded_weight.narrow(input_dim, self.tp_rank * shard, shard)
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        assert param_data.shape == loaded_weight.shape
        param_data.copy_(loaded_weight)

    def weight_loader_v2(self, param: BasevLLMParameter, loaded_weight: torch.Tensor):
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        param.load_row_parallel_weight(loaded_weight, tp_rank=self.tp_rank,
                                       use_presharded_weights=self.use_presharded_weights)

    def forward(self, input_, skip_all_reduce: bool = False):
        if self.input_is_parallel:
            input_parallel = input_
        else:
            input_parallel = split_tensor_along_last_dim(input_, num_partitions=self.tp_size)[self.tp_rank].contiguous()
        # bias is fused into the GEMM epilogue on rank 0 only, so it is added exactly once
        bias_ = None if (self.tp_rank > 0 or self.skip_bias_add) else self.bias
        if self.reduce_results and self.tp_size > 1 and not skip_all_reduce:
            if hasattr(self.quant_method, "apply_all_reduce"):  # linear + collective in one call where the method has it
                output = self.quant_method.apply_all_reduce(self, input_parallel, bias=bias_)
            else:
                output = tensor_model_parallel_all_reduce(self.quant_method.apply(self, input_parallel, bias=bias_))
        else:
            output = self.quant_method.apply(self, input_parallel, bias=bias_)
        return output, (self.bias if self.skip_bias_add else None)

    def extra_repr(self) -> str:
        return (f"input_features={self.input_size_per_partition}, output_features={self.output_size}, "
                f"bias={self.bias is not None}, tp_size={self.tp_size}, reduce_results={self.reduce_results}")
