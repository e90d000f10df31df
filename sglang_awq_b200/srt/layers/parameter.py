"""TP-aware parameter classes used by the AWQ linear path.

Same names, attributes and loading semantics as the reference's
python/sglang/srt/layers/parameter.py (BasevLLMParameter :30, _ColumnvLLMParameter :74-223,
RowvLLMParameter :226-286, GroupQuantScaleParameter, PackedvLLMParameter :433-474,
_adjust_shard_indexes_for_packing :539-550), restricted to what AWQ needs: a checkpoint tensor
is narrowed to this rank's shard along the output (column-parallel) or input (row-parallel)
dimension; for packed int32 tensors the offsets along the packed dimension are divided by the
pack factor (8 nibbles per word).
"""
from __future__ import annotations

from fractions import Fraction
from typing import Callable, Optional, Union

import torch
from torch.nn import Parameter

__all__ = [
    "BasevLLMParameter",
    "PackedvLLMParameter",
    "ModelWeightParameter",
    "GroupQuantScaleParameter",
    "RowvLLMParameter",
]


def _shard(t: torch.Tensor, dim: int, index: int, size: int) -> torch.Tensor:
    """Rank-local slice [index*size, (index+1)*size) of `t` along `dim`."""
    start = index * size
    if start + size > t.shape[dim]:
        raise ValueError(f"shard [{start}, {start + size}) exceeds dimension {dim} of size {t.shape[dim]}")
    return t.narrow(dim, start, size)


class BasevLLMParameter(Parameter):
    """A Parameter that carries its own weight loader (reference parameter.py:30-70)."""

    def __new__(cls, data: torch.Tensor, **kwargs):
        return super().__new__(cls, data=data, requires_grad=False)

    def __init__(self, data: torch.Tensor, weight_loader: Callable, **kwargs):
        self._weight_loader = weight_loader

    @property
    def weight_loader(self) -> Callable:
        return self._weight_loader

    def _copy_checked(self, dst: torch.Tensor, src: torch.Tensor) -> None:
        if dst.shape != src.shape:
            raise AssertionError(f"shape mismatch loading weight: param {tuple(dst.shape)} vs tensor {tuple(src.shape)}")
        dst.copy_(src)

    def load_column_parallel_weight(self, loaded_weight: torch.Tensor, **kwargs):
        self._copy_checked(self.data, loaded_weight)

    def load_row_parallel_weight(self, loaded_weight: torch.Tensor, **kwargs):
        self._copy_checked(self.data, loaded_weight)

    def load_merged_column_weight(self, loaded_weight: torch.Tensor, **kwargs):
        self._copy_checked(self.data, loaded_weight)

    def load_qkv_weight(self, loaded_weight: torch.Tensor, **kwargs):
        self._copy_checked(self.data, loaded_weight)


class _ColumnvLLMParameter(BasevLLMParameter):
    """Sharded along `output_dim` (reference parameter.py:74-223)."""

    def __init__(self, output_dim: int, **kwargs):
        self._output_dim = output_dim
        super().__init__(**kwargs)

    @property
    def output_dim(self) -> int:
        return self._output_dim

    def _packed_on_output(self) -> bool:
        return getattr(self, "packed_dim", None) == self.output_dim

    # reference :93-124
    def load_column_parallel_weight(self, loaded_weight: torch.Tensor, tp_rank: int = 0,
                                    use_presharded_weights: bool = False):
        if not use_presharded_weights:
            loaded_weight = _shard(loaded_weight, self.output_dim, tp_rank, self.data.shape[self.output_dim])
        self._copy_checked(self.data, loaded_weight)

    # reference :126-173 -- one logical matrix (gate or up) of a merged layer
    def load_merged_column_weight(self, loaded_weight: torch.Tensor, **kwargs):
        shard_offset, shard_size = kwargs.get("shard_offset"), kwargs.get("shard_size")
        tp_rank = kwargs.get("tp_rank", 0)
        if self._packed_on_output():
            shard_size, shard_offset = self.adjust_shard_indexes_for_packing(shard_offset=shard_offset,
                                                                             shard_size=shard_size)
        dst = self.data.narrow(self.output_dim, shard_offset, shard_size)
        if not kwargs.get("use_presharded_weights"):
            loaded_weight = _shard(loaded_weight, self.output_dim, tp_rank, shard_size)
        self._copy_checked(dst, loaded_weight)

    # reference :175-223 -- q, k or v of a fused QKV layer; kv heads may be replicated
    def load_qkv_weight(self, loaded_weight: torch.Tensor, tp_rank: int = 0, use_presharded_weights: bool = False,
                        **kwargs):
        shard_offset, shard_size = kwargs.get("shard_offset"), kwargs.get("shard_size")
        shard_id, num_heads = kwargs.get("shard_id"), kwargs.get("num_heads")
        if self._packed_on_output():
            shard_size, shard_offset = self.adjust_shard_indexes_for_packing(shard_offset=shard_offset,
                                                                             shard_size=shard_size)
        dst = self.data.narrow(self.output_dim, shard_offset, shard_size)
        index = tp_rank if shard_id == "q" else tp_rank // num_heads  # num_heads = kv-head replicas
        if not use_presharded_weights:
            loaded_weight = _shard(loaded_weight, self.output_dim, index, shard_size)
        self._copy_checked(dst, loaded_weight)


class RowvLLMParameter(BasevLLMParameter):
    """Sharded along `input_dim` (reference parameter.py:226-286)."""

    def __init__(self, input_dim: int, **kwargs):
        self._input_dim = input_dim
        super().__init__(**kwargs)

    @property
    def input_dim(self) -> int:
        return self._input_dim

    def load_row_parallel_weight(self, loaded_weight: torch.Tensor, tp_rank: int = 0,
                                 use_presharded_weights: bool = False):
        if not use_presharded_weights:
            loaded_weight = _shard(loaded_weight, self.input_dim, tp_rank, self.data.shape[self.input_dim])
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        self._copy_checked(self.data, loaded_weight)


class ModelWeightParameter(_ColumnvLLMParameter, RowvLLMParameter):
    """A weight that can be split either way."""


class GroupQuantScaleParameter(_ColumnvLLMParameter, RowvLLMParameter):
    """Per-group scales [K/G, N]: rows follow the input split, columns the output split."""


def _adjust_shard_indexes_for_packing(shard_size, shard_offset, packed_factor, marlin_tile_size=None):
    """reference parameter.py:539-550."""
    shard_size = shard_size // packed_factor
    shard_offset = shard_offset // packed_factor
    if marlin_tile_size is not None:
        return shard_size * marlin_tile_size, shard_offset * marlin_tile_size
    return shard_size, shard_offset


class PackedvLLMParameter(ModelWeightParameter):
    """int32 words holding `packed_factor` values along `packed_dim` (reference :433-474)."""

    def __init__(self, packed_factor: Union[int, Fraction], packed_dim: int,
                 marlin_tile_size: Optional[int] = None, **kwargs):
        self._packed_factor = packed_factor
        self._packed_dim = packed_dim
        self._marlin_tile_size = marlin_tile_size
        super().__init__(**kwargs)

    @property
    def packed_dim(self) -> int:
        return self._packed_dim

    @property
    def packed_factor(self):
        return self._packed_factor

    @property
    def marlin_tile_size(self):
        return self._marlin_tile_size

    def adjust_shard_indexes_for_packing(self, shard_size, shard_offset):
        return _adjust_shard_indexes_for_packing(shard_size=shard_size, shard_offset=shard_offset,
                                                 packed_factor=self.packed_factor,
                                                 marlin_tile_size=self.marlin_tile_size)
