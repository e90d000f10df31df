"""Small helpers of the TP plumbing (reference: python/sglang/srt/distributed/utils.py:21-60)."""
from typing import Sequence

import torch


def ensure_divisibility(numerator: int, denominator: int) -> None:
    assert numerator % denominator == 0, f"{numerator} is not divisible by {denominator}"


def divide(numerator: int, denominator: int) -> int:
    ensure_divisibility(numerator, denominator)
    return numerator // denominator


def split_tensor_along_last_dim(tensor: torch.Tensor, num_partitions: int,
                                contiguous_split_chunks: bool = False) -> Sequence[torch.Tensor]:
    width = divide(tensor.size(-1), num_partitions)
    chunks = torch.split(tensor, width, dim=-1)
    return tuple(c.contiguous() for c in chunks) if contiguous_split_chunks else chunks
