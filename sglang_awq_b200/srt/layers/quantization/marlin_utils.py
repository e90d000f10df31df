"""Support checks of the AWQ-Marlin front-end (reference:
python/sglang/srt/layers/quantization/marlin_utils.py:52-57 constants, :115-166 type / group-size
checks, :169-233 shape checks).  Same names, argument meaning and error text: the reference uses
them to decide whether a checkpoint resolves to `awq_marlin` and whether a layer can take the
Marlin kernel.  Here the kernel underneath is the B200 tile kernel, so the constraints that remain
are the reference's own (they define which layers a user of `awq_marlin` expects to be covered);
8-bit weights, which Marlin also serves, are not implemented."""
from __future__ import annotations

from typing import Optional, Tuple

GPTQ_MARLIN_TILE = 16
GPTQ_MARLIN_MIN_THREAD_N = 64
GPTQ_MARLIN_MIN_THREAD_K = 128
MARLIN_SUPPORTED_GROUP_SIZES = [-1, 32, 64, 128]
SUPPORTED_WEIGHT_BITS_WITH_ZP = [4]  # reference: scalar_types.uint4 only when has_zp (marlin_utils.py:101-103)


def _check_marlin_supported(weight_bits: int, group_size: Optional[int], has_zp: bool) -> Tuple[bool, Optional[str]]:
    if not has_zp or weight_bits not in SUPPORTED_WEIGHT_BITS_WITH_ZP:
        return (False, f"Marlin does not support weight_bits = {weight_bits}. Only types = uint4 are supported "
                       f"(for group_size = {group_size}, zp = {has_zp}).")
    if group_size is None or group_size not in MARLIN_SUPPORTED_GROUP_SIZES:
        return (False, f"Marlin does not support group_size = {group_size}. "
                       f"Only group_sizes = {MARLIN_SUPPORTED_GROUP_SIZES} are supported.")
    return True, None


def check_marlin_supported(weight_bits: int, group_size: int, has_zp: bool = False) -> bool:
    return _check_marlin_supported(weight_bits, group_size, has_zp)[0]


def verify_marlin_supported(weight_bits: int, group_size: int, has_zp: bool = False) -> None:
    ok, msg = _check_marlin_supported(weight_bits, group_size, has_zp)
    if not ok:
        raise ValueError(msg)


def verify_marlin_supports_shape(output_size_per_partition: int, input_size_per_partition: int, input_size: int,
                                 group_size: int) -> None:
    if output_size_per_partition % GPTQ_MARLIN_MIN_THREAD_N != 0:
        raise ValueError(f"Weight output_size_per_partition = {output_size_per_partition} is not divisible by "
                         f" min_thread_n = {GPTQ_MARLIN_MIN_THREAD_N}. Consider reducing tensor_parallel_size or "
                         "running with --quantization gptq.")
    if input_size_per_partition % GPTQ_MARLIN_MIN_THREAD_K != 0:
        raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition} is not divisible "
                         f"by min_thread_k = {GPTQ_MARLIN_MIN_THREAD_K}. Consider reducing tensor_parallel_size or "
                         "running with --quantization gptq.")
    if group_size < input_size and input_size_per_partition % group_size != 0:
        raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition} is not divisible by "
                         f"group_size = {group_size}. Consider reducing tensor_parallel_size or running with "
                         "--quantization gptq.")


def check_marlin_supports_shape(output_size_per_partition: int, input_size_per_partition: int, input_size: int,
                                group_size: int) -> Tuple[bool, Optional[str]]:
    try:
        verify_marlin_supports_shape(output_size_per_partition, input_size_per_partition, input_size, group_size)
    except ValueError as e:
        return False, str(e)
    return True, None


def check_marlin_supports_layer(layer, group_size: int) -> bool:
    out_pp = getattr(layer, "output_size_per_partition", None) or layer.output_size
    in_pp = getattr(layer, "input_size_per_partition", None) or layer.input_size
    return check_marlin_supports_shape(out_pp, in_pp, layer.input_size, group_size)[0]
