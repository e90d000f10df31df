"""AWQ quantization plugin: AWQConfig + AWQLinearMethod on the B200 kernels.

Drop-in for the reference's python/sglang/srt/layers/quantization/awq.py:89-179 (AWQConfig)
and :352-451 (AWQLinearMethod): same class names (the layers pick `weight_loader_v2` by the
string "AWQLinearMethod", linear.py:51-69), same parameters registered by `create_weights`
(qweight / qzeros / scales with the same shapes, dtypes and loader attributes), same `apply`
contract.  What changes is underneath:

  reference apply  = awq_dequantize -> torch.matmul -> add_        (3 launches, 4.5 B/weight)
  this apply       = ONE fused kernel reading the packed tiles      (0.53 B/weight)

`process_weights_after_loading` -- a no-op re-wrap in the reference (:429-432) -- performs the
one-time repack into the tile layout (sglang_awq_b200/csrc/awq_repack.cu).

AWQMarlinConfig / AWQMarlinLinearMethod (reference awq.py:182-349, :454-600) -- what an AWQ
checkpoint resolves to BY DEFAULT in the reference (`override_quantization_method`) -- are the same
front-end over the same B200 kernels: where the reference repacks to Marlin's layout and calls
gptq_marlin_gemm (mma.sync), this repacks to the tile records and calls the fused kernel.  Out of
scope here: AWQMoE*, Ascend variants, 8-bit Marlin weights.
"""
from __future__ import annotations

import logging
from typing import Any, Dict, List, Optional

import torch

from ..parameter import GroupQuantScaleParameter, PackedvLLMParameter
from .base_config import LinearMethodBase, QuantizationConfig
from .marlin_utils import (check_marlin_supported, check_marlin_supports_layer, verify_marlin_supported,
                           verify_marlin_supports_shape)
from .unquant import UnquantizedLinearMethod

logger = logging.getLogger(__name__)

_TILE_BYTES = 8576
_TILE_GROUP = 128


def _kernels():
    """The compiled ops.  Imported lazily so that building a model definition (create_weights,
    weight loading) works in a CPU-only process; anything that computes needs the library and
    raises if it is missing -- there is no PyTorch/CPU fallback."""
    from sglang_awq_b200 import sgl_kernel

    return sgl_kernel


def is_layer_skipped_awq(prefix: str, modules_to_not_convert: List[str]) -> bool:
    return any(module_name in prefix for module_name in modules_to_not_convert)


class AWQConfig(QuantizationConfig):
    """Config class for AWQ (https://arxiv.org/abs/2306.00978); reference awq.py:89-179."""

    def __init__(self, weight_bits: int, group_size: int, zero_point: bool,
                 modules_to_not_convert: Optional[List[str]] = None) -> None:
        super().__init__()
        self.weight_bits = weight_bits
        self.group_size = group_size
        self.zero_point = zero_point
        self.modules_to_not_convert = modules_to_not_convert or []
        if self.weight_bits != 4:
            raise ValueError("Currently, only 4-bit weight quantization is supported for "
                             f"AWQ, but got {self.weight_bits} bits.")
        self.pack_factor = 32 // self.weight_bits

    def __repr__(self) -> str:
        return (f"AWQConfig(weight_bits={self.weight_bits}, group_size={self.group_size}, "
                f"zero_point={self.zero_point}, modules_to_not_convert={self.modules_to_not_convert})")

    def get_scaled_act_names(self) -> List[str]:
        return []

    def get_name(self) -> str:
        return "awq"

    def get_supported_act_dtypes(self) -> List[torch.dtype]:
        return [torch.float16]

    @classmethod
    def get_min_capability(cls) -> int:
        # the reference's CUDA path needs Turing (75); these kernels are sm_100a only
        return 100

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quant_config.json", "quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "AWQConfig":
        weight_bits = cls.get_from_keys(config, ["w_bit", "bits"])
        group_size = cls.get_from_keys(config, ["q_group_size", "group_size"])
        zero_point = cls.get_from_keys(config, ["zero_point"])
        modules_to_not_convert = cls.get_from_keys_or(config, ["modules_to_not_convert"], None)
        return cls(weight_bits, group_size, zero_point, modules_to_not_convert)

    def get_quant_method(self, layer: torch.nn.Module, prefix: str) -> Optional[LinearMethodBase]:
        from ..linear import LinearBase

        if isinstance(layer, LinearBase):
            if is_layer_skipped_awq(prefix, self.modules_to_not_convert):
                return UnquantizedLinearMethod()
            return AWQLinearMethod(self)
        return None


class AWQLinearMethod(LinearMethodBase):
    """Linear method for AWQ; reference awq.py:352-451."""

    def __init__(self, quant_config: AWQConfig):
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        group_size = self.quant_config.group_size
        if group_size == -1:
            group_size = input_size_per_partition
        if input_size_per_partition % group_size != 0:
            raise ValueError("The input size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        output_size_per_partition = sum(output_partition_sizes)
        pack = self.quant_config.pack_factor
        if output_size_per_partition % pack != 0:
            raise ValueError("The output size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        weight_loader = extra_weight_attrs.get("weight_loader")
        groups = input_size_per_partition // group_size
        qweight = PackedvLLMParameter(
            data=torch.empty(input_size_per_partition, output_size_per_partition // pack, dtype=torch.int32),
            input_dim=0, output_dim=1, packed_dim=1, packed_factor=pack, weight_loader=weight_loader)
        qzeros = PackedvLLMParameter(
            data=torch.empty(groups, output_size_per_partition // pack, dtype=torch.int32),
            input_dim=0, output_dim=1, packed_dim=1, packed_factor=pack, weight_loader=weight_loader)
        scales = GroupQuantScaleParameter(
            data=torch.empty(groups, output_size_per_partition, dtype=params_dtype),
            input_dim=0, output_dim=1, weight_loader=weight_loader)
        layer.register_parameter("qweight", qweight)
        layer.register_parameter("qzeros", qzeros)
        layer.register_parameter("scales", scales)
        layer.awq_in_features = input_size_per_partition
        layer.awq_out_features = output_size_per_partition
        layer.awq_layout = "awq"

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        """One-time repack into the B200 tile layout (kernel 4).

        The three parameter NAMES survive (other code and the dummy-weight initialiser walk
        them): `qweight` becomes the record array int32 [T, 2144]; `scales` and `qzeros` become
        strided VIEWS into the same storage, so a float-only in-place initialiser
        (weight_utils.py:1108-1137, run AFTER this hook by the dummy loader) still lands in the
        right bytes.  Idempotent: a second call on an already tiled layer is a no-op.

        Because `scales` ([T, 8, 16] 16-bit) and `qzeros` ([T, 8, 16] uint8, one byte per zero
        point) ALIAS `qweight`'s storage, they are not the reference's shapes any more: a
        state_dict of a tiled layer only loads back into a tiled layer of the same shape (in
        place, which keeps the aliasing; tests/test_gpu_layers.py), and `.to()` / deepcopy of a
        tiled layer must be followed by re-deriving the views (`retile_views`)."""
        if getattr(layer, "awq_layout", "awq") == "tiled":
            return
        qweight, qzeros, scales = layer.qweight.data, layer.qzeros.data, layer.scales.data
        K, N = layer.awq_in_features, layer.awq_out_features
        tileable = (qweight.is_cuda and K % _TILE_GROUP == 0 and scales.shape[0] * _TILE_GROUP == K
                    and scales.dtype in (torch.float16, torch.bfloat16))
        if not tileable:
            # other group sizes (or a CPU-resident model that is never run): keep the reference
            # layout; apply() then uses awq_dequantize + matmul exactly as the reference does
            if qweight.is_cuda and scales.shape[0] * _TILE_GROUP != K:
                logger.warning("AWQ group size %d: using the unfused dequant+matmul path (the fused B200 "
                               "kernels need group size 128)", K // max(1, scales.shape[0]))
            layer.qweight = torch.nn.Parameter(qweight, requires_grad=False)
            layer.qzeros = torch.nn.Parameter(qzeros, requires_grad=False)
            layer.scales = torch.nn.Parameter(scales, requires_grad=False)
            return
        # MLP gate_up_proj whose owner asked for the fused activation (layer.awq_fuse_silu_mul, set by the model
        # before this hook): gate / up rows interleaved by 8 so the kernels' epilogue holds both halves of a column
        sizes = list(getattr(layer, "output_partition_sizes", []) or [])
        gate_up = bool(getattr(layer, "awq_fuse_silu_mul", False)) and len(sizes) == 2 and sizes[0] == sizes[1] \
            and sizes[0] % 8 == 0 and getattr(layer, "bias", None) is None
        layer.awq_gate_up = gate_up
        repack = _kernels().awq_b200_repack_gate_up if gate_up else _kernels().awq_b200_repack
        tiles = repack(qweight.contiguous(), scales.contiguous(), qzeros.contiguous())
        T = tiles.shape[0]
        as16 = tiles.view(scales.dtype)   # [T, 4288]
        as8 = tiles.view(torch.uint8)     # [T, 8576]
        scales_view = as16.as_strided((T, 8, 16), (_TILE_BYTES // 2, 24, 1), 8192 // 2)
        zeros_view = as8.as_strided((T, 8, 16), (_TILE_BYTES, 48, 1), 8192 + 32)
        layer.qweight = torch.nn.Parameter(tiles, requires_grad=False)
        layer.scales = torch.nn.Parameter(scales_view, requires_grad=False)
        layer.qzeros = torch.nn.Parameter(zeros_view, requires_grad=False)
        layer.awq_layout = "tiled"
        layer.awq_dtype = scales.dtype  # the records carry raw 16-bit scale bits with no type tag

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        k = _kernels()
        if getattr(layer, "awq_layout", "awq") == "tiled":
            if x.dtype != layer.awq_dtype:
                # the reference fails here too (torch.matmul of mismatched dtypes, awq.py:447)
                raise RuntimeError(f"expected activations of dtype {layer.awq_dtype} (the dtype the layer's scales "
                                   f"were repacked with), got {x.dtype}")
            if getattr(layer, "awq_gate_up", False):
                raise RuntimeError("this gate_up_proj was repacked for the fused SiluAndMul epilogue "
                                   "(awq_fuse_silu_mul): call forward_silu_mul / apply_silu_mul")
            # fused dequant + GEMV/GEMM (+ bias); the M-regime switch is inside the library
            return k.awq_b200_linear(x, layer.qweight, bias, layer.awq_out_features)
        # reference path (awq.py:434-451) on our bit-exact dequant kernel
        qweight, scales, qzeros = layer.qweight, layer.scales, layer.qzeros
        out_shape = x.shape[:-1] + (qweight.shape[-1] * self.quant_config.pack_factor,)
        out = torch.matmul(x.reshape(-1, x.shape[-1]), k.awq_dequantize(qweight, scales, qzeros))
        if bias is not None:
            out.add_(bias)
        return out.reshape(out_shape)

    def apply_silu_mul(self, layer: torch.nn.Module, x: torch.Tensor) -> torch.Tensor:
        """gate_up_proj followed by SiluAndMul (reference: python/sglang/srt/models/llama.py:94-106 calls the linear,
        then `self.act_fn`, activation.py SiluAndMul): one launch when the layer was repacked for it, else the
        linear and the activation back to back.  Returns [..., intermediate / tp]."""
        if getattr(layer, "awq_layout", "awq") == "tiled" and getattr(layer, "awq_gate_up", False):
            if x.dtype != layer.awq_dtype:
                raise RuntimeError(f"expected activations of dtype {layer.awq_dtype}, got {x.dtype}")
            return _kernels().awq_b200_linear_silu_mul(x, layer.qweight, layer.awq_out_features)
        gu = self.apply(layer, x, None)
        d = gu.shape[-1] // 2
        return (torch.nn.functional.silu(gu[..., :d].float()) * gu[..., d:].float()).to(gu.dtype)

    @staticmethod
    def gate_up_perm(layer: torch.nn.Module) -> torch.Tensor:
        """perm[n'] = logical output column of physical row n' of a gate_up-interleaved tile array."""
        inter = layer.awq_out_features // 2
        n = torch.arange(2 * inter, device=layer.qweight.device)
        s_, p_ = n // 16, n % 16
        return torch.where(p_ < 8, 8 * s_ + p_, inter + 8 * s_ + (p_ - 8))

    @staticmethod
    def retile_views(layer: torch.nn.Module) -> None:
        """Re-derive the `scales` / `qzeros` views from `qweight` after an operation that
        de-aliased them (module.to(), copy.deepcopy)."""
        if getattr(layer, "awq_layout", "awq") != "tiled":
            return
        tiles = layer.qweight.data
        T = tiles.shape[0]
        layer.scales = torch.nn.Parameter(
            tiles.view(layer.awq_dtype).as_strided((T, 8, 16), (_TILE_BYTES // 2, 24, 1), 8192 // 2), requires_grad=False)
        layer.qzeros = torch.nn.Parameter(
            tiles.view(torch.uint8).as_strided((T, 8, 16), (_TILE_BYTES, 48, 1), 8192 + 32), requires_grad=False)

    @staticmethod
    def dequantize(layer: torch.nn.Module) -> torch.Tensor:
        """Dense [K, N] weight of a layer in either layout (for consumers such as kv_b_proj
        absorption, reference deepseek_v2.py:3436-3449, that called awq_dequantize directly)."""
        k = _kernels()
        if getattr(layer, "awq_layout", "awq") == "tiled":
            W = k.awq_b200_dequantize_tiled(layer.qweight, layer.awq_in_features, layer.awq_out_features,
                                            layer.scales.dtype)
            if getattr(layer, "awq_gate_up", False):  # physical (slab) order -> logical [gate | up] columns
                out = torch.empty_like(W)
                out[:, AWQLinearMethod.gate_up_perm(layer)] = W
                return out
            return W
        return k.awq_dequantize(layer.qweight, layer.scales, layer.qzeros)


class AWQMarlinConfig(QuantizationConfig):
    """Config class for AWQ Marlin; reference awq.py:182-349.  Same keys, name, override rule and
    layer routing; the linear method it hands out runs on the B200 tile kernels."""

    SUPPORTED_BITS = (4,)  # reference TYPE_MAP also lists 8 (uint8), which only its no-zero-point path serves

    def __init__(self, weight_bits: int, group_size: int, zero_point: bool, lm_head_quantized: bool,
                 modules_to_not_convert: Optional[List[str]], full_config: Dict[str, Any]) -> None:
        super().__init__()
        self.pack_factor = 32 // weight_bits
        self.group_size = group_size
        self.zero_point = zero_point
        self.lm_head_quantized = lm_head_quantized
        self.weight_bits = weight_bits
        self.modules_to_not_convert = modules_to_not_convert or []
        self.full_config = full_config
        if self.weight_bits not in self.SUPPORTED_BITS:
            raise ValueError(f"Unsupported num_bits = {self.weight_bits}. Supported num_bits = {list(self.SUPPORTED_BITS)}")
        verify_marlin_supported(self.weight_bits, group_size=self.group_size, has_zp=self.zero_point)

    def __repr__(self) -> str:
        return (f"AWQMarlinConfig(quant_type=uint{self.weight_bits}, group_size={self.group_size}, "
                f"zero_point={self.zero_point}, lm_head_quantized={self.lm_head_quantized}, "
                f"modules_to_not_convert={self.modules_to_not_convert})")

    def get_scaled_act_names(self) -> List[str]:
        return []

    @classmethod
    def get_name(cls) -> str:
        return "awq_marlin"

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.half, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100  # reference: 80 (mma.sync Marlin); these kernels are sm_100a only

    @classmethod
    def get_config_filenames(cls) -> List[str]:
        return ["quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "AWQMarlinConfig":
        weight_bits = cls.get_from_keys(config, ["bits"])
        group_size = cls.get_from_keys(config, ["group_size"])
        zero_point = cls.get_from_keys(config, ["zero_point"])
        lm_head_quantized = cls.get_from_keys_or(config, ["lm_head"], default=False)
        modules_to_not_convert = cls.get_from_keys_or(config, ["modules_to_not_convert"], None)
        return cls(weight_bits, group_size, zero_point, lm_head_quantized, modules_to_not_convert, config)

    @classmethod
    def override_quantization_method(cls, hf_quant_cfg, user_quant) -> Optional[str]:
        can_convert = cls.is_awq_marlin_compatible(hf_quant_cfg)
        is_valid_user_quant = user_quant is None or user_quant == "marlin" or user_quant == "awq_marlin"
        if can_convert and is_valid_user_quant:
            logger.info("The model is convertible to %s during runtime. Using %s kernel.", cls.get_name(), cls.get_name())
            return cls.get_name()
        if can_convert and user_quant == "awq":
            logger.info("Detected that the model can run with awq_marlin, however you specified quantization=awq "
                        "explicitly, so forcing awq. Use quantization=awq_marlin for faster inference")
        return None

    def get_quant_method(self, layer: torch.nn.Module, prefix: str) -> Optional[LinearMethodBase]:
        from ..linear import LinearBase

        if isinstance(layer, LinearBase):  # (ParallelLMHead / FusedMoE branches of the reference: out of scope)
            if is_layer_skipped_awq(prefix, self.modules_to_not_convert):
                return UnquantizedLinearMethod()
            if not check_marlin_supports_layer(layer, self.group_size):
                logger.warning("Layer '%s' is not supported by AWQMarlin. Falling back to unoptimized AWQ kernels.", prefix)
                return AWQConfig.from_config(self.full_config).get_quant_method(layer, prefix)
            return AWQMarlinLinearMethod(self)
        return None

    @classmethod
    def is_awq_marlin_compatible(cls, quant_config: Dict[str, Any]) -> bool:
        quant_method = quant_config.get("quant_method", "").lower()
        num_bits = quant_config.get("bits")
        group_size = quant_config.get("group_size")
        zero_point = quant_config.get("zero_point")
        if not torch.cuda.is_available():
            return False
        if quant_method != "awq":
            return False
        if num_bits is None or group_size is None or zero_point is None:
            return False
        if num_bits not in cls.SUPPORTED_BITS:
            return False
        return check_marlin_supported(num_bits, group_size=group_size, has_zp=zero_point)


class AWQMarlinLinearMethod(AWQLinearMethod):
    """Linear method for AWQ Marlin; reference awq.py:454-600.  Registers the same three parameters
    with the same loader attributes and shape checks; the post-load hook repacks into the B200 tile
    records instead of Marlin's layout (awq_marlin_repack / marlin_permute_scales /
    awq_to_marlin_zero_points), and `apply` is the fused kernel instead of gptq_marlin_gemm."""

    def __init__(self, quant_config: AWQMarlinConfig) -> None:
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs) -> None:
        output_size_per_partition = sum(output_partition_sizes)
        group_size = self.quant_config.group_size if self.quant_config.group_size != -1 else input_size
        verify_marlin_supports_shape(output_size_per_partition=output_size_per_partition,
                                     input_size_per_partition=input_size_per_partition, input_size=input_size,
                                     group_size=group_size)
        weight_loader = extra_weight_attrs.get("weight_loader")
        pack = self.quant_config.pack_factor
        num_groups = input_size_per_partition // group_size
        qweight = PackedvLLMParameter(
            data=torch.empty(input_size_per_partition, output_size_per_partition // pack, dtype=torch.int32),
            input_dim=0, output_dim=1, packed_dim=1, packed_factor=pack, weight_loader=weight_loader)
        qzeros = PackedvLLMParameter(
            data=torch.empty(num_groups, output_size_per_partition // pack, dtype=torch.int32),
            input_dim=0, output_dim=1, packed_dim=1, packed_factor=pack, weight_loader=weight_loader)
        scales = GroupQuantScaleParameter(
            data=torch.empty(num_groups, output_size_per_partition, dtype=params_dtype),
            input_dim=0, output_dim=1, weight_loader=weight_loader)
        layer.register_parameter("qweight", qweight)
        layer.register_parameter("qzeros", qzeros)
        layer.register_parameter("scales", scales)
        layer.input_size_per_partition = input_size_per_partition
        layer.output_size_per_partition = output_size_per_partition
        layer.num_groups = num_groups
        layer.awq_in_features = input_size_per_partition
        layer.awq_out_features = output_size_per_partition
        layer.awq_layout = "awq"
    # process_weights_after_loading / apply / dequantize: inherited (tile repack, fused kernel)
