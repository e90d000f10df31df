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
        from ..moe import FusedMoE

        if isinstance(layer, FusedMoE):
            return AWQMoEMethod(self)
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
        G = K // max(1, scales.shape[0])
        tileable = (qweight.is_cuda and K % _TILE_GROUP == 0 and scales.shape[0] * G == K
                    and (G in (32, 64) or G % _TILE_GROUP == 0) and scales.dtype in (torch.float16, torch.bfloat16))
        if not tileable:
            # other group sizes (or a CPU-resident model that is never run): keep the reference
            # layout; apply() then uses awq_dequantize + matmul exactly as the reference does
            if qweight.is_cuda:
                logger.warning("AWQ group size %d with K = %d: using the unfused dequant+matmul path (the fused B200 "
                               "kernels take group sizes 32, 64 and multiples of 128, K %% 128 == 0)", G, K)
            layer.qweight = torch.nn.Parameter(qweight, requires_grad=False)
            layer.qzeros = torch.nn.Parameter(qzeros, requires_grad=False)
            layer.scales = torch.nn.Parameter(scales, requires_grad=False)
            return
        # MLP gate_up_proj whose owner asked for the fused activation (layer.awq_fuse_silu_mul, set by the model
        # before this hook): gate / up rows interleaved by 8 so the kernels' epilogue holds both halves of a column
        sizes = list(getattr(layer, "output_partition_sizes", []) or [])
        gate_up = bool(getattr(layer, "awq_fuse_silu_mul", False)) and len(sizes) == 2 and sizes[0] == sizes[1] \
            and sizes[0] % 8 == 0 and getattr(layer, "bias", None) is None and G == _TILE_GROUP
        layer.awq_gate_up = gate_up
        repack = _kernels().awq_b200_repack_gate_up if gate_up else _kernels().awq_b200_repack
        tiles = repack(qweight.contiguous(), scales.contiguous(), qzeros.contiguous())
        scales_view, zeros_view = AWQLinearMethod._meta_views(tiles, scales.dtype)
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

    def apply_all_reduce(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`apply` followed by the TP all-reduce, as RowParallelLinear.forward runs them (reference linear.py:1388-1414).
        Decode-sized results of tiled layers go through awq_b200_linear_allreduce (partial sums pushed to the peers from
        the linear's epilogue); everything else is apply -> tensor_model_parallel_all_reduce."""
        from ...distributed import get_tp_group

        tp = get_tp_group()
        N = layer.awq_out_features
        if (tp.world_size > 1 and tp.peer_ar is not None and getattr(layer, "awq_layout", "awq") == "tiled"
                and not getattr(layer, "awq_gate_up", False) and layer.qweight.shape[1] == _TILE_BYTES // 4
                and x.dtype == layer.awq_dtype):
            x2 = x.reshape(-1, x.shape[-1])
            M = x2.shape[0]
            # the gate of PeerAllReduce.should_use, on quantities every rank shares (M, N, dtype)
            if 0 < M * N * x.element_size() <= tp.peer_ar.max_message_bytes and (M * N) % 8 == 0:
                y = tp.peer_ar.linear_all_reduce(x2.contiguous(), layer.qweight, bias, N)
                return y.reshape(x.shape[:-1] + (N,))
        return tp.all_reduce(self.apply(layer, x, bias))

    def apply_silu_mul(self, layer: torch.nn.Module, x: torch.Tensor) -> torch.Tensor:
        """gate_up_proj followed by SiluAndMul (reference: python/sglang/srt/models/llama.py:94-106 calls the linear,
        then `self.act_fn`, activation.py SiluAndMul): one launch when the layer was repacked for it, else the
        linear and the activation back to back.  Returns [..., intermediate / tp]."""
        if getattr(layer, "awq_layout", "awq") == "tiled" and getattr(layer, "awq_gate_up", False):
            if x.dtype != layer.awq_dtype:
                raise RuntimeError(f"expected activations of dtype {layer.awq_dtype}, got {x.dtype}")
            return _kernels().awq_b200_linear_silu_mul(x, layer.qweight, layer.awq_out_features)
        return _kernels().silu_and_mul(self.apply(layer, x, None).contiguous())

    @staticmethod
    def gate_up_perm(layer: torch.nn.Module) -> torch.Tensor:
        """perm[n'] = logical output column of physical row n' of a gate_up-interleaved tile array."""
        inter = layer.awq_out_features // 2
        n = torch.arange(2 * inter, device=layer.qweight.device)
        s_, p_ = n // 16, n % 16
        return torch.where(p_ < 8, 8 * s_ + p_, inter + 8 * s_ + (p_ - 8))

    @staticmethod
    def _meta_views(tiles: torch.Tensor, dtype: torch.dtype):
        """strided views of the scales ([T, 8 * sections, 16] 16-bit) and zero points ([T, 8 * sections, 16] uint8) inside
        the records: one 48-byte block per 16 rows per section (sections = 1 for group sizes >= 128, 2 / 4 for 64 / 32)"""
        T, rec = tiles.shape[0], tiles.shape[1] * 4
        blocks = (rec - 8192) // 48
        return (tiles.view(dtype).as_strided((T, blocks, 16), (rec // 2, 24, 1), 8192 // 2),
                tiles.view(torch.uint8).as_strided((T, blocks, 16), (rec, 48, 1), 8192 + 32))

    @staticmethod
    def retile_views(layer: torch.nn.Module) -> None:
        """Re-derive the `scales` / `qzeros` views from `qweight` after an operation that
        de-aliased them (module.to(), copy.deepcopy)."""
        if getattr(layer, "awq_layout", "awq") != "tiled":
            return
        sv, zv = AWQLinearMethod._meta_views(layer.qweight.data, layer.awq_dtype)
        layer.scales = torch.nn.Parameter(sv, requires_grad=False)
        layer.qzeros = torch.nn.Parameter(zv, requires_grad=False)

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

        from ..moe import FusedMoE

        if isinstance(layer, FusedMoE):  # reference awq.py:300-305
            return AWQMoEMethod(self)
        if isinstance(layer, LinearBase):  # (the ParallelLMHead branch of the reference: out of scope)
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


class AWQMoEMethod:
    """AWQ mixture-of-experts method; reference awq.py:661-852 (`AWQMoEMethod(FusedMoEMethodBase)`: same parameter names,
    shapes and loader attributes; the Marlin repack / permutes of its process_weights_after_loading and the
    fused_marlin_moe runner of its apply are replaced by the B200 tile repack and the kernels of csrc/awq_moe.cu)."""

    def __init__(self, quant_config):
        self.quant_config = quant_config
        if self.quant_config.weight_bits != 4:
            raise ValueError("AWQMoEMethod only supports 4bit now.")

    def create_weights(self, layer: torch.nn.Module, num_experts: int, hidden_size: int,
                       intermediate_size_per_partition: int, params_dtype: torch.dtype, **extra_weight_attrs):
        pack, G = self.quant_config.pack_factor, self.quant_config.group_size
        if hidden_size % G or intermediate_size_per_partition % G:
            raise ValueError("The input size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        extra_weight_attrs.update({"is_transposed": True, "quant_method": "group"})
        ipp = intermediate_size_per_partition

        def reg(name, shape, dtype):
            prm = torch.nn.Parameter(torch.empty(*shape, dtype=dtype), requires_grad=False)
            layer.register_parameter(name, prm)
            for k, v in extra_weight_attrs.items():
                setattr(prm, k, v)

        reg("w13_qweight", (num_experts, hidden_size, 2 * ipp // pack), torch.int32)
        reg("w2_qweight", (num_experts, ipp, hidden_size // pack), torch.int32)
        reg("w13_scales", (num_experts, hidden_size // G, 2 * ipp), params_dtype)
        reg("w2_scales", (num_experts, ipp // G, hidden_size), params_dtype)
        reg("w13_qzeros", (num_experts, hidden_size // G, 2 * ipp // pack), torch.int32)
        reg("w2_qzeros", (num_experts, ipp // G, hidden_size // pack), torch.int32)
        layer.awq_moe_sizes = (num_experts, hidden_size, ipp)

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        """per expert: W13 -> tile records with gate / up rows interleaved (SiluAndMul becomes the epilogue), W2 -> tile
        records; stacked [E, T, 2144].  `*_scales` / `*_qzeros` become strided views into the records (the dummy
        loader's in-place float initialisation still lands in the bytes the kernels read).  Idempotent."""
        if getattr(layer, "awq_layout", "awq") == "tiled":
            return
        if self.quant_config.group_size != _TILE_GROUP or not layer.w13_qweight.is_cuda:
            raise NotImplementedError("AWQ MoE on B200 needs group size 128 and CUDA-resident weights")
        k = _kernels()
        E = layer.w13_qweight.shape[0]
        dt = layer.w13_scales.dtype
        for name, repack in (("w13", k.awq_b200_repack_gate_up), ("w2", k.awq_b200_repack)):
            qw, sc, qz = (getattr(layer, f"{name}_{s}").data for s in ("qweight", "scales", "qzeros"))
            tiles = torch.stack([repack(qw[e].contiguous(), sc[e].contiguous(), qz[e].contiguous()) for e in range(E)])
            T = tiles.shape[1]
            setattr(layer, f"{name}_qweight", torch.nn.Parameter(tiles, requires_grad=False))
            setattr(layer, f"{name}_scales", torch.nn.Parameter(
                tiles.view(dt).as_strided((E, T, 8, 16), (T * _TILE_BYTES // 2, _TILE_BYTES // 2, 24, 1), 8192 // 2),
                requires_grad=False))
            setattr(layer, f"{name}_qzeros", torch.nn.Parameter(
                tiles.view(torch.uint8).as_strided((E, T, 8, 16), (T * _TILE_BYTES, _TILE_BYTES, 48, 1), 8192 + 32),
                requires_grad=False))
        layer.awq_layout = "tiled"
        layer.awq_dtype = dt

    # up to this many tokens (and 1024 pairs) take the streaming kernel (device-side routing, capturable); larger batches
    # gather each expert's tokens on the host and run the dense kernels (tcgen05 GEMM) on them
    STREAMING_MAX_TOKENS = 64

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor) -> torch.Tensor:
        """x [M, hidden], topk_weights / topk_ids [M, top_k] (the router's output; reference: StandardDispatchOutput.
        topk_output) -> [M, hidden] = sum_k w_k * W2_e(silu(W13g_e x) * W13u_e x), partial sums under TP."""
        assert getattr(layer, "awq_layout", "awq") == "tiled", "process_weights_after_loading has not run"
        if x.dtype != layer.awq_dtype:
            raise RuntimeError(f"expected activations of dtype {layer.awq_dtype}, got {x.dtype}")
        k = _kernels()
        E, hidden, ipp = layer.awq_moe_sizes
        M, topk = topk_ids.shape
        x = x.reshape(M, hidden).contiguous()
        ids = topk_ids.reshape(-1).to(torch.int32).contiguous()
        w = topk_weights.reshape(-1).to(torch.float32).contiguous()
        if M * topk <= 1024 and (M <= self.STREAMING_MAX_TOKENS or torch.cuda.is_current_stream_capturing()):
            h = k.awq_b200_moe_linear(x, layer.w13_qweight, ids, None, topk, 2 * ipp, True)
            o = k.awq_b200_moe_linear(h, layer.w2_qweight, ids, w, 1, hidden, False)
            return k.awq_b200_moe_sum(o, topk)
        # grouped: tokens of one expert at a time through the dense kernels (tcgen05 GEMM for the big groups)
        order = torch.argsort(ids, stable=True)
        sorted_ids = ids[order]
        counts = torch.bincount(sorted_ids.clamp_min(0), minlength=E)
        skipped = int((sorted_ids < 0).sum())  # not routed to this rank (expert parallelism)
        counts[0] -= skipped
        counts = counts.tolist()
        xs = x[order // topk]
        o_sorted = torch.zeros((M * topk, hidden), dtype=x.dtype, device=x.device)
        lo = skipped
        for e, n in enumerate(counts):
            if n == 0:
                continue
            seg = slice(lo, lo + n)
            h = k.awq_b200_linear_silu_mul(xs[seg], layer.w13_qweight[e], 2 * ipp)
            o_sorted[seg] = (k.awq_b200_linear(h, layer.w2_qweight[e], None, hidden).float() * w[order[seg]].unsqueeze(1)).to(x.dtype)
            lo += n
        o = torch.empty_like(o_sorted)
        o[order] = o_sorted
        return k.awq_b200_moe_sum(o, topk)
