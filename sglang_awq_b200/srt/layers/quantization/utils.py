"""Helpers shared by the quantization plugins (reference:
python/sglang/srt/utils/common.py:1180-1197 set_weight_attrs,
python/sglang/srt/layers/quantization/utils.py:170-191 replace_parameter)."""
from typing import Any, Dict, Optional, Union

import torch


def set_weight_attrs(weight: torch.Tensor, weight_attrs: Optional[Dict[str, Any]]) -> None:
    """Attach loader metadata to a parameter; refuses to overwrite an existing attribute."""
    if not weight_attrs:
        return
    for key, value in weight_attrs.items():
        assert not hasattr(weight, key), f"Overwriting existing tensor attribute: {key}"
        setattr(weight, key, value)


def replace_parameter(mod: torch.nn.Module, name: str, new: Union[torch.Tensor, torch.nn.Parameter]) -> None:
    """Swap `mod.<name>` for `new`: in place when type, dtype and byte size match, else re-register."""
    old = getattr(mod, name)
    if (type(old) is type(new) and old.dtype == new.dtype
            and old.untyped_storage().nbytes() == new.untyped_storage().nbytes()):
        old.copy_(new)
        return
    data = new.data if isinstance(new, torch.nn.Parameter) else new
    mod.register_parameter(name, torch.nn.Parameter(data, requires_grad=False))
