"""`--load-format dummy` for modules built from the AWQ linear layers.

Mirrors the two steps of the reference's DummyModelLoader that touch this path
(python/sglang/srt/model_loader/loader.py:1211-1218): every module's
`quant_method.process_weights_after_loading` runs FIRST, then `initialize_dummy_weights`
(python/sglang/srt/model_loader/weight_utils.py:1108-1137) fills every FLOATING-POINT entry of
`state_dict()` in place with uniform(-1e-3, 1e-3) from a per-parameter generator seeded with the
same seed, so the values depend only on a parameter's element count and dtype (identical on all
TP ranks).  Integer parameters (qweight, qzeros) stay whatever `torch.empty` left in them in the
reference; here they are seeded explicitly, BEFORE the hook, so that runs are reproducible.
"""
from __future__ import annotations

import torch


def initialize_dummy_weights(model: torch.nn.Module, low: float = -1e-3, high: float = 1e-3, seed: int = 1234) -> None:
    """reference weight_utils.py:1108-1137 (the < 16-bit branch is FP8-only and not on this path)."""
    for param in model.state_dict().values():
        if torch.is_floating_point(param):
            generator = torch.Generator(device=param.data.device)
            generator.manual_seed(seed)
            param.uniform_(low, high, generator=generator)


def seed_integer_params(model: torch.nn.Module, seed: int = 1234) -> None:
    """Full-range int32 words for every integer parameter (uniform nibbles incl. the sign bit),
    from one generator per parameter seeded with `seed` + the parameter's position."""
    for i, (_, param) in enumerate(model.named_parameters()):
        if not torch.is_floating_point(param):
            g = torch.Generator(device=param.device)
            g.manual_seed(seed + i)
            param.data.copy_(torch.randint(-(2**31), 2**31, param.shape, generator=g, device=param.device,
                                           dtype=torch.int64).to(param.dtype))


def load_dummy_model(model: torch.nn.Module, seed: int = 1234) -> torch.nn.Module:
    """seed integers -> process_weights_after_loading on every module -> dummy float init;
    the order of loader.py:1211-1218."""
    seed_integer_params(model, seed)
    for _, module in model.named_modules():
        quant_method = getattr(module, "quant_method", None)
        if quant_method is not None:
            quant_method.process_weights_after_loading(module)
    initialize_dummy_weights(model, seed=seed)
    return model.eval()
