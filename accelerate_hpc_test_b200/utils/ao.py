"""torchao float8 adapter (reference src/accelerate/utils/ao.py:32-143); optional library back-end."""

from functools import partial
from typing import Callable, Optional

import torch

from .imports import is_torchao_available


def find_first_last_linear_layers(model: torch.nn.Module):
    """Names of the first and last ``nn.Linear`` in definition order."""
    first_linear, last_linear = None, None
    for name, module in model.named_modules():
        if isinstance(module, torch.nn.Linear):
            if first_linear is None:
                first_linear = name
            last_linear = name
    return first_linear, last_linear


def filter_linear_layers(module, fqn: str, layers_to_filter: list) -> bool:
    """True for linears that can run in float8: dims multiples of 16 and not in ``layers_to_filter``."""
    if isinstance(module, torch.nn.Linear):
        if module.in_features % 16 != 0 or module.out_features % 16 != 0:
            return False
    if fqn in layers_to_filter:
        return False
    return True


def filter_first_and_last_linear_layers(module, fqn: str) -> bool:
    raise NotImplementedError("use `functools.partial(filter_linear_layers, layers_to_filter=[first, last])`")


def has_ao_layers(model: torch.nn.Module) -> bool:
    if not is_torchao_available():
        return False
    from torchao.float8.float8_linear import Float8Linear

    return any(isinstance(m, Float8Linear) for m in model.modules())


def convert_model_to_fp8_ao(model: torch.nn.Module, config=None, module_filter_func: Optional[Callable] = filter_first_and_last_linear_layers):
    """``torchao.float8.convert_to_float8_training`` skipping the first/last linear by default."""
    if not is_torchao_available():
        raise ImportError("torchao is not installed; the native MXFP8 path (`ACCELERATE_FP8_BACKEND=NATIVE`) needs no extra package.")
    from torchao.float8 import convert_to_float8_training

    first_linear, last_linear = find_first_last_linear_layers(model)
    if module_filter_func is None or module_filter_func is filter_first_and_last_linear_layers:
        module_filter_func = partial(filter_linear_layers, layers_to_filter=[first_linear, last_linear])
    convert_to_float8_training(model, module_filter_fn=module_filter_func, config=config)
