"""Adapter for the optional ``torchao`` float8 training back-end (``ACCELERATE_FP8_BACKEND=AO``).

The native fp8 path of this framework is the MXFP8 block-scaled tcgen05 GEMM in ``parallel/precision.py``; this module is for users who
ask for torchao explicitly.  torchao swaps ``nn.Linear`` for ``Float8Linear``; by convention the first and the last linear of a network
(input embedding projection / output head) stay in high precision, and so does any layer whose dimensions the fp8 GEMMs cannot tile.
Behavioural parity: reference ``src/accelerate/utils/ao.py:32-143``.
"""

from __future__ import annotations

from typing import Callable, Optional

import torch

from .imports import is_torchao_available

_FP8_TILE = 16  # fp8 tensor-core GEMMs want every matrix dimension to be a multiple of 16


def find_first_last_linear_layers(model: torch.nn.Module):
    """``(first, last)`` qualified names of the ``nn.Linear`` layers in registration order (both None when there is none)."""
    linears = [name for name, module in model.named_modules() if isinstance(module, torch.nn.Linear)]
    return (linears[0], linears[-1]) if linears else (None, None)


def filter_linear_layers(module, fqn: str, layers_to_filter: list) -> bool:
    """Should ``module`` (named ``fqn``) be converted?  No for the excluded names and for linears with a dimension off the fp8 tile."""
    if fqn in layers_to_filter:
        return False
    if isinstance(module, torch.nn.Linear):
        return module.in_features % _FP8_TILE == 0 and module.out_features % _FP8_TILE == 0
    return True


def filter_first_and_last_linear_layers(module, fqn: str) -> bool:
    """Marker for "the default filter": the first and last linear are only known once a model is given, so ``convert_model_to_fp8_ao``
    resolves this name to ``filter_linear_layers`` bound to that model's first and last linear."""
    raise NotImplementedError("resolved by `convert_model_to_fp8_ao`; call `filter_linear_layers(module, fqn, [first, last])` directly")


def has_ao_layers(model: torch.nn.Module) -> bool:
    if not is_torchao_available():
        return False
    from torchao.float8.float8_linear import Float8Linear

    return any(isinstance(module, Float8Linear) for module in model.modules())


def convert_model_to_fp8_ao(model: torch.nn.Module, config=None, module_filter_func: Optional[Callable] = filter_first_and_last_linear_layers):
    """In-place ``torchao.float8.convert_to_float8_training``.  ``module_filter_func(module, fqn) -> bool`` picks the layers; the
    default keeps the first and the last linear (and badly shaped ones) in high precision."""
    if not is_torchao_available():
        raise ImportError("torchao is not installed; the native MXFP8 path (`ACCELERATE_FP8_BACKEND=NATIVE`) needs no extra package.")
    from torchao.float8 import convert_to_float8_training

    if module_filter_func in (None, filter_first_and_last_linear_layers):
        keep_precise = list(find_first_last_linear_layers(model))

        def module_filter_func(module, fqn):
            return filter_linear_layers(module, fqn, keep_precise)

    convert_to_float8_training(model, module_filter_fn=module_filter_func, config=config)
