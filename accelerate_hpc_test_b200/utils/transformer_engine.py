"""Adapter for the optional TransformerEngine fp8 back-end (``ACCELERATE_FP8_BACKEND=TE``).

TE is a library route to fp8; the native route is the MXFP8 tcgen05 GEMM in ``parallel/precision.py`` and needs nothing installed.  What
TE needs from a framework is small: swap ``nn.Linear`` / ``nn.LayerNorm`` for the TE modules that understand ``fp8_autocast`` (and back,
for export), and run the model's forward inside that context with the user's recipe.  Without TE installed every entry point raises an
ImportError that says so.  Behavioural parity: reference ``src/accelerate/utils/transformer_engine.py:26-186``.
"""

from __future__ import annotations

from types import MethodType

import torch.nn as nn

from .imports import is_transformer_engine_available

_FP8_TILE = 16


def _require_te():
    if not is_transformer_engine_available():
        raise ImportError("Using `convert_model` requires transformer_engine to be installed; the native fp8 path "
                          "(`ACCELERATE_FP8_BACKEND=NATIVE`) needs no extra package.")
    import transformer_engine.pytorch as te

    return te


def _copy_parameters(source: nn.Module, target: nn.Module) -> nn.Module:
    for name in ("weight", "bias"):
        value = getattr(source, name, None)
        if value is not None:
            getattr(target, name).data.copy_(value.data)
    return target


def _swap_rules(te, to_transformer_engine: bool, linear: bool, norm: bool) -> list:
    """``[(class to replace, builder of the replacement)]`` for one direction of the conversion."""
    rules = []
    if to_transformer_engine:
        if linear:
            rules.append((nn.Linear, lambda m: te.Linear(m.in_features, m.out_features, bias=m.bias is not None, params_dtype=m.weight.dtype)))
        if norm:
            rules.append((nn.LayerNorm, lambda m: te.LayerNorm(m.normalized_shape[0], eps=m.eps, params_dtype=m.weight.dtype)))
    else:
        if linear:
            rules.append((te.Linear, lambda m: nn.Linear(m.in_features, m.out_features, bias=m.bias is not None, dtype=m.weight.dtype,
                                                         device=m.weight.device)))
        if norm:
            rules.append((te.LayerNorm, lambda m: nn.LayerNorm(m.normalized_shape[0], eps=m.eps, dtype=m.weight.dtype, device=m.weight.device)))
    return rules


def convert_model(model, to_transformer_engine=True, _convert_linear=True, _convert_ln=True):
    """Replace, in place and recursively, torch linears / layer norms by their TE counterparts (or the reverse).  As in the reference, a
    linear whose weight has a dimension that is not a multiple of 16 stops the conversion of the container it is met in."""
    te = _require_te()
    rules = _swap_rules(te, to_transformer_engine, _convert_linear, _convert_ln)
    for name, child in model.named_children():
        builder = next((build for kind, build in rules if isinstance(child, kind)), None)
        if builder is None:
            convert_model(child, to_transformer_engine=to_transformer_engine, _convert_linear=_convert_linear, _convert_ln=_convert_ln)
            continue
        if to_transformer_engine and isinstance(child, nn.Linear) and any(dim % _FP8_TILE for dim in child.weight.shape):
            return
        setattr(model, name, _copy_parameters(child, builder(child)))


def has_transformer_engine_layers(model) -> bool:
    if not is_transformer_engine_available():
        return False
    import transformer_engine.pytorch as te

    return any(isinstance(module, (te.LayerNorm, te.Linear, te.TransformerLayer)) for module in model.modules())


def contextual_fp8_autocast(model_forward, fp8_recipe, use_during_eval=False):
    """A forward that runs ``model_forward`` under ``te.fp8_autocast`` — on while training, and in eval only when asked."""
    te = _require_te()

    def forward(self, *args, **kwargs):
        with te.fp8_autocast(enabled=use_during_eval or self.training, fp8_recipe=fp8_recipe):
            return model_forward(*args, **kwargs)

    forward.__wrapped__ = model_forward  # `unwrap_model` peels this off again
    return forward


def apply_fp8_autowrap(model, fp8_recipe_handler):
    """Turn the user's ``TERecipeKwargs`` into a TE recipe (delayed scaling, or MXFP8 block scaling on Blackwell) and make
    ``model.forward`` run under it."""
    _require_te()
    import transformer_engine.common.recipe as te_recipe

    options = fp8_recipe_handler.to_kwargs() if fp8_recipe_handler is not None else {}
    if "fp8_format" in options:
        options["fp8_format"] = getattr(te_recipe.Format, options["fp8_format"])
    use_during_eval = options.pop("use_autocast_during_eval", False)
    if options.pop("use_mxfp8_block_scaling", False):
        recipe = te_recipe.MXFP8BlockScaling(**{key: options[key] for key in ("fp8_format", "margin") if key in options})
    else:
        recipe = te_recipe.DelayedScaling(**options)
    wrapped = contextual_fp8_autocast(model.forward, recipe, use_during_eval)
    model.forward = MethodType(wrapped, model) if hasattr(model.forward, "__func__") else wrapped
    return model
