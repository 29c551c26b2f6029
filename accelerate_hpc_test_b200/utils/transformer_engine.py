"""TransformerEngine adapter (reference src/accelerate/utils/transformer_engine.py:26-186).

TE is an optional library back-end for fp8; the native MXFP8 path lives in :mod:`..parallel.precision`.
With TE absent every entry point raises an informative ImportError instead of failing on import.
"""

from types import MethodType

import torch.nn as nn

from .imports import is_transformer_engine_available


def _require_te():
    if not is_transformer_engine_available():
        raise ImportError("Using `convert_model` requires transformer_engine to be installed; the native fp8 path "
                          "(`ACCELERATE_FP8_BACKEND=NATIVE`) needs no extra package.")
    import transformer_engine.pytorch as te

    return te


def convert_model(model, to_transformer_engine=True, _convert_linear=True, _convert_ln=True):
    """Swap ``nn.Linear`` / ``nn.LayerNorm`` for their TE versions (dims must be multiples of 16), or back."""
    te = _require_te()
    import torch

    for name, module in model.named_children():
        if isinstance(module, nn.Linear) and to_transformer_engine and _convert_linear:
            has_bias = module.bias is not None
            if any(p % 16 != 0 for p in module.weight.shape):
                return
            te_module = te.Linear(module.in_features, module.out_features, bias=has_bias, params_dtype=module.weight.dtype)
            te_module.weight.copy_(module.weight) if False else te_module.weight.data.copy_(module.weight.data)
            if has_bias:
                te_module.bias.data.copy_(module.bias.data)
            setattr(model, name, te_module)
        elif isinstance(module, nn.LayerNorm) and to_transformer_engine and _convert_ln:
            te_module = te.LayerNorm(module.normalized_shape[0], eps=module.eps, params_dtype=module.weight.dtype)
            te_module.weight.data.copy_(module.weight.data)
            te_module.bias.data.copy_(module.bias.data)
            setattr(model, name, te_module)
        elif isinstance(module, te.Linear) and not to_transformer_engine and _convert_linear:
            has_bias = module.bias is not None
            new_module = nn.Linear(module.in_features, module.out_features, bias=has_bias, dtype=module.weight.dtype, device=module.weight.device)
            new_module.weight.data.copy_(module.weight.data)
            if has_bias:
                new_module.bias.data.copy_(module.bias.data)
            setattr(model, name, new_module)
        elif isinstance(module, te.LayerNorm) and not to_transformer_engine and _convert_ln:
            new_module = nn.LayerNorm(module.normalized_shape[0], eps=module.eps, dtype=module.weight.dtype, device=module.weight.device)
            new_module.weight.data.copy_(module.weight.data)
            new_module.bias.data.copy_(module.bias.data)
            setattr(model, name, new_module)
        else:
            convert_model(module, to_transformer_engine=to_transformer_engine, _convert_linear=_convert_linear, _convert_ln=_convert_ln)


def has_transformer_engine_layers(model) -> bool:
    if not is_transformer_engine_available():
        return False
    import transformer_engine.pytorch as te

    return any(isinstance(m, (te.LayerNorm, te.Linear, te.TransformerLayer)) for m in model.modules())


def contextual_fp8_autocast(model_forward, fp8_recipe, use_during_eval=False):
    """Wrap ``model_forward`` in ``te.fp8_autocast`` (disabled in eval mode unless asked)."""
    te = _require_te()

    def forward(self, *args, **kwargs):
        enabled = use_during_eval or self.training
        with te.fp8_autocast(enabled=enabled, fp8_recipe=fp8_recipe):
            return model_forward(*args, **kwargs)

    forward.__wrapped__ = model_forward
    return forward


def apply_fp8_autowrap(model, fp8_recipe_handler):
    """Build the TE recipe (DelayedScaling or MXFP8BlockScaling) and wrap ``model.forward`` with it."""
    te = _require_te()
    import transformer_engine.common.recipe as te_recipe

    kwargs = fp8_recipe_handler.to_kwargs() if fp8_recipe_handler is not None else {}
    if "fp8_format" in kwargs:
        kwargs["fp8_format"] = getattr(te_recipe.Format, kwargs["fp8_format"])
    use_during_eval = kwargs.pop("use_autocast_during_eval", False)
    use_mxfp8_block_scaling = kwargs.pop("use_mxfp8_block_scaling", False)
    if use_mxfp8_block_scaling:
        fp8_recipe = te_recipe.MXFP8BlockScaling(**{k: v for k, v in kwargs.items() if k in ("fp8_format", "margin")})
    else:
        fp8_recipe = te_recipe.DelayedScaling(**kwargs)
    new_forward = contextual_fp8_autocast(model.forward, fp8_recipe, use_during_eval)
    if hasattr(model.forward, "__func__"):
        model.forward = MethodType(new_forward, model)
    else:
        model.forward = new_forward
    return model
