"""Loading a model in 4 / 8 bits through the optional ``bitsandbytes`` package.

bitsandbytes' int8 / nf4 kernels are a library route to small-memory inference; they are not part of this framework's B200 hot path
(whose low-precision route is MXFP8 on the tcgen05 tensor cores).  What a framework has to provide around them is the plumbing: decide
which modules stay in high precision (output head, tied weights, user exclusions), swap the remaining ``nn.Linear`` for the bnb classes on
the meta device, infer a device map with the quantised sizes, stream the checkpoint in and dispatch.  Everything is import-guarded.
Behavioural parity: reference ``src/accelerate/utils/bnb.py:44-469``.
"""

from __future__ import annotations

import logging
from copy import deepcopy

import torch
import torch.nn as nn

from .dataclasses import BnbQuantizationConfig
from .imports import is_4bit_bnb_available, is_8bit_bnb_available, is_bnb_available

logger = logging.getLogger(__name__)

_AUTO_MAPS = ("auto", "balanced", "balanced_low_0", "sequential")


def _require_bnb(cfg: BnbQuantizationConfig) -> None:
    if cfg.load_in_8bit and not is_8bit_bnb_available():
        raise ImportError("You have a version of `bitsandbytes` that is not compatible with 8bit quantization, make sure you have the "
                          "latest version of `bitsandbytes` installed.")
    if cfg.load_in_4bit and not is_4bit_bnb_available():
        raise ValueError("You have a version of `bitsandbytes` that is not compatible with 4bit quantization, make sure you have the "
                         "latest version of `bitsandbytes` installed.")


def _excluded(qualified_name: str, exclusions) -> bool:
    """``exclusions`` name modules either exactly or by a path component (``"lm_head"`` excludes ``model.lm_head`` and what is below it)."""
    return any(key == qualified_name or (key in qualified_name and f"{key}." in qualified_name) for key in exclusions)


def get_keys_to_not_convert(model) -> list:
    """Modules that should stay un-quantised when the user did not say: whatever takes part in weight tying, and the last top-level
    child (the output head, by convention).  A bare base model without ties has neither."""
    from .modeling import find_tied_parameters

    probe = deepcopy(model)  # tie_weights() mutates
    if hasattr(probe, "tie_weights"):
        probe.tie_weights()
    tied = {name for group in find_tied_parameters(probe) for name in group}
    prefix = getattr(model, "base_model_prefix", None)
    if not tied and prefix is not None and not hasattr(model, prefix):
        return []
    children = [name for name, _ in model.named_children()]
    keep = [name.replace(".weight", "").replace(".bias", "") for name in sorted(tied)] + children[-1:]
    return list(dict.fromkeys(keep))


def _bnb_linear(bnb, source: nn.Linear, cfg: BnbQuantizationConfig):
    has_bias = source.bias is not None
    if cfg.load_in_8bit:
        return bnb.nn.Linear8bitLt(source.in_features, source.out_features, has_bias, has_fp16_weights=False, threshold=cfg.llm_int8_threshold)
    if cfg.load_in_4bit:
        return bnb.nn.Linear4bit(source.in_features, source.out_features, has_bias, cfg.bnb_4bit_compute_dtype,
                                 compress_statistics=cfg.bnb_4bit_use_double_quant, quant_type=cfg.bnb_4bit_quant_type)
    raise ValueError("load_in_8bit and load_in_4bit can't be both False")


def replace_with_bnb_layers(model, bnb_quantization_config, modules_to_not_convert=None, current_key_name=None):
    """Swap every ``nn.Linear`` that is not excluded for its bitsandbytes counterpart, in place.  Returns ``(model, replaced_any)``.
    (``current_key_name`` is the path prefix when called on a sub-module.)"""
    import bitsandbytes as bnb

    exclusions = list(modules_to_not_convert or ())
    prefix = ".".join(current_key_name) + "." if current_key_name else ""
    targets = [(name, module) for name, module in model.named_modules()
               if isinstance(module, nn.Linear) and name and name.rsplit(".", 1)[-1] not in exclusions and not _excluded(prefix + name, exclusions)]
    for name, linear in targets:
        replacement = _bnb_linear(bnb, linear, bnb_quantization_config)
        replacement.weight.data = linear.weight.data
        if linear.bias is not None:
            replacement.bias.data = linear.bias.data
        replacement.requires_grad_(False)
        parent_path, _, leaf = name.rpartition(".")
        setattr(model.get_submodule(parent_path) if parent_path else model, leaf, replacement)
    return model, bool(targets)


def has_4bit_bnb_layers(model) -> bool:
    if not is_bnb_available():
        return False
    import bitsandbytes as bnb

    return any(isinstance(module, bnb.nn.Linear4bit) for module in model.modules())


def _quantize_loaded_model(model, cfg: BnbQuantizationConfig, exclusions, device: torch.device):
    """The discouraged path: the weights are already in memory.  Swap the layers, cast, and move to the GPU (quantisation happens there)."""
    logger.warning("It is not recommended to quantize a loaded model. The model should be instantiated under the `init_empty_weights` "
                   "context manager.")
    model, _ = replace_with_bnb_layers(model, cfg, modules_to_not_convert=exclusions)
    for name, tensor in model.state_dict().items():
        if any(key in name for key in cfg.keep_in_fp32_modules):
            tensor.data = tensor.data.to(torch.float32)
        elif torch.is_floating_point(tensor):
            tensor.data = tensor.data.to(cfg.torch_dtype)
    if not torch.cuda.is_available():
        raise RuntimeError("No GPU found. A GPU is needed for quantization.")
    model.to(torch.cuda.current_device())
    if device.type == "cuda":
        torch.cuda.empty_cache()
    logger.info(f"The model device type is {device.type}. However, gpu is needed for quantization. We move the model to gpu.")
    return model


def _auto_device_map(model, cfg: BnbQuantizationConfig, strategy: str, no_split_module_classes, max_memory):
    from .modeling import get_balanced_memory, infer_auto_device_map

    if strategy not in _AUTO_MAPS:
        raise ValueError("If passing a string for `device_map`, please choose 'auto', 'balanced', 'balanced_low_0' or 'sequential'.")
    # excluded modules keep the load dtype, keep_in_fp32 ones fp32; everything else counts with the quantised size
    special = {name: cfg.torch_dtype for name, _ in model.named_parameters() if any(key in name for key in cfg.skip_modules)}
    special.update({name: torch.float32 for name, _ in model.named_parameters() if any(key in name for key in cfg.keep_in_fp32_modules)})
    sizing = dict(special_dtypes=special, no_split_module_classes=no_split_module_classes, dtype=cfg.target_dtype)
    if strategy != "sequential":
        max_memory = get_balanced_memory(model, low_zero=strategy == "balanced_low_0", max_memory=max_memory, **sizing)
    return infer_auto_device_map(model, max_memory=max_memory, **sizing)


def get_quantized_model_device_map(model, bnb_quantization_config: BnbQuantizationConfig, device_map=None, max_memory=None,
                                   no_split_module_classes=None):
    """The device map a quantised ``model`` will be loaded with: the current GPU when none is asked for, an inferred one for
    ``"auto"`` / ``"balanced"`` / ``"balanced_low_0"`` / ``"sequential"`` (sized with the quantised bytes), or the given dict — checked:
    a 4-bit layer cannot live on cpu / disk (its entries must be in ``skip_modules``), an 8-bit one can and is quantised on the way."""
    cfg = bnb_quantization_config
    if device_map is None:
        if not torch.cuda.is_available():
            raise RuntimeError("No GPU found. A GPU is needed for quantization.")
        device_map = {"": torch.cuda.current_device()}
        logger.info("The device_map was not initialized.Setting device_map to `{'':torch.cuda.current_device()}`.")
    if isinstance(device_map, str):
        device_map = _auto_device_map(model, cfg, device_map, no_split_module_classes, max_memory)
    if isinstance(device_map, dict):
        left_whole = set((cfg.skip_modules or []) + (cfg.keep_in_fp32_modules or []))
        offloaded = [key for key, where in device_map.items() if key not in left_whole and where in ("cpu", "disk")]
        if offloaded and cfg.load_in_4bit:
            raise ValueError(
                "Some modules are dispatched on the CPU or the disk. Make sure you have enough GPU RAM to fit the quantized model. "
                "If you want to dispatch the model on the CPU or the disk while keeping these modules in `torch_dtype`, you need to "
                f"pass a custom `device_map` to `load_and_quantize_model` (offloaded here: {offloaded}).")
        if offloaded:
            logger.info("Some modules are are offloaded to the CPU or the disk. Note that these modules will be converted to 8-bit")
    return device_map


def get_parameter_device(parameter: torch.nn.Module):
    """Device of a module's first parameter."""
    return next(parameter.parameters()).device


def quantize_and_offload_8bit(model, param, param_name, new_dtype, offload_folder, offload_index, fp16_statistics):
    """Store an 8-bit weight that will live on cpu / disk: quantise it by passing it through GPU 0 once (unless the checkpoint
    already holds int8 data with its ``SCB`` statistics), write the int8 data and the statistics to ``offload_folder``, and leave a
    meta tensor in the model."""
    from .modeling import _owner_of, set_module_tensor_to_device
    from .offload import offload_weight

    statistics_name = param_name.replace("weight", "SCB")
    if fp16_statistics is None:
        set_module_tensor_to_device(model, param_name, 0, dtype=new_dtype, value=param)
        owner, leaf = _owner_of(model, param_name)
        quantised = owner._parameters[leaf]
        quantised.requires_grad = False
        offload_weight(quantised, param_name, offload_folder, index=offload_index)
        if hasattr(quantised, "SCB"):
            offload_weight(quantised.SCB, statistics_name, offload_folder, index=offload_index)
    else:
        offload_weight(param, param_name, offload_folder, index=offload_index)
        offload_weight(fp16_statistics, statistics_name, offload_folder, index=offload_index)
    set_module_tensor_to_device(model, param_name, "meta", dtype=new_dtype, value=torch.empty(*param.size()))


def load_and_quantize_model(model: torch.nn.Module, bnb_quantization_config: BnbQuantizationConfig, weights_location=None, device_map=None,
                            no_split_module_classes=None, max_memory=None, offload_folder=None, offload_state_dict: bool = False):
    """Quantise ``model`` with bitsandbytes.  The intended use: ``model`` was built under ``init_empty_weights`` and
    ``weights_location`` holds its checkpoint — layers are swapped on the meta device, a device map is inferred (or taken), the weights
    are streamed to their devices (quantised as they land on the GPU) and the model is dispatched."""
    cfg = bnb_quantization_config
    _require_bnb(cfg)
    from ..big_modeling import dispatch_model, init_empty_weights
    from .modeling import load_checkpoint_in_model

    if cfg.skip_modules is None:
        cfg.skip_modules = get_keys_to_not_convert(model)
    if cfg.load_in_4bit and isinstance(device_map, dict):
        # 4-bit layers cannot be offloaded: whatever the map sends to cpu / disk stays un-quantised
        cfg.skip_modules.extend(key for key, where in device_map.items() if where in ("cpu", "disk"))
    if cfg.keep_in_fp32_modules is None:
        cfg.keep_in_fp32_modules = []
    exclusions = cfg.skip_modules
    exclusions.extend(cfg.keep_in_fp32_modules)
    model.is_loaded_in_4bit, model.is_loaded_in_8bit = cfg.load_in_4bit, cfg.load_in_8bit

    first = next(model.parameters(), None)
    device = first.device if first is not None else torch.device("cpu")
    if device.type != "meta":
        return _quantize_loaded_model(model, cfg, exclusions, device)
    if weights_location is None:
        raise RuntimeError(f"`weights_location` needs to be the folder path containing the weights of the model, but we found {weights_location} ")
    with init_empty_weights():
        model, _ = replace_with_bnb_layers(model, cfg, modules_to_not_convert=exclusions)
    device_map = get_quantized_model_device_map(model, cfg, device_map, max_memory=max_memory, no_split_module_classes=no_split_module_classes)
    load_checkpoint_in_model(model, weights_location, device_map, dtype=cfg.torch_dtype, offload_folder=offload_folder,
                             offload_state_dict=offload_state_dict, keep_in_fp32_modules=cfg.keep_in_fp32_modules,
                             offload_8bit_bnb=cfg.load_in_8bit and offload_folder is not None)
    return dispatch_model(model, device_map=device_map, offload_dir=offload_folder)
