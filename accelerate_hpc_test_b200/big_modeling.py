"""Big-model inference: meta-device initialisation, device-map dispatch, CPU / disk offload.

API parity with reference src/accelerate/big_modeling.py: ``init_empty_weights`` :60-94,
``init_on_device`` :97-171, ``cpu_offload`` :174-221, ``cpu_offload_with_hook`` :224-262,
``disk_offload`` :265-307, ``dispatch_model`` :310-510, ``load_checkpoint_and_dispatch`` :513-651,
``attach_layerwise_casting_hooks`` :654-750, ``_attach_context_parallel_hooks`` :753-790.
"""

from __future__ import annotations

import logging
import os
import re
from contextlib import contextmanager
from functools import wraps
from typing import Optional, Union

import torch
import torch.nn as nn

from .hooks import (
    AlignDevicesHook,
    CpuOffload,
    LayerwiseCastingHook,
    UserCpuOffloadHook,
    add_hook_to_module,
    attach_align_device_hook,
    attach_align_device_hook_on_blocks,
)
from .utils.modeling import (
    check_device_map,
    find_tied_parameters,
    get_balanced_memory,
    infer_auto_device_map,
    load_checkpoint_in_model,
    retie_parameters,
)
from .utils.offload import OffloadedWeightsLoader, extract_submodules_state_dict, offload_state_dict
from .utils.environment import parse_flag_from_env

logger = logging.getLogger(__name__)


@contextmanager
def init_empty_weights(include_buffers: bool = None):
    """Inside the block modules are created on the ``meta`` device (no memory, no init cost)."""
    if include_buffers is None:
        include_buffers = parse_flag_from_env("ACCELERATE_INIT_INCLUDE_BUFFERS", False)
    with init_on_device(torch.device("meta"), include_buffers=include_buffers) as f:
        yield f


@contextmanager
def init_on_device(device: torch.device, include_buffers: bool = None):
    """Inside the block every registered parameter (and optionally buffer / factory tensor) lands on ``device``."""
    if include_buffers is None:
        include_buffers = parse_flag_from_env("ACCELERATE_INIT_INCLUDE_BUFFERS", False)
    if include_buffers:
        with device:
            yield
        return
    old_register_parameter = nn.Module.register_parameter
    if include_buffers:
        old_register_buffer = nn.Module.register_buffer

    def register_empty_parameter(module, name, param):
        old_register_parameter(module, name, param)
        if param is not None:
            param_cls = type(module._parameters[name])
            kwargs = module._parameters[name].__dict__
            kwargs["requires_grad"] = param.requires_grad
            module._parameters[name] = param_cls(module._parameters[name].to(device), **kwargs)

    def register_empty_buffer(module, name, buffer, persistent=True):
        old_register_buffer(module, name, buffer, persistent=persistent)
        if buffer is not None:
            module._buffers[name] = module._buffers[name].to(device)

    if include_buffers:
        tensor_constructors_to_patch = {n: getattr(torch, n) for n in ["empty", "zeros", "ones", "full"]}
    else:
        tensor_constructors_to_patch = {}

    def patch_tensor_constructor(fn):
        def wrapper(*args, **kwargs):
            kwargs["device"] = device
            return fn(*args, **kwargs)

        return wrapper

    try:
        nn.Module.register_parameter = register_empty_parameter
        if include_buffers:
            nn.Module.register_buffer = register_empty_buffer
        for torch_function_name in tensor_constructors_to_patch.keys():
            setattr(torch, torch_function_name, patch_tensor_constructor(getattr(torch, torch_function_name)))
        yield
    finally:
        nn.Module.register_parameter = old_register_parameter
        if include_buffers:
            nn.Module.register_buffer = old_register_buffer
        for torch_function_name, old_torch_function in tensor_constructors_to_patch.items():
            setattr(torch, torch_function_name, old_torch_function)


def cpu_offload(model: nn.Module, execution_device: Optional[torch.device] = None, offload_buffers: bool = False,
                state_dict: Optional[dict] = None, preload_module_classes: Optional[list] = None):
    """Keep all weights on CPU; each sub-module's weights visit ``execution_device`` only during its forward."""
    if execution_device is None:
        execution_device = next(iter(model.parameters())).device
    if state_dict is None:
        state_dict = {n: p.to("cpu") for n, p in model.state_dict().items()}
    add_hook_to_module(model, AlignDevicesHook(io_same_device=True), append=True)
    attach_align_device_hook(model, execution_device=execution_device, offload=True, offload_buffers=offload_buffers,
                             weights_map=state_dict, preload_module_classes=preload_module_classes)
    return model


def cpu_offload_with_hook(model: torch.nn.Module, execution_device: Optional[Union[int, str, torch.device]] = None,
                          prev_module_hook: Optional[UserCpuOffloadHook] = None):
    """Whole-model offload for pipelines of models: the model moves to the device when called and stays there until
    the returned hook's ``offload()`` (or the next model in the chain) sends it back."""
    hook = CpuOffload(execution_device=execution_device, prev_module_hook=prev_module_hook)
    add_hook_to_module(model, hook, append=True)
    user_hook = UserCpuOffloadHook(model, hook)
    return model, user_hook


def disk_offload(model: nn.Module, offload_dir: Union[str, os.PathLike], execution_device: Optional[torch.device] = None,
                 offload_buffers: bool = False, preload_module_classes: Optional[list] = None):
    """Like :func:`cpu_offload` with the weights memory-mapped from ``offload_dir``."""
    if not os.path.isdir(offload_dir) or not os.path.isfile(os.path.join(offload_dir, "index.json")):
        offload_state_dict(offload_dir, model.state_dict())
    if execution_device is None:
        execution_device = next(iter(model.parameters())).device
    weights_map = OffloadedWeightsLoader(save_folder=offload_dir)
    add_hook_to_module(model, AlignDevicesHook(io_same_device=True), append=True)
    attach_align_device_hook(model, execution_device=execution_device, offload=True, offload_buffers=offload_buffers,
                             weights_map=weights_map, preload_module_classes=preload_module_classes)
    return model


def dispatch_model(model: nn.Module, device_map: dict, main_device: Optional[torch.device] = None, state_dict: Optional[dict] = None,
                   offload_dir: Optional[Union[str, os.PathLike]] = None, offload_index: Optional[dict] = None,
                   offload_buffers: bool = False, skip_keys: Optional[Union[str, list]] = None,
                   preload_module_classes: Optional[list] = None, force_hooks: bool = False):
    """Place ``model`` according to ``device_map`` (module name -> gpu index | "cpu" | "disk") and hook it so a
    plain ``model(x)`` works: inputs follow the weights, offloaded weights are paged in per module."""
    check_device_map(model, device_map)
    is_bnb_quantized = (getattr(model, "is_quantized", False) or getattr(model, "is_loaded_in_8bit", False)) and getattr(model, "quantization_method", "bitsandbytes") == "bitsandbytes"
    if len(set(device_map.values())) > 1 or is_bnb_quantized or force_hooks:
        if main_device is None:
            if set(device_map.values()) == {"cpu"} or set(device_map.values()) == {"cpu", "disk"}:
                main_device = "cpu"
            else:
                main_device = [d for d in device_map.values() if d not in ["cpu", "disk"]][0]
        if main_device != "cpu":
            cpu_modules = [name for name, device in device_map.items() if device == "cpu"]
            if state_dict is None and len(cpu_modules) > 0:
                state_dict = extract_submodules_state_dict(model.state_dict(), cpu_modules)
        disk_modules = [name for name, device in device_map.items() if device == "disk"]
        if offload_dir is None and offload_index is None and len(disk_modules) > 0:
            raise ValueError(f"We need an `offload_dir` to dispatch this model according to this `device_map`, the following submodules need to be offloaded: {', '.join(disk_modules)}.")
        if len(disk_modules) > 0 and offload_index is None and (not os.path.isdir(offload_dir) or not os.path.isfile(os.path.join(offload_dir, "index.json"))):
            disk_state_dict = extract_submodules_state_dict(model.state_dict(), disk_modules)
            offload_state_dict(offload_dir, disk_state_dict)
        execution_device = {name: main_device if device in ["cpu", "disk"] else device for name, device in device_map.items()}
        execution_device[""] = main_device
        offloaded_devices = ["disk"] if main_device == "cpu" or main_device == "mps" else ["cpu", "disk"]
        offload = {name: device in offloaded_devices for name, device in device_map.items()}
        save_folder = offload_dir if len(disk_modules) > 0 else None
        if state_dict is not None or save_folder is not None or offload_index is not None:
            device = main_device if offload_index is not None else None
            weights_map = OffloadedWeightsLoader(state_dict=state_dict, save_folder=save_folder, index=offload_index, device=device)
        else:
            weights_map = None
        # tied parameters: materialise once per execution device
        tied_params = find_tied_parameters(model)
        tied_params_map = {}
        for group in tied_params:
            for param_name in group:
                data_ptr = recursive_getattr_safe(model, param_name).data_ptr()
                tied_params_map[data_ptr] = {}
        attach_align_device_hook_on_blocks(model, execution_device=execution_device, offload=offload, offload_buffers=offload_buffers,
                                           weights_map=weights_map, skip_keys=skip_keys, preload_module_classes=preload_module_classes,
                                           tied_params_map=tied_params_map)
        offloaded_devices_str = " and ".join([device for device in set(device_map.values()) if device in ("cpu", "disk")])
        if len(offloaded_devices_str) > 0:
            logger.warning(f"Some parameters are on the meta device because they were offloaded to the {offloaded_devices_str}.")
        retie_parameters(model, tied_params)

        def add_warning(fn, model):
            @wraps(fn)
            def wrapper(*args, **kwargs):
                warning_msg = "You shouldn't move a model that is dispatched using accelerate hooks."
                if str(fn.__name__) == "to":
                    to_device = torch._C._nn._parse_to(*args, **kwargs)[0]
                    if to_device is not None:
                        logger.warning(warning_msg)
                else:
                    logger.warning(warning_msg)
                for param in model.parameters():
                    if param.device == torch.device("meta"):
                        raise RuntimeError("You can't move a model that has some modules offloaded to cpu or disk.")
                return fn(*args, **kwargs)

            return wrapper

        model.to = add_warning(model.to, model)
        model.cuda = add_warning(model.cuda, model)
        use_multi_gpu = len([device for device in set(device_map.values()) if device not in ("cpu", "disk")]) > 1
        if use_multi_gpu and not model._hf_hook.__dict__.get("io_same_device", True) is False:
            pass
    else:
        device = list(device_map.values())[0]
        if isinstance(device, int):
            device = f"cuda:{device}"
        if device != "disk":
            model.to(device)
        else:
            raise ValueError("You are trying to offload the whole model to the disk. Please use the `disk_offload` function instead.")
    model.hf_device_map = dict(device_map)
    return model


def recursive_getattr_safe(obj, attr: str):
    for a in attr.split("."):
        obj = getattr(obj, a)
    return obj


def load_checkpoint_and_dispatch(model: nn.Module, checkpoint: Union[str, os.PathLike], device_map=None, max_memory: Optional[dict] = None,
                                 no_split_module_classes: Optional[list] = None, offload_folder=None, offload_buffers: bool = False,
                                 dtype=None, offload_state_dict: Optional[bool] = None, skip_keys=None, preload_module_classes=None,
                                 force_hooks: bool = False, strict: bool = False, full_state_dict: bool = True,
                                 broadcast_from_rank0: bool = False):
    """``infer_auto_device_map`` (for ``"auto" | "balanced" | "balanced_low_0" | "sequential"``) +
    ``load_checkpoint_in_model`` + ``dispatch_model``."""
    if isinstance(device_map, str) and device_map not in ["auto", "balanced", "balanced_low_0", "sequential"]:
        raise ValueError("If passing a string for `device_map`, please choose 'auto', 'balanced', 'balanced_low_0' or 'sequential'.")
    if isinstance(device_map, str):
        if device_map != "sequential":
            max_memory = get_balanced_memory(model, max_memory=max_memory, no_split_module_classes=no_split_module_classes, dtype=dtype,
                                             low_zero=(device_map == "balanced_low_0"))
        device_map = infer_auto_device_map(model, max_memory=max_memory, no_split_module_classes=no_split_module_classes, dtype=dtype,
                                           offload_buffers=offload_buffers)
    if offload_state_dict is None and device_map is not None and "disk" in device_map.values():
        offload_state_dict = True
    load_checkpoint_in_model(model, checkpoint, device_map=device_map, offload_folder=offload_folder, dtype=dtype,
                             offload_state_dict=bool(offload_state_dict), offload_buffers=offload_buffers, strict=strict,
                             full_state_dict=full_state_dict, broadcast_from_rank0=broadcast_from_rank0)
    if device_map is None:
        return model
    return dispatch_model(model, device_map=device_map, offload_dir=offload_folder, offload_buffers=offload_buffers, skip_keys=skip_keys,
                          preload_module_classes=preload_module_classes, force_hooks=force_hooks)


_SUPPORTED_CAST_LAYERS = (nn.Conv1d, nn.Conv2d, nn.Conv3d, nn.ConvTranspose1d, nn.ConvTranspose2d, nn.ConvTranspose3d, nn.Linear)
_DEFAULT_SKIP_CAST_PATTERNS = ("pos_embed", "patch_embed", "norm", "^proj_in$", "^proj_out$")


def attach_layerwise_casting_hooks(module: torch.nn.Module, storage_dtype: torch.dtype, compute_dtype: torch.dtype,
                                   skip_modules_pattern: Union[str, tuple] = None, skip_modules_classes: Optional[tuple] = None,
                                   non_blocking: bool = False) -> None:
    """Store the weights of conv / linear layers in ``storage_dtype`` and cast them up per layer for compute."""
    _attach_layerwise_casting_hooks(module, storage_dtype, compute_dtype, skip_modules_pattern, skip_modules_classes, non_blocking)


def _attach_layerwise_casting_hooks(module, storage_dtype, compute_dtype, skip_modules_pattern=None, skip_modules_classes=None,
                                    non_blocking=False, _prefix=""):
    should_skip = (skip_modules_classes is not None and isinstance(module, skip_modules_classes)) or (
        skip_modules_pattern is not None and any(re.search(pattern, _prefix) for pattern in skip_modules_pattern))
    if should_skip:
        logger.debug(f'Skipping layerwise casting for layer "{_prefix}"')
        return
    if isinstance(module, _SUPPORTED_CAST_LAYERS) or type(module).__name__ == "B200Linear":
        logger.debug(f'Applying layerwise casting to layer "{_prefix}"')
        add_hook_to_module(module, LayerwiseCastingHook(storage_dtype, compute_dtype, non_blocking), append=True)
        return
    for name, submodule in module.named_children():
        layer_name = f"{_prefix}.{name}" if _prefix else name
        _attach_layerwise_casting_hooks(submodule, storage_dtype, compute_dtype, skip_modules_pattern, skip_modules_classes,
                                        non_blocking, _prefix=layer_name)


def _attach_context_parallel_hooks(model: nn.Module):
    """Context parallelism supports causal attention only: drop ``attention_mask`` and force ``is_causal=True``
    on every ``*.self_attn`` module (reference big_modeling.py:753-790)."""

    def _self_attn_pre_forward_hook(_module, module_args, module_kwargs):
        if "attention_mask" in module_kwargs:
            module_kwargs["attention_mask"] = None
            module_kwargs["is_causal"] = True
        return module_args, module_kwargs

    for name, module in model.named_modules():
        if name.endswith("self_attn"):
            module.register_forward_pre_hook(_self_attn_pre_forward_hook, with_kwargs=True, prepend=True)
