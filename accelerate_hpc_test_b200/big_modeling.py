"""Big-model inference: build a model without allocating its weights, then run it from wherever the weights fit.

* ``init_empty_weights`` / ``init_on_device`` — module construction with parameters redirected to ``meta`` (or any device).
* ``cpu_offload`` / ``disk_offload`` — every leaf module streams its weights in for its own forward.
* ``cpu_offload_with_hook`` — whole-model ping-pong for pipelines of models.
* ``dispatch_model`` — realise a device map (module path → GPU index | ``"cpu"`` | ``"disk"``) with hooks so ``model(x)`` just works.
* ``load_checkpoint_and_dispatch`` — device-map inference + checkpoint streaming + dispatch in one call.
* ``attach_layerwise_casting_hooks`` — low-precision weight storage with per-layer up-casting.

Same public API as the reference's ``big_modeling.py``; the device-map analysis lives in
``_Placement`` and the construction-time redirection in ``_ConstructionRedirect``.
"""

from __future__ import annotations

import functools
import logging
import os
import re
from contextlib import contextmanager
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
from .utils.environment import parse_flag_from_env
from .utils.modeling import (
    check_device_map,
    find_tied_parameters,
    get_balanced_memory,
    infer_auto_device_map,
    load_checkpoint_in_model,
    retie_parameters,
)
from .utils.offload import INDEX_FILE, OffloadedWeightsLoader, extract_submodules_state_dict, offload_state_dict
from .utils.other import recursive_getattr

logger = logging.getLogger(__name__)
_OFF_DEVICE = ("cpu", "disk")
_AUTO_MAPS = ("auto", "balanced", "balanced_low_0", "sequential")


# =========================================================================== construction on another device
class _ConstructionRedirect:
    """While active, ``nn.Module.register_parameter`` re-homes every new parameter on ``device`` (keeping its class, its
    ``requires_grad`` and any extra attributes a Parameter subclass carries).  With ``buffers=True`` new buffers and the
    ``torch.empty/zeros/ones/full`` factories follow too."""

    _FACTORIES = ("empty", "zeros", "ones", "full")

    def __init__(self, device: torch.device, buffers: bool):
        self.device = device
        self.buffers = buffers
        self._saved = {}

    def _param_hook(self, original):
        device = self.device

        def register_parameter(module, name, param):
            original(module, name, param)
            if param is None:
                return
            held = module._parameters[name]
            extras = dict(held.__dict__)
            extras["requires_grad"] = param.requires_grad
            module._parameters[name] = type(held)(held.to(device), **extras)

        return register_parameter

    def _buffer_hook(self, original):
        device = self.device

        def register_buffer(module, name, buffer, persistent=True):
            original(module, name, buffer, persistent=persistent)
            if buffer is not None:
                module._buffers[name] = module._buffers[name].to(device)

        return register_buffer

    def _factory_hook(self, original):
        device = self.device

        @functools.wraps(original)
        def make(*args, **kwargs):
            kwargs["device"] = device
            return original(*args, **kwargs)

        return make

    def __enter__(self):
        self._saved["register_parameter"] = nn.Module.register_parameter
        nn.Module.register_parameter = self._param_hook(nn.Module.register_parameter)
        if self.buffers:
            self._saved["register_buffer"] = nn.Module.register_buffer
            nn.Module.register_buffer = self._buffer_hook(nn.Module.register_buffer)
            for fname in self._FACTORIES:
                self._saved[fname] = getattr(torch, fname)
                setattr(torch, fname, self._factory_hook(getattr(torch, fname)))
        return self

    def __exit__(self, *exc):
        nn.Module.register_parameter = self._saved.pop("register_parameter")
        if "register_buffer" in self._saved:
            nn.Module.register_buffer = self._saved.pop("register_buffer")
        for fname, fn in self._saved.items():
            setattr(torch, fname, fn)
        self._saved = {}
        return False


def _buffers_too(flag: Optional[bool]) -> bool:
    return parse_flag_from_env("ACCELERATE_INIT_INCLUDE_BUFFERS", False) if flag is None else bool(flag)


@contextmanager
def init_on_device(device: torch.device, include_buffers: bool = None):
    """Modules created inside the block get their parameters (and with ``include_buffers`` their buffers) on ``device``."""
    if _buffers_too(include_buffers):
        # torch's own device context already covers parameters, buffers and factory calls
        with device:
            yield
        return
    with _ConstructionRedirect(device, buffers=False):
        yield


@contextmanager
def init_empty_weights(include_buffers: bool = None):
    """Modules created inside the block cost no memory and no initialisation time: their parameters live on ``meta``."""
    with init_on_device(torch.device("meta"), include_buffers=_buffers_too(include_buffers)) as ctx:
        yield ctx


# =========================================================================== per-module streaming offload
def _first_param_device(model: nn.Module) -> torch.device:
    return next(iter(model.parameters())).device


def _stream_from(model: nn.Module, weights, execution_device, offload_buffers, preload_module_classes) -> nn.Module:
    if execution_device is None:
        execution_device = _first_param_device(model)
    # the root hook only sends the outputs back to where the inputs came from
    add_hook_to_module(model, AlignDevicesHook(io_same_device=True), append=True)
    attach_align_device_hook(model, execution_device=execution_device, offload=True, offload_buffers=offload_buffers,
                             weights_map=weights, preload_module_classes=preload_module_classes)
    return model


def cpu_offload(model: nn.Module, execution_device: Optional[torch.device] = None, offload_buffers: bool = False,
                state_dict: Optional[dict] = None, preload_module_classes: Optional[list] = None):
    """All weights rest in CPU memory; each tensor-owning sub-module borrows its weights for the duration of its forward."""
    target = _first_param_device(model) if execution_device is None else execution_device
    if state_dict is None:
        state_dict = {name: tensor.to("cpu") for name, tensor in model.state_dict().items()}
    return _stream_from(model, state_dict, target, offload_buffers, preload_module_classes)


def disk_offload(model: nn.Module, offload_dir: Union[str, os.PathLike], execution_device: Optional[torch.device] = None,
                 offload_buffers: bool = False, preload_module_classes: Optional[list] = None):
    """Same as :func:`cpu_offload` with the weights memory-mapped from ``offload_dir`` (written first if it has no index)."""
    target = _first_param_device(model) if execution_device is None else execution_device
    if not os.path.isfile(os.path.join(offload_dir, INDEX_FILE)):
        offload_state_dict(offload_dir, model.state_dict())
    return _stream_from(model, OffloadedWeightsLoader(save_folder=offload_dir), target, offload_buffers, preload_module_classes)


def cpu_offload_with_hook(model: nn.Module, execution_device: Optional[Union[int, str, torch.device]] = None,
                          prev_module_hook: Optional[UserCpuOffloadHook] = None):
    """Returns ``(model, handle)``.  The model jumps to the device when called and stays until ``handle.offload()`` — or until
    the next model of the chain (built with ``prev_module_hook=handle``) is called."""
    resident = CpuOffload(execution_device=execution_device, prev_module_hook=prev_module_hook)
    add_hook_to_module(model, resident, append=True)
    return model, UserCpuOffloadHook(model, resident)


# =========================================================================== device-map dispatch
class _Placement:
    """Everything ``dispatch_model`` derives from a device map."""

    def __init__(self, device_map: dict, main_device=None):
        self.device_map = device_map
        values = list(device_map.values())
        self.kinds = set(values)
        self.on_cpu = [name for name, d in device_map.items() if d == "cpu"]
        self.on_disk = [name for name, d in device_map.items() if d == "disk"]
        accelerators = [d for d in values if d not in _OFF_DEVICE]
        if main_device is None:
            main_device = accelerators[0] if accelerators else "cpu"
        self.main_device = main_device
        self.n_accelerators = len(set(accelerators))

    @property
    def needs_hooks(self) -> bool:
        return len(self.kinds) > 1

    def execution_devices(self) -> dict:
        table = {name: (self.main_device if d in _OFF_DEVICE else d) for name, d in self.device_map.items()}
        table[""] = self.main_device
        return table

    def streamed(self) -> dict:
        """Which blocks stream their weights: disk always; CPU too unless the CPU *is* where things run."""
        away = ("disk",) if self.main_device in ("cpu", "mps") else _OFF_DEVICE
        return {name: d in away for name, d in self.device_map.items()}

    def describe_offload(self) -> str:
        return " and ".join(sorted(k for k in self.kinds if k in _OFF_DEVICE))


def _guard_against_moves(model: nn.Module) -> None:
    """``model.to(...)`` / ``model.cuda()`` on a dispatched model warn, and refuse when weights are parked on ``meta``."""
    note = "You shouldn't move a model that is dispatched using accelerate hooks."

    def guarded(method):
        @functools.wraps(method)
        def call(*args, **kwargs):
            changes_device = True
            if method.__name__ == "to":
                changes_device = torch._C._nn._parse_to(*args, **kwargs)[0] is not None
            if changes_device:
                logger.warning(note)
            if any(p.device.type == "meta" for p in model.parameters()):
                raise RuntimeError("You can't move a model that has some modules offloaded to cpu or disk.")
            return method(*args, **kwargs)

        return call

    model.to = guarded(model.to)
    model.cuda = guarded(model.cuda)


def _is_bnb_quantized(model) -> bool:
    flagged = getattr(model, "is_quantized", False) or getattr(model, "is_loaded_in_8bit", False)
    return bool(flagged) and getattr(model, "quantization_method", "bitsandbytes") == "bitsandbytes"


def dispatch_model(model: nn.Module, device_map: dict, main_device: Optional[torch.device] = None, state_dict: Optional[dict] = None,
                   offload_dir: Optional[Union[str, os.PathLike]] = None, offload_index: Optional[dict] = None,
                   offload_buffers: bool = False, skip_keys: Optional[Union[str, list]] = None,
                   preload_module_classes: Optional[list] = None, force_hooks: bool = False):
    """Put every block of ``model`` where ``device_map`` says and hook the model so a plain call works: inputs follow the
    weights from block to block, CPU / disk blocks page their weights in, the final output returns to the caller's device."""
    check_device_map(model, device_map)
    plan = _Placement(device_map, main_device)
    if not (plan.needs_hooks or _is_bnb_quantized(model) or force_hooks):
        (only,) = plan.kinds
        if only == "disk":
            raise ValueError("You are trying to offload the whole model to the disk. Please use the `disk_offload` function instead.")
        model.to(f"cuda:{only}" if isinstance(only, int) else only)
        model.hf_device_map = dict(device_map)
        return model

    # ---- where do the parked weights come from
    if plan.main_device != "cpu" and state_dict is None and plan.on_cpu:
        state_dict = extract_submodules_state_dict(model.state_dict(), plan.on_cpu)
    folder = None
    if plan.on_disk:
        if offload_dir is None and offload_index is None:
            raise ValueError("We need an `offload_dir` to dispatch this model according to this `device_map`, the following submodules "
                             f"need to be offloaded: {', '.join(plan.on_disk)}.")
        if offload_index is None and not os.path.isfile(os.path.join(offload_dir, INDEX_FILE)):
            offload_state_dict(offload_dir, extract_submodules_state_dict(model.state_dict(), plan.on_disk))
        folder = offload_dir
    weights = None
    if state_dict is not None or folder is not None or offload_index is not None:
        weights = OffloadedWeightsLoader(state_dict=state_dict, save_folder=folder, index=offload_index,
                                         device=plan.main_device if offload_index is not None else None)

    # ---- tied weights are materialised once per device and shared between their modules
    tied_groups = find_tied_parameters(model)
    shared = {recursive_getattr(model, name).data_ptr(): {} for group in tied_groups for name in group}

    attach_align_device_hook_on_blocks(model, execution_device=plan.execution_devices(), offload=plan.streamed(),
                                       offload_buffers=offload_buffers, weights_map=weights, skip_keys=skip_keys,
                                       preload_module_classes=preload_module_classes, tied_params_map=shared)
    parked = plan.describe_offload()
    if parked:
        logger.warning(f"Some parameters are on the meta device because they were offloaded to the {parked}.")
    retie_parameters(model, tied_groups)
    _guard_against_moves(model)
    model.hf_device_map = dict(device_map)
    return model


def load_checkpoint_and_dispatch(model: nn.Module, checkpoint: Union[str, os.PathLike], device_map=None, max_memory: Optional[dict] = None,
                                 no_split_module_classes: Optional[list] = None, offload_folder=None, offload_buffers: bool = False,
                                 dtype=None, offload_state_dict: Optional[bool] = None, skip_keys=None, preload_module_classes=None,
                                 force_hooks: bool = False, strict: bool = False, full_state_dict: bool = True,
                                 broadcast_from_rank0: bool = False):
    """Load ``checkpoint`` into a (typically meta-initialised) model directly at its final placement, then dispatch it.
    ``device_map`` may be a mapping or one of ``"auto" | "balanced" | "balanced_low_0" | "sequential"``."""
    if isinstance(device_map, str):
        if device_map not in _AUTO_MAPS:
            raise ValueError("If passing a string for `device_map`, please choose 'auto', 'balanced', 'balanced_low_0' or 'sequential'.")
        budget = max_memory
        if device_map != "sequential":
            budget = get_balanced_memory(model, max_memory=max_memory, no_split_module_classes=no_split_module_classes, dtype=dtype,
                                         low_zero=device_map == "balanced_low_0")
        device_map = infer_auto_device_map(model, max_memory=budget, no_split_module_classes=no_split_module_classes, dtype=dtype,
                                           offload_buffers=offload_buffers)
    if offload_state_dict is None:
        offload_state_dict = device_map is not None and "disk" in device_map.values()
    load_checkpoint_in_model(model, checkpoint, device_map=device_map, offload_folder=offload_folder, dtype=dtype,
                             offload_state_dict=bool(offload_state_dict), offload_buffers=offload_buffers, strict=strict,
                             full_state_dict=full_state_dict, broadcast_from_rank0=broadcast_from_rank0)
    if device_map is None:
        return model
    return dispatch_model(model, device_map=device_map, offload_dir=offload_folder, offload_buffers=offload_buffers, skip_keys=skip_keys,
                          preload_module_classes=preload_module_classes, force_hooks=force_hooks)


# =========================================================================== storage-dtype casting
SUPPORTED_PYTORCH_LAYERS_FOR_UPCASTING = (nn.Conv1d, nn.Conv2d, nn.Conv3d, nn.ConvTranspose1d, nn.ConvTranspose2d, nn.ConvTranspose3d, nn.Linear)
_CASTABLE = SUPPORTED_PYTORCH_LAYERS_FOR_UPCASTING


def attach_layerwise_casting_hooks(module: nn.Module, storage_dtype: torch.dtype, compute_dtype: torch.dtype,
                                   skip_modules_pattern: Union[str, tuple] = None, skip_modules_classes: Optional[tuple] = None,
                                   non_blocking: bool = False) -> None:
    """Linear / convolution layers keep their weights in ``storage_dtype`` and are cast to ``compute_dtype`` only while they
    run.  A sub-tree is left alone when its class is in ``skip_modules_classes`` or its path matches one of the regular
    expressions in ``skip_modules_pattern`` (one expression or several).  Only those layer kinds are ever touched — embeddings and
    normalisation layers keep their dtype even when nothing is skipped (the reference casts the module as a whole in that case)."""
    patterns = (skip_modules_pattern,) if isinstance(skip_modules_pattern, str) else skip_modules_pattern

    def excluded(mod, path) -> bool:
        if skip_modules_classes is not None and isinstance(mod, skip_modules_classes):
            return True
        return patterns is not None and any(re.search(rx, path) for rx in patterns)

    stack = [(module, "")]
    while stack:
        mod, path = stack.pop()
        if excluded(mod, path):
            logger.debug(f'Skipping layerwise casting for layer "{path}"')
            continue
        if isinstance(mod, _CASTABLE) or type(mod).__name__ == "B200Linear":
            logger.debug(f'Applying layerwise casting to layer "{path}"')
            add_hook_to_module(mod, LayerwiseCastingHook(storage_dtype, compute_dtype, non_blocking), append=True)
            continue
        stack.extend((child, f"{path}.{name}" if path else name) for name, child in mod.named_children())


def _attach_context_parallel_hooks(model: nn.Module) -> None:
    """Context parallelism computes causal attention over sequence shards: a padding mask cannot be honoured, so every
    ``*.self_attn`` module has ``attention_mask`` dropped and ``is_causal=True`` forced before it runs."""

    def causal_only(_module, args, kwargs):
        if "attention_mask" in kwargs:
            kwargs["attention_mask"] = None
            kwargs["is_causal"] = True
        return args, kwargs

    for path, sub in model.named_modules():
        if path.endswith("self_attn"):
            sub.register_forward_pre_hook(causal_only, with_kwargs=True, prepend=True)
