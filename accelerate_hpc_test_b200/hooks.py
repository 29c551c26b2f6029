"""Module hooks: code that runs around ``module.forward`` to put weights and inputs where the computation happens.

Public surface (behaviour parity with reference ``src/accelerate/hooks.py``; the implementation is organised
differently):

* ``ModelHook`` — the four-callback protocol (``init_hook`` / ``pre_forward`` / ``post_forward`` / ``detach_hook``,
  plus the ``no_grad`` switch).  ``SequentialHook`` chains several.
* ``add_hook_to_module`` / ``remove_hook_from_module`` — install / undo.  The original forward is kept on the module
  as ``_old_forward`` and the active hook as ``_hf_hook`` (names other libraries look for).
* ``AlignDevicesHook`` — device placement.  Internally it picks one of two *placement policies*:
  ``_Resident`` (weights are moved to the execution device once) or ``_Streamed`` (weights are parked on the
  ``meta`` device, fetched from a ``weights_map`` right before every forward and released right after — CPU / disk
  offload).  Inputs always follow the weights; outputs can be sent back to where the inputs came from.
* ``attach_execution_device_hook``, ``attach_align_device_hook``, ``attach_align_device_hook_on_blocks`` — walk a
  model and install the hooks that realise a device map.
* ``CpuOffload`` + ``UserCpuOffloadHook`` — whole-model ping-pong between CPU and one device, for pipelines of models.
* ``LayerwiseCastingHook`` — keep weights in a small storage dtype, up-cast for the forward.
"""

from __future__ import annotations

import functools
from collections.abc import Mapping
from typing import Iterator, Optional, Union

import torch
import torch.nn as nn

from .state import PartialState
from .utils.modeling import get_non_persistent_buffers, named_module_tensors, set_module_tensor_to_device
from .utils.operations import find_device, send_to_device
from .utils.offload import PrefixedDataset
from .utils.other import recursive_getattr

META = torch.device("meta")
# methods `dispatch_model` shadows on the instance to warn about moving a dispatched model; removed with the hooks
_accelerate_added_attributes = ("to", "cuda")


# =========================================================================== protocol
class ModelHook:
    """Base class: override what you need.  With ``no_grad = True`` the wrapped forward runs under ``torch.no_grad()``."""

    no_grad = False

    def init_hook(self, module: nn.Module) -> nn.Module:
        """Called once when the hook is installed; may replace the module."""
        return module

    def pre_forward(self, module: nn.Module, *args, **kwargs):
        """Called before every forward; returns the (possibly changed) ``(args, kwargs)``."""
        return args, kwargs

    def post_forward(self, module: nn.Module, output):
        """Called after every forward; returns the (possibly changed) output."""
        return output

    def detach_hook(self, module: nn.Module) -> nn.Module:
        """Called when the hook is removed."""
        return module


class SequentialHook(ModelHook):
    """Runs its hooks first-to-last in every phase."""

    def __init__(self, *hooks: ModelHook):
        self.hooks = hooks

    def _thread_module(self, method: str, module):
        for h in self.hooks:
            module = getattr(h, method)(module)
        return module

    def init_hook(self, module):
        return self._thread_module("init_hook", module)

    def detach_hook(self, module):
        return self._thread_module("detach_hook", module)

    def pre_forward(self, module, *args, **kwargs):
        for h in self.hooks:
            args, kwargs = h.pre_forward(module, *args, **kwargs)
        return args, kwargs

    def post_forward(self, module, output):
        for h in self.hooks:
            output = h.post_forward(module, output)
        return output


# =========================================================================== installation
def _is_fx_graph_module(module) -> bool:
    # torch.fx generates a per-instance class and looks `forward` up on the class
    return "GraphModuleImpl" in str(type(module))


def _hooked_forward(module: nn.Module, *args, **kwargs):
    hook = module._hf_hook
    args, kwargs = hook.pre_forward(module, *args, **kwargs)
    ctx = torch.no_grad() if hook.no_grad else _NULL
    with ctx:
        result = module._old_forward(*args, **kwargs)
    return hook.post_forward(module, result)


class _NullContext:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NULL = _NullContext()


def _set_forward(module: nn.Module, fn) -> None:
    target = type(module) if _is_fx_graph_module(module) else module
    target.forward = fn


def add_hook_to_module(module: nn.Module, hook: ModelHook, append: bool = False) -> nn.Module:
    """Install ``hook``.  A hook already present is replaced, or kept in front of the new one with ``append=True``."""
    existing = getattr(module, "_hf_hook", None)
    if append and existing is not None:
        remove_hook_from_module(module)
        hook = SequentialHook(existing, hook)
    if not (hasattr(module, "_hf_hook") and hasattr(module, "_old_forward")):
        module._old_forward = module.forward
    original = module._old_forward
    module = hook.init_hook(module)
    module._hf_hook = hook
    _set_forward(module, functools.update_wrapper(functools.partial(_hooked_forward, module), original))
    return module


def remove_hook_from_module(module: nn.Module, recurse: bool = False) -> nn.Module:
    hook = module.__dict__.pop("_hf_hook", None)
    if hook is not None:
        hook.detach_hook(module)
    original = module.__dict__.pop("_old_forward", None)
    if original is not None:
        _set_forward(module, original)
    for name in _accelerate_added_attributes:
        module.__dict__.pop(name, None)
    if recurse:
        for child in module.children():
            remove_hook_from_module(child, recurse=True)
    return module


def remove_hook_from_submodules(module: nn.Module) -> None:
    for m in module.modules():
        remove_hook_from_module(m)


# =========================================================================== placement policies
def _as_device(d):
    return torch.device("cuda", d) if isinstance(d, int) else d


class _Resident:
    """Weights live on the execution device from installation on."""

    def __init__(self, hook: "AlignDevicesHook"):
        self.h = hook

    def install(self, module):
        h = self.h
        if h.execution_device is None:
            return
        for name, _ in named_module_tensors(module, recurse=h.place_submodules):
            set_module_tensor_to_device(module, name, h.execution_device, tied_params_map=h.tied_params_map)

    def before(self, module):
        pass

    def after(self, module):
        pass

    def restore(self, module):
        pass


class _Streamed:
    """Weights are parked on ``meta``; each forward fetches them from ``weights_map`` and drops them again.

    Tied weights (one stored tensor used by several modules) are materialised once per device: ``tied_params_map``
    maps the stored tensor's ``data_ptr`` to ``{device: live tensor}``; the module that created an entry deletes it
    after its forward, the others just borrow it.
    """

    def __init__(self, hook: "AlignDevicesHook"):
        self.h = hook
        self._created: set = set()

    def _streamed_names(self, module) -> Iterator[str]:
        h = self.h
        for name, _ in named_module_tensors(module, include_buffers=h.offload_buffers, recurse=h.place_submodules,
                                            remove_non_persistent=True):
            yield name

    def install(self, module):
        h = self.h
        h.original_devices = {n: t.device for n, t in named_module_tensors(module, recurse=h.place_submodules)}
        if h.weights_map is None:
            h.weights_map = {n: t.to("cpu") for n, t in named_module_tensors(module, include_buffers=h.offload_buffers,
                                                                           recurse=h.place_submodules)}
        for name in list(self._streamed_names(module)):
            if h.tied_params_map is not None and recursive_getattr(module, name).data_ptr() in h.tied_params_map:
                h.tied_params_names.add(name)
            set_module_tensor_to_device(module, name, "meta")
        if h.execution_device is None:
            return
        # buffers that are not streamed must still sit next to the computation
        if h.offload_buffers:
            keep = get_non_persistent_buffers(module, recurse=h.place_submodules)
        else:
            keep = [n for n, _ in module.named_buffers(recurse=h.place_submodules)]
        for name in keep:
            set_module_tensor_to_device(module, name, h.execution_device, tied_params_map=h.tied_params_map)

    def _int8_statistics(self, name: str, stored):
        # bitsandbytes 8-bit layers keep their quantisation statistics under "<...>SCB" next to the int8 weight
        if "weight" not in name or stored.dtype != torch.int8:
            return None
        return self.h.weights_map[name.replace("weight", "SCB")] if name.replace("weight", "SCB") in self.h.weights_map.keys() else None

    def before(self, module):
        h = self.h
        self._created = set()
        shared = h.tied_params_map
        for name in self._streamed_names(module):
            stored = h.weights_map[name]
            if shared is not None and stored is not None:
                key = stored.data_ptr()
                if name in h.tied_params_names:
                    shared.setdefault(key, {})
                if key in shared and h.execution_device not in shared[key]:
                    self._created.add((key, h.execution_device))
            set_module_tensor_to_device(module, name, h.execution_device, value=stored,
                                        fp16_statistics=self._int8_statistics(name, stored), tied_params_map=shared)

    def after(self, module):
        h = self.h
        for name in self._streamed_names(module):
            set_module_tensor_to_device(module, name, "meta")
        if type(module).__name__ == "Linear8bitLt":
            module.state.SCB = None
            module.state.CxB = None
        for key, device in self._created:
            live = h.tied_params_map.get(key, {})
            live.pop(_as_device(device), None)
            live.pop(device, None)
        self._created = set()

    def restore(self, module):
        h = self.h
        for name, device in h.original_devices.items():
            if device != META:
                set_module_tensor_to_device(module, name, device, value=h.weights_map.get(name, None))


class AlignDevicesHook(ModelHook):
    """Make a module compute on ``execution_device`` whatever device its inputs arrive on.

    Args mirror the reference: ``offload`` (stream weights from ``weights_map``), ``io_same_device`` (send the
    output back to the device of the first tensor found in the inputs), ``offload_buffers``, ``place_submodules``
    (treat the whole sub-tree as one unit instead of only the module's own tensors), ``skip_keys`` (kwargs left where
    they are) and ``tied_params_map`` (see ``_Streamed``).
    """

    def __init__(self, execution_device: Optional[Union[int, str, torch.device]] = None, offload: bool = False,
                 io_same_device: bool = False, weights_map: Optional[Mapping] = None, offload_buffers: bool = False,
                 place_submodules: bool = False, skip_keys: Optional[Union[str, list]] = None,
                 tied_params_map: Optional[dict] = None):
        self.execution_device = execution_device
        self.offload = offload
        self.io_same_device = io_same_device
        self.weights_map = weights_map
        self.offload_buffers = offload_buffers
        self.place_submodules = place_submodules
        self.skip_keys = skip_keys
        self.tied_params_map = tied_params_map
        self.tied_params_names: set = set()
        self.original_devices: dict = {}
        self.input_device = None
        self._policy = _Streamed(self) if offload else _Resident(self)

    def __repr__(self):
        fields = ("execution_device", "offload", "io_same_device", "offload_buffers", "place_submodules")
        body = ", ".join(f"{f}={getattr(self, f)}" for f in fields)
        return f"AlignDevicesHook({body}, skip_keys={self.skip_keys!r})"

    def init_hook(self, module):
        if self.execution_device is not None and torch.device(_as_device(self.execution_device)) == META:
            self.tied_params_map = None  # nothing is ever materialised on meta: no sharing to track
        self._policy.install(module)
        return module

    def pre_forward(self, module, *args, **kwargs):
        if self.io_same_device:
            self.input_device = find_device([args, kwargs])
        self._policy.before(module)
        moved_args = send_to_device(args, self.execution_device)
        moved_kwargs = send_to_device(kwargs, self.execution_device, skip_keys=self.skip_keys)
        return moved_args, moved_kwargs

    def post_forward(self, module, output):
        self._policy.after(module)
        if self.io_same_device and self.input_device is not None:
            return send_to_device(output, self.input_device, skip_keys=self.skip_keys)
        return output

    def detach_hook(self, module):
        self._policy.restore(module)
        return module


# =========================================================================== realising a device map
def _child_path(parent: str, child: str) -> str:
    return f"{parent}.{child}" if parent else child


def _is_preloaded(module, preload_module_classes) -> bool:
    return preload_module_classes is not None and type(module).__name__ in preload_module_classes


def attach_execution_device_hook(module: nn.Module, execution_device, skip_keys=None, preload_module_classes=None,
                                 tied_params_map=None) -> None:
    """Every sub-module that owns state and has no hook yet gets an input-moving hook for ``execution_device``.
    Sub-trees rooted at a class named in ``preload_module_classes`` are treated as one unit (not descended into)."""
    pending = [module]
    while pending:
        current = pending.pop()
        if not hasattr(current, "_hf_hook") and len(current.state_dict()) > 0:
            add_hook_to_module(current, AlignDevicesHook(execution_device, skip_keys=skip_keys, tied_params_map=tied_params_map))
        if _is_preloaded(current, preload_module_classes):
            continue
        pending.extend(current.children())


def attach_align_device_hook(module: nn.Module, execution_device=None, offload: bool = False,
                             weights_map: Optional[Mapping] = None, offload_buffers: bool = False, module_name: str = "",
                             skip_keys=None, preload_module_classes=None, tied_params_map=None) -> None:
    """Install one ``AlignDevicesHook`` on every module that owns tensors *directly* (so weights are fetched at the
    finest granularity).  A class in ``preload_module_classes`` is offloaded as a whole instead."""
    pending = [(module, module_name)]
    while pending:
        current, path = pending.pop()
        whole_tree = offload and _is_preloaded(current, preload_module_classes)
        owns_tensors = next(iter(named_module_tensors(current)), None) is not None
        if owns_tensors or whole_tree:
            local_map = None if weights_map is None else PrefixedDataset(weights_map, prefix=f"{path}.")
            add_hook_to_module(
                current,
                AlignDevicesHook(execution_device=execution_device, offload=offload, weights_map=local_map,
                                 offload_buffers=offload_buffers, place_submodules=whole_tree, skip_keys=skip_keys,
                                 tied_params_map=tied_params_map),
                append=True,
            )
        if whole_tree:
            continue
        pending.extend((child, _child_path(path, name)) for name, child in current.named_children())


def attach_align_device_hook_on_blocks(module: nn.Module, execution_device=None, offload=False, weights_map: Mapping = None,
                                       offload_buffers: bool = False, module_name: str = "", skip_keys=None,
                                       preload_module_classes=None, tied_params_map=None) -> None:
    """Realise a device map.  ``execution_device`` and ``offload`` may be single values (whole model) or mappings
    ``module path -> value``.  A block named in the map becomes one placement unit:

    * not offloaded → its weights move to its device now; every state-owning descendant gets an input-moving hook;
    * offloaded → its tensor-owning descendants stream their own weights; the block itself at least moves inputs.

    The root additionally returns outputs to the caller's device (``io_same_device``).
    """
    uniform = not isinstance(execution_device, Mapping) and not isinstance(offload, dict)
    if uniform:
        if offload:
            attach_align_device_hook(module, execution_device=execution_device, offload=True, weights_map=weights_map,
                                     offload_buffers=offload_buffers, module_name=module_name, skip_keys=skip_keys,
                                     tied_params_map=tied_params_map)
        else:
            add_hook_to_module(module, AlignDevicesHook(execution_device=execution_device, io_same_device=True, skip_keys=skip_keys,
                                                        place_submodules=True, tied_params_map=tied_params_map))
        return
    devices = execution_device if isinstance(execution_device, Mapping) else dict.fromkeys(offload, execution_device)
    offloaded = offload if isinstance(offload, Mapping) else dict.fromkeys(devices, offload)

    pending = [(module, module_name)]
    while pending:
        current, path = pending.pop()
        is_root = path == ""
        if path in devices and path in offloaded:
            device = devices[path]
            if offloaded[path]:
                attach_align_device_hook(current, execution_device=device, offload=True, weights_map=weights_map,
                                         offload_buffers=offload_buffers, module_name=path, skip_keys=skip_keys,
                                         preload_module_classes=preload_module_classes, tied_params_map=tied_params_map)
                if not hasattr(current, "_hf_hook"):
                    add_hook_to_module(current, AlignDevicesHook(execution_device=device, io_same_device=is_root, skip_keys=skip_keys,
                                                                 tied_params_map=tied_params_map))
                attach_execution_device_hook(current, device, preload_module_classes=preload_module_classes, skip_keys=skip_keys,
                                             tied_params_map=tied_params_map)
            else:
                add_hook_to_module(current, AlignDevicesHook(execution_device=device, offload_buffers=offload_buffers,
                                                             io_same_device=is_root, place_submodules=True, skip_keys=skip_keys,
                                                             tied_params_map=tied_params_map))
                attach_execution_device_hook(current, device, skip_keys=skip_keys, tied_params_map=tied_params_map)
        elif is_root:
            add_hook_to_module(current, AlignDevicesHook(execution_device=devices.get(""), io_same_device=True, skip_keys=skip_keys,
                                                         tied_params_map=tied_params_map))
        pending.extend((child, _child_path(path, name)) for name, child in current.named_children())


# =========================================================================== whole-model CPU offload
class CpuOffload(ModelHook):
    """The model rests on the CPU and is moved to ``execution_device`` when it is called.  If ``prev_module_hook`` is
    given, the model behind that handle is sent back to the CPU first (so a chain of models shares one device)."""

    def __init__(self, execution_device=None, prev_module_hook: Optional["UserCpuOffloadHook"] = None):
        self.prev_module_hook = prev_module_hook
        self.execution_device = PartialState().default_device if execution_device is None else execution_device

    def init_hook(self, module):
        return module.to("cpu")

    def _evict_previous(self) -> None:
        prev = self.prev_module_hook
        if not isinstance(prev, UserCpuOffloadHook):
            return
        if next(prev.model.parameters()).device.type != "cpu":
            prev.offload()

    def pre_forward(self, module, *args, **kwargs):
        self._evict_previous()
        if next(module.parameters()).device == self.execution_device:
            return args, kwargs
        module.to(self.execution_device)
        return send_to_device(args, self.execution_device), send_to_device(kwargs, self.execution_device)


class UserCpuOffloadHook:
    """What ``cpu_offload_with_hook`` hands back: ``offload()`` returns the model to the CPU, ``remove()`` uninstalls."""

    def __init__(self, model: nn.Module, hook: ModelHook):
        self.model = model
        self.hook = hook

    def offload(self) -> None:
        self.hook.init_hook(self.model)

    def remove(self) -> None:
        remove_hook_from_module(self.model)


# =========================================================================== storage-dtype casting
class LayerwiseCastingHook(ModelHook):
    """Weights rest in ``storage_dtype`` (e.g. ``torch.float8_e4m3fn``) and are cast to ``compute_dtype`` only while the
    module runs."""

    _is_stateful = False

    def __init__(self, storage_dtype: torch.dtype, compute_dtype: torch.dtype, non_blocking: bool) -> None:
        self.storage_dtype = storage_dtype
        self.compute_dtype = compute_dtype
        self.non_blocking = non_blocking

    def _cast(self, module: nn.Module, dtype: torch.dtype) -> None:
        module.to(dtype=dtype, non_blocking=self.non_blocking)

    def init_hook(self, module):
        self._cast(module, self.storage_dtype)
        return module

    def pre_forward(self, module, *args, **kwargs):
        self._cast(module, self.compute_dtype)
        return args, kwargs

    def post_forward(self, module, output):
        self._cast(module, self.storage_dtype)
        return output
