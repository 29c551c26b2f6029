"""Assorted helpers: model unwrapping, rank-aware save/load, small system checks.

Reference parity: src/accelerate/utils/other.py — ``extract_model_from_parallel`` :174-260,
``save`` :354-385 (main process only, shared-tensor de-dup :319-351), ``load`` :388-433
(``weights_only`` with a numpy allow-list), ``is_port_in_use`` :478-487, ``convert_bytes`` :490-498,
``check_os_kernel`` :501-516, ``recursive_getattr`` :519-535, ``get_module_children_bottom_up``
:538-564, regional compilation :102-171.
"""

from __future__ import annotations

import collections
import logging
import os
import platform
import re
import socket
from contextlib import contextmanager
from functools import partial, reduce
from types import MethodType
from typing import Callable, OrderedDict

import numpy as np
import torch
from packaging.version import Version

from ._lazy_state import PartialState
from .constants import FSDP_PYTORCH_VERSION
from .dataclasses import DistributedType
from .imports import is_safetensors_available
from .versions import is_torch_version

logger = logging.getLogger(__name__)


@contextmanager
def clear_environment_for_defaults():
    """Used by ``KwargsHandler.to_kwargs``: build the default instance without ACCELERATE_* overrides."""
    saved = {k: v for k, v in os.environ.items() if k.startswith(("ACCELERATE_", "FSDP_", "PARALLELISM_CONFIG_", "MEGATRON_LM_"))}
    for k in saved:
        del os.environ[k]
    try:
        yield
    finally:
        os.environ.update(saved)


def is_compiled_module(module: torch.nn.Module) -> bool:
    return hasattr(torch, "_dynamo") and isinstance(module, torch._dynamo.eval_frame.OptimizedModule)


def has_compiled_regions(module: torch.nn.Module) -> bool:
    if not hasattr(torch, "_dynamo"):
        return False
    if module._modules:
        for submodule in module.modules():
            if isinstance(submodule, torch._dynamo.eval_frame.OptimizedModule):
                return True
    return False


def is_repeated_blocks(module: torch.nn.Module) -> bool:
    return isinstance(module, torch.nn.ModuleList) and all(isinstance(m, module[0].__class__) for m in module)


def has_repeated_blocks(module: torch.nn.Module) -> bool:
    if module._modules:
        for submodule in module.modules():
            if is_repeated_blocks(submodule):
                return True
    return False


def compile_regions(module: torch.nn.Module, **compile_kwargs) -> torch.nn.Module:
    """Compile repeated blocks (ModuleLists of one class) separately and the rest as a whole.

    Available for user modules that want it; the framework's own hot path does not use torch.compile.
    """
    new_module = _compile_regions(module, **compile_kwargs)
    if "_orig_mod" not in new_module.__dict__:
        new_module.__dict__["_orig_mod"] = module
    return new_module


def _compile_regions(module: torch.nn.Module, **compile_kwargs) -> torch.nn.Module:
    if is_repeated_blocks(module):
        new_module = torch.nn.ModuleList()
        for submodule in module:
            new_module.append(torch.compile(submodule, **compile_kwargs))
    elif has_repeated_blocks(module):
        new_module = module.__class__.__new__(module.__class__)
        new_module.__dict__.update(module.__dict__)
        new_module._modules = {}
        for name, submodule in module.named_children():
            new_module.add_module(name, _compile_regions(submodule, **compile_kwargs))
    else:
        new_module = torch.compile(module, **compile_kwargs)
    return new_module


def compile_regions_deepspeed(module: torch.nn.Module, **compile_kwargs):
    if is_repeated_blocks(module):
        for submodule in module:
            submodule.compile(**compile_kwargs)
    elif has_repeated_blocks(module):
        for child in module.children():
            compile_regions_deepspeed(child, **compile_kwargs)
    else:
        module.compile(**compile_kwargs)


def model_has_dtensor(model: torch.nn.Module) -> bool:
    from torch.distributed.tensor import DTensor

    return any(isinstance(p, DTensor) for p in model.parameters())


def _parallel_wrapper_types():
    from ..parallel.ddp import B200DataParallel
    from ..parallel.fsdp import B200ShardedModule

    return (torch.nn.parallel.DistributedDataParallel, torch.nn.DataParallel, B200DataParallel, B200ShardedModule)


def extract_model_from_parallel(model, keep_fp32_wrapper: bool = True, keep_torch_compile: bool = True, recursive: bool = False):
    """Strip the distributed wrappers (torch DDP, native DDP / sharded engines, torch.compile) off ``model``."""
    options = _parallel_wrapper_types()
    is_compiled = is_compiled_module(model)
    has_compiled = has_compiled_regions(model)
    compiled_model = None
    if is_compiled:
        compiled_model = model
        model = model._orig_mod
    elif has_compiled:
        compiled_model = model
        model = model.__dict__["_orig_mod"]
    while isinstance(model, options):
        model = model.module
    if recursive:
        def _recursive_unwrap(module):
            if hasattr(module, "module") and isinstance(module, options):
                unwrapped = _recursive_unwrap(module.module)
            else:
                unwrapped = module
            for name, child in unwrapped.named_children():
                setattr(unwrapped, name, _recursive_unwrap(child))
            return unwrapped

        model = _recursive_unwrap(model)
    if not keep_fp32_wrapper:
        forward = model.forward
        original_forward = model.__dict__.pop("_original_forward", None)
        if original_forward is not None:
            while hasattr(forward, "__wrapped__"):
                forward = forward.__wrapped__
                if forward == original_forward:
                    break
            model.forward = MethodType(forward, model)
        if getattr(model, "_converted_to_transformer_engine", False):
            from .transformer_engine import convert_model

            convert_model(model, to_transformer_engine=False)
    if keep_torch_compile and compiled_model is not None:
        if is_compiled:
            compiled_model._orig_mod = model
            model = compiled_model
        elif has_compiled:
            compiled_model.__dict__["_orig_mod"] = model
            model = compiled_model
    return model


def wait_for_everyone():
    """Barrier; equivalent to ``PartialState().wait_for_everyone()``."""
    PartialState().wait_for_everyone()


def _storage_id(t: torch.Tensor):
    try:
        return (t.device, t.untyped_storage().data_ptr(), t.storage_offset(), tuple(t.shape), tuple(t.stride()))
    except Exception:
        return id(t)


def clean_state_dict_for_safetensors(state_dict: dict):
    """safetensors refuses aliased tensors: keep one name per storage and make everything contiguous."""
    seen = {}
    drop = []
    for name, tensor in state_dict.items():
        if not isinstance(tensor, torch.Tensor):
            continue
        key = _storage_id(tensor)
        if key in seen:
            drop.append(name)
        else:
            seen[key] = name
    if drop:
        logger.warning(f"Removed shared tensor {set(drop)} while saving. This should be OK, but check by verifying that you don't receive any warning while reloading")
    out = {k: v for k, v in state_dict.items() if k not in drop}
    return {k: (v.contiguous() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}


def save(obj, f, save_on_each_node: bool = False, safe_serialization: bool = False):
    """Write ``obj`` to ``f`` once per job (or once per node): torch.save, or safetensors for state dicts."""
    if safe_serialization:
        from safetensors.torch import save_file

        save_func = partial(save_file, metadata={"format": "pt"})
        if isinstance(obj, (OrderedDict, dict)):
            obj = clean_state_dict_for_safetensors(obj)
    else:
        save_func = torch.save
    state = PartialState()
    if state.is_main_process and not save_on_each_node:
        save_func(obj, f)
    elif state.is_local_main_process and save_on_each_node:
        save_func(obj, f)


# numpy globals that legacy RNG-state pickles reference; allowed under weights_only=True
def _numpy_safe_globals():
    out = [np.dtype, np.ndarray]
    core = getattr(np, "_core", None) or getattr(np, "core")
    out.append(core.multiarray._reconstruct)
    try:
        out.append(type(np.dtype(np.uint32)))
        out.append(type(np.dtype(np.float64)))
        out.append(type(np.dtype(np.int64)))
    except Exception:  # pragma: no cover
        pass
    return out


def load(f, map_location=None, **kwargs):
    """``torch.load`` with ``weights_only=True`` by default and numpy reconstruction allowed."""
    if "weights_only" not in kwargs:
        kwargs["weights_only"] = True
    old = None
    try:
        if kwargs["weights_only"]:
            old = torch.serialization.get_safe_globals()
            torch.serialization.add_safe_globals(_numpy_safe_globals())
        return torch.load(f, map_location=map_location, **kwargs)
    finally:
        if old is not None:
            torch.serialization.clear_safe_globals()
            if old:
                torch.serialization.add_safe_globals(old)


def get_pretty_name(obj) -> str:
    if not hasattr(obj, "__qualname__") and not hasattr(obj, "__name__"):
        obj = getattr(obj, "__class__", obj)
    if hasattr(obj, "__qualname__"):
        return obj.__qualname__
    if hasattr(obj, "__name__"):
        return obj.__name__
    return str(obj)


def merge_dicts(source: dict, destination: dict) -> dict:
    """Recursively merge ``source`` into ``destination`` (source wins)."""
    for key, value in source.items():
        if isinstance(value, dict):
            node = destination.setdefault(key, {})
            merge_dicts(value, node)
        else:
            destination[key] = value
    return destination


def is_port_in_use(port: int = None) -> bool:
    if port is None:
        port = 29500
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        return s.connect_ex(("localhost", port)) == 0


def get_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("", 0))
        return s.getsockname()[1]


def convert_bytes(size) -> str:
    for unit in ["bytes", "KB", "MB", "GB", "TB"]:
        if size < 1024.0:
            return f"{round(size, 2)} {unit}"
        size /= 1024.0
    return f"{round(size, 2)} PB"


def check_os_kernel():
    """Warn on Linux kernels older than 5.5, which can hang multi-process jobs."""
    info = platform.uname()
    if info.system != "Linux":
        return
    m = re.search(r"(\d+\.\d+\.\d+)", info.release)
    if m is None:
        return
    if Version(m.group(1)) < Version("5.5.0"):
        logger.warning(
            f"Detected kernel version {m.group(1)}, which is below the recommended minimum of 5.5.0; this can cause "
            "the process to hang. It is recommended to upgrade the kernel to the minimum version or higher."
        )


def recursive_getattr(obj, attr: str):
    """``recursive_getattr(model, "layers.0.weight")``."""
    return reduce(lambda o, a: getattr(o, a), [obj] + attr.split("."))


def get_module_children_bottom_up(model: torch.nn.Module, return_fqns: bool = False):
    """All sub-modules, children before parents, root last."""
    out = []

    def visit(mod, name):
        for cname, child in mod.named_children():
            visit(child, f"{name}.{cname}" if name else cname)
        out.append((name, mod) if return_fqns else mod)

    visit(model, "")
    return out


def write_basic_config(mixed_precision="no", save_location: str = None):
    """Non-interactive default config file (``accelerate config default``)."""
    from ..commands.config.default import write_basic_config as _w

    return _w(mixed_precision=mixed_precision, save_location=save_location) if save_location else _w(mixed_precision=mixed_precision)
