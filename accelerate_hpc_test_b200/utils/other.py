"""Assorted helpers: model unwrapping, rank-aware save/load, small system checks.

Reference parity: src/accelerate/utils/other.py — ``extract_model_from_parallel`` :174-260,
``save`` :354-385 (main process only, shared-tensor de-dup :319-351), ``load`` :388-433
(``weights_only`` with a numpy allow-list), ``is_port_in_use`` :478-487, ``convert_bytes`` :490-498,
``check_os_kernel`` :501-516, ``recursive_getattr`` :519-535, ``get_module_children_bottom_up``
:538-564, regional compilation :102-171.
"""

from __future__ import annotations

import logging
import os
import platform
import re
import socket
from contextlib import contextmanager
from functools import reduce
from types import MethodType

import numpy as np
import torch
from packaging.version import Version

from ._lazy_state import PartialState

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


def _optimized_module_type():
    dynamo = getattr(torch, "_dynamo", None)
    return dynamo.eval_frame.OptimizedModule if dynamo is not None else ()


def is_compiled_module(module: torch.nn.Module) -> bool:
    """Is ``module`` what ``torch.compile(model)`` returned?"""
    return isinstance(module, _optimized_module_type())


def has_compiled_regions(module: torch.nn.Module) -> bool:
    """Does ``module`` contain compiled sub-modules (the result of ``compile_regions``)?"""
    kind = _optimized_module_type()
    return bool(kind) and bool(module._modules) and any(isinstance(sub, kind) for sub in module.modules())


def is_repeated_blocks(module: torch.nn.Module) -> bool:
    """A ``ModuleList`` whose members all have the class of the first — a stack of decoder layers, typically."""
    return isinstance(module, torch.nn.ModuleList) and all(isinstance(member, type(module[0])) for member in module)


def has_repeated_blocks(module: torch.nn.Module) -> bool:
    return bool(module._modules) and any(is_repeated_blocks(sub) for sub in module.modules())


def compile_regions(module: torch.nn.Module, **compile_kwargs) -> torch.nn.Module:
    """Regional compilation for user models: every member of a stack of identical blocks is compiled on its own (the compiler's
    cache makes the 2nd..Nth nearly free), everything else as one piece.  The original stays reachable as ``_orig_mod``.  (This
    framework's own hot path is hand-written kernels; nothing here depends on ``torch.compile``.)"""
    compiled = _compile_regions(module, **compile_kwargs)
    compiled.__dict__.setdefault("_orig_mod", module)
    return compiled


def _compile_regions(module: torch.nn.Module, **compile_kwargs) -> torch.nn.Module:
    if is_repeated_blocks(module):
        return torch.nn.ModuleList(torch.compile(block, **compile_kwargs) for block in module)
    if not has_repeated_blocks(module):
        return torch.compile(module, **compile_kwargs)
    # a shell of the same class sharing the attributes, with every child replaced by its regionally compiled version
    shell = type(module).__new__(type(module))
    shell.__dict__.update(module.__dict__)
    shell._modules = {}
    for name, child in module.named_children():
        shell.add_module(name, _compile_regions(child, **compile_kwargs))
    return shell


def compile_regions_deepspeed(module: torch.nn.Module, **compile_kwargs) -> None:
    """In-place variant (``nn.Module.compile``) for engines that hold on to the module objects."""
    if is_repeated_blocks(module):
        targets = list(module)
    elif has_repeated_blocks(module):
        for child in module.children():
            compile_regions_deepspeed(child, **compile_kwargs)
        return
    else:
        targets = [module]
    for target in targets:
        target.compile(**compile_kwargs)


def model_has_dtensor(model: torch.nn.Module) -> bool:
    from torch.distributed.tensor import DTensor

    return any(isinstance(p, DTensor) for p in model.parameters())


def _parallel_wrapper_types():
    from ..parallel.ddp import B200DataParallel
    from ..parallel.fsdp import B200ShardedModule

    return (torch.nn.parallel.DistributedDataParallel, torch.nn.DataParallel, B200DataParallel, B200ShardedModule)


def _peel(model, wrappers):
    while isinstance(model, wrappers):
        model = model.module
    return model


def _peel_everywhere(module, wrappers):
    module = _peel(module, wrappers)
    for name, child in module.named_children():
        setattr(module, name, _peel_everywhere(child, wrappers))
    return module


def _restore_plain_forward(model) -> None:
    """Undo what ``prepare`` did to ``model.forward`` (autocast / fp32-output decorators, TransformerEngine layers)."""
    original = model.__dict__.pop("_original_forward", None)
    if original is not None:
        forward = model.forward
        while hasattr(forward, "__wrapped__") and forward != original:
            forward = forward.__wrapped__
        model.forward = MethodType(forward, model)
    if getattr(model, "_converted_to_transformer_engine", False):
        from .transformer_engine import convert_model

        convert_model(model, to_transformer_engine=False)


def extract_model_from_parallel(model, keep_fp32_wrapper: bool = True, keep_torch_compile: bool = True, recursive: bool = False):
    """The user's module back out of the wrappers ``prepare`` put around it: torch DDP / DataParallel, the native data-parallel and
    sharded engines, ``torch.compile`` (kept on top unless ``keep_torch_compile=False``).  ``recursive`` also unwraps wrapped
    sub-modules; ``keep_fp32_wrapper=False`` additionally restores the undecorated ``forward``."""
    wrappers = _parallel_wrapper_types()
    compiled_outer = model if (is_compiled_module(model) or has_compiled_regions(model)) else None
    whole_model_compiled = compiled_outer is not None and is_compiled_module(model)
    if compiled_outer is not None:
        model = model._orig_mod if whole_model_compiled else model.__dict__["_orig_mod"]
    model = _peel_everywhere(model, wrappers) if recursive else _peel(model, wrappers)
    if not keep_fp32_wrapper:
        _restore_plain_forward(model)
    if compiled_outer is None or not keep_torch_compile:
        return model
    if whole_model_compiled:
        compiled_outer._orig_mod = model
    else:
        compiled_outer.__dict__["_orig_mod"] = model
    return compiled_outer


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
    """Write ``obj`` to ``f`` from one process only — the job's main process, or each node's with ``save_on_each_node`` (node-local
    disks).  ``safe_serialization`` writes safetensors (state dicts are de-aliased first), otherwise ``torch.save``."""
    state = PartialState()
    writer = state.is_local_main_process if save_on_each_node else state.is_main_process
    if safe_serialization and isinstance(obj, dict):
        obj = clean_state_dict_for_safetensors(obj)
    if not writer:
        return
    if safe_serialization:
        from safetensors.torch import save_file

        save_file(obj, f, metadata={"format": "pt"})
    else:
        torch.save(obj, f)


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


# numpy globals that RNG-state pickles reference; allowed while `load` unpickles under weights_only=True
TORCH_SAFE_GLOBALS = _numpy_safe_globals()


def load(f, map_location=None, **kwargs):
    """``torch.load`` with ``weights_only=True`` by default and numpy reconstruction allowed."""
    if "weights_only" not in kwargs:
        kwargs["weights_only"] = True
    old = None
    try:
        if kwargs["weights_only"]:
            old = torch.serialization.get_safe_globals()
            torch.serialization.add_safe_globals(TORCH_SAFE_GLOBALS)
        return torch.load(f, map_location=map_location, **kwargs)
    finally:
        if old is not None:
            torch.serialization.clear_safe_globals()
            if old:
                torch.serialization.add_safe_globals(old)


def get_pretty_name(obj) -> str:
    """Qualified name of a function / class, or of the class of an instance."""
    for candidate in (obj, type(obj)):
        for attribute in ("__qualname__", "__name__"):
            if hasattr(candidate, attribute):
                return getattr(candidate, attribute)
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
        return s.connect_ex(("127.0.0.1", port)) == 0


def get_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("", 0))
        return s.getsockname()[1]


def convert_bytes(size) -> str:
    """``1536`` -> ``"1.5 KB"`` (binary steps, decimal names, as the reference prints them)."""
    units = ("bytes", "KB", "MB", "GB", "TB", "PB")
    step = 0
    while size >= 1024.0 and step < len(units) - 1:
        size /= 1024.0
        step += 1
    return f"{round(size, 2)} {units[step]}"


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
