"""Sharded-training helpers under the reference's names (``accelerate.utils.fsdp_utils``).

Two kinds of sharded models reach these functions:

* the native engine (:class:`..parallel.fsdp.B200ShardedModule`) — the normal case on B200 nodes;
* a torch ``fully_shard`` model — the fallback ``Accelerator`` takes for jobs the native engine does not cover
  (multi-node groups, fp32 compute, CPU offload).

Both write the same files: ``pytorch_model_fsdp_{i}/`` and ``optimizer_{i}/`` are ``torch.distributed.checkpoint`` directories
holding ``{"model": ...}`` / ``{"optimizer": ...}`` (SHARDED_STATE_DICT), ``pytorch_model_fsdp.bin`` / ``optimizer.bin`` are rank-0
``torch.save`` files (FULL_STATE_DICT) — the layout of reference ``utils/fsdp_utils.py:103-335``, so checkpoints interchange.
``merge_fsdp_weights`` (``accelerate merge-weights``) folds a sharded model directory into one file without a process group.
"""

from __future__ import annotations

import os
import shutil
from pathlib import Path
from typing import Optional

import torch

from .constants import FSDP_MODEL_NAME, OPTIMIZER_NAME, SAFE_WEIGHTS_NAME, WEIGHTS_NAME
from .dataclasses import get_module_class_from_name  # noqa: F401  (re-exported, reference API)
from .other import get_module_children_bottom_up, save  # noqa: F401


# --------------------------------------------------------------------------- env switches
def enable_fsdp_ram_efficient_loading():
    """Only rank 0 materialises pretrained weights; the other ranks build on ``meta`` and receive their shards by broadcast."""
    os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] = "True"


def disable_fsdp_ram_efficient_loading():
    os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] = "False"


# --------------------------------------------------------------------------- which engine is this?
def _native_engine(model) -> Optional[torch.nn.Module]:
    from ..parallel.fsdp import B200ShardedModule

    seen = model
    for _ in range(8):  # DDP / compile / checkpoint wrappers nest only a few levels
        if isinstance(seen, B200ShardedModule):
            return seen
        inner = getattr(seen, "module", None)
        if inner is None or inner is seen:
            break
        seen = inner
    return None


def _engine(model):
    engine = _native_engine(model)
    if engine is None:
        raise TypeError("expected a model prepared with the sharded engine (B200ShardedModule)")
    return engine


def _sdt(plugin) -> str:
    return str(getattr(plugin, "state_dict_type", "SHARDED_STATE_DICT")).split(".")[-1]


def _file(stem: str, index: int, suffix: str = ".bin") -> str:
    return f"{stem}{suffix}" if index == 0 else f"{stem}_{index}{suffix}"


# --------------------------------------------------------------------------- torch FSDP2 fallback I/O
def _torch_model_io(plugin, model, folder: str, index: int, saving: bool):
    import torch.distributed as dist
    import torch.distributed.checkpoint as dist_cp
    from torch.distributed.checkpoint.state_dict import StateDictOptions, get_model_state_dict, set_model_state_dict

    full = _sdt(plugin) == "FULL_STATE_DICT"
    path = os.path.join(folder, _file(FSDP_MODEL_NAME, index)) if full else os.path.join(folder, f"{FSDP_MODEL_NAME}_{index}")
    if saving:
        sd = get_model_state_dict(model, options=StateDictOptions(full_state_dict=full, cpu_offload=full))
        if full:
            if dist.get_rank() == 0:
                torch.save(sd, path)
            dist.barrier()
        else:
            os.makedirs(path, exist_ok=True)
            dist_cp.save({"model": sd}, storage_writer=dist_cp.FileSystemWriter(path))
        return
    if full or not os.path.isdir(path):
        sd = torch.load(os.path.join(folder, _file(FSDP_MODEL_NAME, index)), map_location="cpu", weights_only=True)
        set_model_state_dict(model, sd, options=StateDictOptions(full_state_dict=True, broadcast_from_rank0=False))
        return
    holder = {"model": get_model_state_dict(model)}
    dist_cp.load(holder, storage_reader=dist_cp.FileSystemReader(path))
    set_model_state_dict(model, holder["model"])


def _torch_optimizer_io(plugin, model, optimizer, folder: str, index: int, saving: bool):
    import torch.distributed as dist
    import torch.distributed.checkpoint as dist_cp
    from torch.distributed.checkpoint.state_dict import StateDictOptions, get_optimizer_state_dict, set_optimizer_state_dict

    inner = getattr(optimizer, "optimizer", optimizer)
    full = _sdt(plugin) == "FULL_STATE_DICT"
    path = os.path.join(folder, _file(OPTIMIZER_NAME, index)) if full else os.path.join(folder, f"{OPTIMIZER_NAME}_{index}")
    if saving:
        osd = get_optimizer_state_dict(model, inner, options=StateDictOptions(full_state_dict=full, cpu_offload=full))
        if full:
            if dist.get_rank() == 0:
                torch.save(osd, path)
            dist.barrier()
        else:
            os.makedirs(path, exist_ok=True)
            dist_cp.save({"optimizer": osd}, storage_writer=dist_cp.FileSystemWriter(path))
        return
    if full or not os.path.isdir(path):
        osd = torch.load(os.path.join(folder, _file(OPTIMIZER_NAME, index)), map_location="cpu", weights_only=False)
        set_optimizer_state_dict(model, inner, osd, options=StateDictOptions(full_state_dict=True))
        return
    holder = {"optimizer": get_optimizer_state_dict(model, inner)}
    dist_cp.load(holder, storage_reader=dist_cp.FileSystemReader(path))
    set_optimizer_state_dict(model, inner, holder["optimizer"])


# --------------------------------------------------------------------------- public save / load
def save_fsdp_model(fsdp_plugin, accelerator, model, output_dir, model_index=0, adapter_only=False):
    os.makedirs(output_dir, exist_ok=True)
    engine = _native_engine(model)
    if engine is not None:
        engine.save_sharded_state(output_dir, model_index, fsdp_plugin)
    else:
        _torch_model_io(fsdp_plugin, model, output_dir, model_index, saving=True)


def load_fsdp_model(fsdp_plugin, accelerator, model, input_dir, model_index=0, adapter_only=False):
    accelerator.wait_for_everyone()
    engine = _native_engine(model)
    if engine is not None:
        engine.load_sharded_state(input_dir, model_index, fsdp_plugin)
    else:
        _torch_model_io(fsdp_plugin, model, input_dir, model_index, saving=False)


def save_fsdp_optimizer(fsdp_plugin, accelerator, optimizer, model, output_dir, optimizer_index=0):
    os.makedirs(output_dir, exist_ok=True)
    engine = _native_engine(model)
    if engine is not None:
        engine.save_sharded_optimizer(optimizer, output_dir, optimizer_index, fsdp_plugin)
    else:
        _torch_optimizer_io(fsdp_plugin, model, optimizer, output_dir, optimizer_index, saving=True)


def load_fsdp_optimizer(fsdp_plugin, accelerator, optimizer, model, input_dir, optimizer_index=0, adapter_only=False):
    accelerator.wait_for_everyone()
    engine = _native_engine(model)
    if engine is not None:
        engine.load_sharded_optimizer(optimizer, input_dir, optimizer_index, fsdp_plugin)
    else:
        _torch_optimizer_io(fsdp_plugin, model, optimizer, input_dir, optimizer_index, saving=False)


# --------------------------------------------------------------------------- merge-weights
def _distributed_checkpoint_to_state_dict(checkpoint_dir: str) -> dict:
    """Full ``name -> tensor`` dict of a DCP model directory, read on one process."""
    from ..parallel.fsdp import merge_sharded_checkpoint

    return merge_sharded_checkpoint(checkpoint_dir)


def _locate_model_dir(root: Path) -> Path:
    if (root / ".metadata").exists():
        return root
    nested = root / f"{FSDP_MODEL_NAME}_0"
    if nested.is_dir() and (nested / ".metadata").exists():
        return nested
    raise ValueError(
        f"Tried to load from {root} but couldn't find a valid metadata file, and no `{FSDP_MODEL_NAME}_0` sub-directory either."
    )


def merge_fsdp_weights(checkpoint_dir: str, output_path: str, safe_serialization: bool = True, remove_checkpoint_dir: bool = False):
    """``pytorch_model_fsdp_0/`` (or a ``save_state`` folder that contains it) → ``model.safetensors`` / ``pytorch_model.bin``
    under ``output_path``.  Only the main process works; the others wait."""
    from ..state import PartialState

    root = Path(checkpoint_dir)
    if not root.is_dir():
        raise ValueError(f"Tried to merge the weights from `{root}` but it is not a directory.")
    source = _locate_model_dir(root)
    state = PartialState()
    if state.is_main_process:
        tensors = _distributed_checkpoint_to_state_dict(str(source))
        target_dir = Path(output_path)
        target_dir.mkdir(parents=True, exist_ok=True)
        save(tensors, target_dir / (SAFE_WEIGHTS_NAME if safe_serialization else WEIGHTS_NAME), safe_serialization=safe_serialization)
        if remove_checkpoint_dir:
            shutil.rmtree(source)
    state.wait_for_everyone()


# --------------------------------------------------------------------------- FSDP1 helper kept for API parity
def ensure_weights_retied(param_init_fn, model: torch.nn.Module, device: torch.device):
    """Wrap a per-module ``param_init_fn`` so that weights the model declares tied (``_tied_weights_keys``) end up being the
    *same* Parameter again after every module has been materialised separately."""
    tied_keys = getattr(model, "_tied_weights_keys", None)
    if not tied_keys:
        return param_init_fn
    canonical: dict = {}
    for dotted in tied_keys:
        owner_path, _, leaf = dotted.rpartition(".")
        canonical[id(getattr(model.get_submodule(owner_path), leaf))] = None

    def init_and_retie(module: torch.nn.Module):
        mine = [(name, id(p)) for name, p in module.named_parameters(recurse=False) if id(p) in canonical]
        module = param_init_fn(module)
        for name, key in mine:
            if canonical[key] is None:
                canonical[key] = getattr(module, name)
            else:
                setattr(module, name, canonical[key])
        return module

    return init_and_retie


# --------------------------------------------------------------------------- FSDP2 entry points of the reference
def fsdp2_prepare_model(accelerator, model: torch.nn.Module) -> torch.nn.Module:
    """Shard ``model`` (native engine, or torch ``fully_shard`` when the job is outside its envelope)."""
    return accelerator._wrap_sharded(model)


def fsdp2_load_full_state_dict(accelerator, model: torch.nn.Module, full_sd: dict, cpu_offload: bool = False):
    """Distribute a full state dict over the shards.  With ``cpu_ram_efficient_loading`` only rank 0 passes values (the other
    ranks pass an empty dict) and every parameter travels by broadcast."""
    engine = _native_engine(model)
    if engine is not None:
        engine.load_full_state_dict(full_sd or {})
        return
    from torch.distributed.checkpoint.state_dict import StateDictOptions, set_model_state_dict

    set_model_state_dict(model, full_sd, options=StateDictOptions(full_state_dict=True, broadcast_from_rank0=True, cpu_offload=cpu_offload))


def fsdp2_switch_optimizer_parameters(optimizer: torch.optim.Optimizer, mapping: dict):
    """Re-point an optimizer built before sharding: ``mapping`` goes from the old parameters' ``data_ptr`` to the sharded ones."""
    for group in optimizer.param_groups:
        try:
            group["params"] = [mapping[p.data_ptr] for p in group["params"]]
        except KeyError:
            raise KeyError(
                "A parameter in the optimizer couldn't be switched to its sharded version. This breaks the training. Please raise an "
                "issue on GitHub."
            ) from None


def fsdp2_apply_ac(accelerator, model: torch.nn.Module):
    """Activation checkpointing on every module the auto-wrap policy selects."""
    from ..parallel.fsdp import apply_activation_checkpointing

    plugin = accelerator.state.fsdp_plugin
    plugin.set_auto_wrap_policy(model)
    apply_activation_checkpointing(model, plugin)
    return model


def fsdp2_prepare_auto_wrap_policy(fsdp2_plugin, model: torch.nn.Module):
    fsdp2_plugin.set_auto_wrap_policy(model)
    return fsdp2_plugin.should_wrap


def get_fsdp2_grad_scaler(**kwargs):
    return torch.amp.GradScaler("cuda" if torch.cuda.is_available() else "cpu", **kwargs)


_WRAPPER_INFIXES = ("._checkpoint_wrapped_module", "._orig_mod")


def fsdp2_canonicalize_names(named_params: dict) -> dict:
    """Parameter names without wrapper noise (activation-checkpoint wrappers, ``torch.compile``'s ``_orig_mod``)."""
    clean = {}
    for name, value in named_params.items():
        for infix in _WRAPPER_INFIXES:
            name = name.replace(infix, "")
        if name.startswith("_orig_mod."):
            name = name[len("_orig_mod.") :]
        clean[name] = value
    return clean


def get_parameters_from_modules(modules, model, device) -> set:
    """The parameters of ``modules`` — a list of modules, or a regular expression matched against the module names of ``model`` (the
    matches are moved to ``device`` first): what a sharding engine must leave whole when they are "ignored modules"."""
    import re

    if modules is None:
        return set()
    if isinstance(modules, str):
        pattern = re.compile(modules)
        modules = [module.to(device) for name, module in model.named_modules() if pattern.fullmatch(name)]
    return {parameter for module in modules for parameter in module.parameters()}
