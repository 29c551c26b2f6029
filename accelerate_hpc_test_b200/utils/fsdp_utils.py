"""Sharded-training utilities with the reference's names (src/accelerate/utils/fsdp_utils.py).

In the reference these wrap torch FSDP / ``torch.distributed.checkpoint``; here they front the native
sharded engine (:class:`..parallel.fsdp.B200ShardedModule`).  The on-disk names are the reference's:
``pytorch_model_fsdp_{i}/`` + ``optimizer_{i}/`` (sharded) and ``pytorch_model_fsdp.bin`` +
``optimizer.bin`` (full) — :103-335; ``merge_fsdp_weights`` :338-418.
"""

from __future__ import annotations

import os
import shutil
from pathlib import Path

import torch

from .constants import FSDP_MODEL_NAME, OPTIMIZER_NAME, SAFE_WEIGHTS_NAME, WEIGHTS_NAME
from .dataclasses import get_module_class_from_name
from .other import get_module_children_bottom_up, save


def enable_fsdp_ram_efficient_loading():
    """Only rank 0 materialises pretrained weights; the engine broadcasts them while sharding."""
    os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] = "True"


def disable_fsdp_ram_efficient_loading():
    os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] = "False"


def _engine(model):
    from ..parallel.fsdp import B200ShardedModule

    m = model
    while m is not None and not isinstance(m, B200ShardedModule):
        m = getattr(m, "module", None) if hasattr(m, "module") and m is not getattr(m, "module", None) else None
    if m is None:
        raise TypeError("expected a model prepared with the sharded engine (B200ShardedModule)")
    return m


def save_fsdp_model(fsdp_plugin, accelerator, model, output_dir, model_index=0, adapter_only=False):
    os.makedirs(output_dir, exist_ok=True)
    _engine(model).save_sharded_state(output_dir, model_index, fsdp_plugin)


def load_fsdp_model(fsdp_plugin, accelerator, model, input_dir, model_index=0, adapter_only=False):
    accelerator.wait_for_everyone()
    _engine(model).load_sharded_state(input_dir, model_index, fsdp_plugin)


def save_fsdp_optimizer(fsdp_plugin, accelerator, optimizer, model, output_dir, optimizer_index=0):
    os.makedirs(output_dir, exist_ok=True)
    _engine(model).save_sharded_optimizer(optimizer, output_dir, optimizer_index, fsdp_plugin)


def load_fsdp_optimizer(fsdp_plugin, accelerator, optimizer, model, input_dir, optimizer_index=0, adapter_only=False):
    accelerator.wait_for_everyone()
    _engine(model).load_sharded_optimizer(optimizer, input_dir, optimizer_index, fsdp_plugin)


def _distributed_checkpoint_to_state_dict(checkpoint_dir: str) -> dict:
    """Full name -> tensor dict of a sharded model directory written by this engine or by torch DCP."""
    from ..parallel.fsdp import merge_sharded_checkpoint

    meta = os.path.join(checkpoint_dir, ".metadata")
    if os.path.isfile(meta):
        try:
            import json

            with open(meta) as f:
                json.load(f)
            return merge_sharded_checkpoint(checkpoint_dir)
        except (UnicodeDecodeError, ValueError):
            pass
    # a torch.distributed.checkpoint directory (e.g. written by the reference): read it without a process group
    import torch.distributed.checkpoint as dist_cp
    import torch.distributed.checkpoint.format_utils as dist_cp_format_utils

    state_dict = {}
    dist_cp_format_utils._load_state_dict(state_dict, storage_reader=dist_cp.FileSystemReader(checkpoint_dir),
                                          planner=dist_cp_format_utils._EmptyStateDictLoadPlanner(), no_dist=True)
    if len(state_dict.keys()) == 1:
        state_dict = state_dict[list(state_dict)[0]]
    return state_dict


def merge_fsdp_weights(checkpoint_dir: str, output_path: str, safe_serialization: bool = True, remove_checkpoint_dir: bool = False):
    """Merge ``pytorch_model_fsdp_0`` (or a directory that contains it) into ``model.safetensors`` / ``pytorch_model.bin``."""
    checkpoint_dir = Path(checkpoint_dir)
    if not checkpoint_dir.is_dir():
        raise ValueError(f"Tried to merge the weights from `{checkpoint_dir}` but it is not a directory.")
    if not (checkpoint_dir / ".metadata").exists():
        candidate = checkpoint_dir / f"{FSDP_MODEL_NAME}_0"
        if candidate.is_dir():
            checkpoint_dir = candidate
        else:
            raise ValueError(f"Tried to load from {checkpoint_dir} but couldn't find a valid metadata file, and no `{FSDP_MODEL_NAME}_0` sub-directory either.")
    from ..state import PartialState

    state = PartialState()
    if state.is_main_process:
        state_dict = _distributed_checkpoint_to_state_dict(str(checkpoint_dir))
        output_path = Path(output_path)
        output_path.mkdir(parents=True, exist_ok=True)
        target = output_path / (SAFE_WEIGHTS_NAME if safe_serialization else WEIGHTS_NAME)
        save(state_dict, target, safe_serialization=safe_serialization)
        if remove_checkpoint_dir:
            shutil.rmtree(checkpoint_dir)
    state.wait_for_everyone()


def ensure_weights_retied(param_init_fn, model: torch.nn.Module, device: torch.device):
    """Wrap ``param_init_fn`` so tied weights are tied again after each module is materialised."""
    _tied_names = getattr(model, "_tied_weights_keys", None)
    if not _tied_names:
        return param_init_fn
    _tied_params = {}
    for name in _tied_names:
        name = name.split(".")
        name, param_name = ".".join(name[:-1]), name[-1]
        mod = model.get_submodule(name)
        param = getattr(mod, param_name)
        _tied_params[id(param)] = None

    def param_init_fn_tied_param(module: torch.nn.Module):
        params_to_tie = {}
        for n, param in module.named_parameters(recurse=False):
            if id(param) in _tied_params:
                params_to_tie.setdefault(id(param), []).append(n)
        module = param_init_fn(module)
        for id_key, _param_names in params_to_tie.items():
            for param_name in _param_names:
                param = _tied_params[id_key]
                if param is None:
                    _tied_params[id_key] = getattr(module, param_name)
                else:
                    setattr(module, param_name, param)
        return module

    return param_init_fn_tied_param


def fsdp2_prepare_model(accelerator, model: torch.nn.Module) -> torch.nn.Module:
    """Shard ``model`` with the native engine (the reference calls torch ``fully_shard`` here, :621-737)."""
    return accelerator._wrap_sharded(model)


def fsdp2_load_full_state_dict(accelerator, model: torch.nn.Module, full_sd: dict, cpu_offload: bool = False):
    """Distribute a full state dict (present on rank 0) over the shards."""
    _engine(model).load_full_state_dict(full_sd)


def fsdp2_switch_optimizer_parameters(optimizer: torch.optim.Optimizer, mapping: dict):
    """Swap optimizer parameters by ``data_ptr`` (kept for API parity; the native engine owns sharded optimizer state)."""
    try:
        for param_group in optimizer.param_groups:
            param_group["params"] = [mapping[p.data_ptr] for p in param_group["params"]]
    except KeyError:
        raise KeyError("A parameter in the optimizer couldn't be switched to its sharded version. This breaks the training. Please raise an issue on GitHub.")


def fsdp2_apply_ac(accelerator, model: torch.nn.Module):
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


def fsdp2_canonicalize_names(named_params: dict) -> dict:
    """Strip wrapper prefixes (``module.``, ``_orig_mod.``, checkpoint wrappers) from parameter names."""
    named_params = {k.replace("._checkpoint_wrapped_module", ""): v for k, v in named_params.items()}
    named_params = {k.replace("_orig_mod.", "") if k.startswith("_orig_mod.") else k: v for k, v in named_params.items()}
    named_params = {k.replace("._orig_mod", ""): v for k, v in named_params.items()}
    return named_params
