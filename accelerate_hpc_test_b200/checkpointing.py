"""``save_state`` / ``load_state`` back-end: the on-disk checkpoint layout.

File names and dict schemas are the reference's (src/accelerate/checkpointing.py:62-331, table in
SURVEY.md §5.4): ``model.safetensors`` (``model_1.safetensors`` ...) or ``pytorch_model.bin``,
``optimizer.bin``, ``scheduler.bin``, ``sampler.bin``, ``dl_state_dict.bin``, ``scaler.pt``,
``random_states_{rank}.pkl`` written by EVERY rank, ``custom_checkpoint_{i}.pkl``.
"""

from __future__ import annotations

import random
from pathlib import Path

import numpy as np
import torch
from safetensors.torch import load_model

from .logging import get_logger
from .state import PartialState
from .utils.constants import MODEL_NAME, OPTIMIZER_NAME, RNG_STATE_NAME, SAFE_MODEL_NAME, SAMPLER_NAME, SCALER_NAME, SCHEDULER_NAME
from .utils.other import get_pretty_name, load, save

logger = get_logger(__name__)


def _suffix(i: int) -> str:
    return "" if i == 0 else f"_{i}"


def save_accelerator_state(output_dir: str, model_states: list, optimizers: list, schedulers: list, dataloaders: list,
                           process_index: int, step: int, scaler=None, save_on_each_node: bool = False,
                           safe_serialization: bool = True):
    """Write model / optimizer / scheduler / sampler / dataloader / scaler / RNG state under ``output_dir``."""
    output_dir = Path(output_dir)
    for i, state in enumerate(model_states):
        name = (SAFE_MODEL_NAME if safe_serialization else MODEL_NAME) + _suffix(i) + (".safetensors" if safe_serialization else ".bin")
        save(state, output_dir.joinpath(name), save_on_each_node=save_on_each_node, safe_serialization=safe_serialization)
        logger.info(f"Model weights saved in {output_dir.joinpath(name)}")
    for i, opt in enumerate(optimizers):
        name = f"{OPTIMIZER_NAME}{_suffix(i)}.bin"
        save(opt.state_dict(), output_dir.joinpath(name), save_on_each_node=save_on_each_node, safe_serialization=False)
        logger.info(f"Optimizer state saved in {output_dir.joinpath(name)}")
    for i, scheduler in enumerate(schedulers):
        name = f"{SCHEDULER_NAME}{_suffix(i)}.bin"
        save(scheduler.state_dict(), output_dir.joinpath(name), save_on_each_node=save_on_each_node, safe_serialization=False)
        logger.info(f"Scheduler state saved in {output_dir.joinpath(name)}")
    for i, dataloader in enumerate(dataloaders):
        from .data_loader import IterableDatasetShard, SeedableRandomSampler

        if isinstance(dataloader.dataset, IterableDatasetShard):
            sampler = None
        else:
            sampler = dataloader.get_sampler() if hasattr(dataloader, "get_sampler") else None
        if isinstance(sampler, SeedableRandomSampler):
            name = f"{SAMPLER_NAME}{_suffix(i)}.bin"
            save(sampler, output_dir.joinpath(name), save_on_each_node=save_on_each_node, safe_serialization=False)
        if getattr(dataloader, "use_stateful_dataloader", False):
            name = "dl_state_dict.bin" if i == 0 else f"dl_state_dict_{i}.bin"
            save(dataloader.state_dict(), output_dir.joinpath(name), save_on_each_node=save_on_each_node, safe_serialization=False)
        logger.info(f"Sampler state for dataloader {i} saved in {output_dir}")
    if scaler is not None:
        save(scaler.state_dict(), output_dir.joinpath(SCALER_NAME), save_on_each_node=save_on_each_node, safe_serialization=False)
        logger.info(f"Gradient scaler state saved in {output_dir.joinpath(SCALER_NAME)}")
    # RNG: every rank writes its own file
    states = {"step": step, "random_state": random.getstate(), "numpy_random_seed": np.random.get_state(),
              "torch_manual_seed": torch.get_rng_state()}
    if torch.cuda.is_available():
        states["torch_cuda_manual_seed"] = torch.cuda.get_rng_state_all()
    output_states_file = output_dir.joinpath(f"{RNG_STATE_NAME}_{process_index}.pkl")
    torch.save(states, output_states_file)
    logger.info(f"Random states saved in {output_states_file}")
    return output_dir


def load_accelerator_state(input_dir, models, optimizers, schedulers, dataloaders, process_index, scaler=None, map_location=None,
                           load_kwargs=None, **load_model_func_kwargs):
    """Inverse of :func:`save_accelerator_state`. Returns ``{"step": int}`` read from the RNG file (if present)."""
    override_attributes = {}
    if map_location not in (None, "cpu", "on_device"):
        raise TypeError("Unsupported optimizer map location passed, please choose one of `None`, `'cpu'`, or `'on_device'`")
    if map_location is None:
        map_location = "cpu"
    elif map_location == "on_device":
        map_location = PartialState().device
    load_kwargs = dict(load_kwargs or {})
    input_dir = Path(input_dir)
    for i, model in enumerate(models):
        safe_path = input_dir.joinpath(f"{SAFE_MODEL_NAME}{_suffix(i)}.safetensors")
        if safe_path.exists():
            load_model(model, str(safe_path), device=str(map_location), **load_model_func_kwargs)
        else:
            bin_path = input_dir.joinpath(f"{MODEL_NAME}{_suffix(i)}.bin")
            state_dict = load(bin_path, map_location=map_location)
            model.load_state_dict(state_dict, **load_model_func_kwargs)
    logger.info("All model weights loaded successfully")
    for i, opt in enumerate(optimizers):
        path = input_dir.joinpath(f"{OPTIMIZER_NAME}{_suffix(i)}.bin")
        opt.load_state_dict(load(path, map_location=map_location, **load_kwargs))
    logger.info("All optimizer states loaded successfully")
    for i, scheduler in enumerate(schedulers):
        path = input_dir.joinpath(f"{SCHEDULER_NAME}{_suffix(i)}.bin")
        scheduler.load_state_dict(load(path, **load_kwargs))
    logger.info("All scheduler states loaded successfully")
    for i, dataloader in enumerate(dataloaders):
        from .data_loader import IterableDatasetShard, SeedableRandomSampler

        path = input_dir.joinpath(f"{SAMPLER_NAME}{_suffix(i)}.bin")
        if isinstance(dataloader.dataset, IterableDatasetShard):
            sampler = None
        else:
            sampler = dataloader.get_sampler() if hasattr(dataloader, "get_sampler") else None
        if isinstance(sampler, SeedableRandomSampler) and path.exists():
            sampler = dataloader.set_sampler(load(path, weights_only=False))
        if getattr(dataloader, "use_stateful_dataloader", False):
            dl_path = input_dir.joinpath("dl_state_dict.bin" if i == 0 else f"dl_state_dict_{i}.bin")
            if dl_path.exists():
                dataloader.load_state_dict(load(dl_path, **load_kwargs))
    logger.info("All dataloader sampler states loaded successfully")
    if scaler is not None:
        scaler.load_state_dict(load(input_dir.joinpath(SCALER_NAME)))
        logger.info("GradScaler state loaded successfully")
    try:
        states = load(input_dir.joinpath(f"{RNG_STATE_NAME}_{process_index}.pkl"), weights_only=False)
        if "step" in states:
            override_attributes["step"] = states["step"]
        random.setstate(states["random_state"])
        np.random.set_state(states["numpy_random_seed"])
        torch.set_rng_state(states["torch_manual_seed"])
        if torch.cuda.is_available() and "torch_cuda_manual_seed" in states:
            torch.cuda.set_rng_state_all(states["torch_cuda_manual_seed"])
        logger.info("All random states loaded successfully")
    except Exception:
        logger.info("Could not load random states")
    return override_attributes


def save_custom_state(obj, path, index: int = 0, save_on_each_node: bool = False):
    """``custom_checkpoint_{index}.pkl`` holding ``obj.state_dict()``."""
    save_location = Path(path) / f"custom_checkpoint_{index}.pkl"
    logger.info(f"Saving the state of {get_pretty_name(obj)} to {save_location}")
    save(obj.state_dict(), save_location, save_on_each_node=save_on_each_node)


def load_custom_state(obj, path, index: int = 0):
    load_location = f"{path}/custom_checkpoint_{index}.pkl"
    logger.info(f"Loading the state of {get_pretty_name(obj)} from {load_location}")
    obj.load_state_dict(load(load_location, map_location="cpu", weights_only=False))
