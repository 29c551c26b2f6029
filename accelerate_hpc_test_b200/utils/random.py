"""Seeding and cross-rank RNG synchronisation (reference src/accelerate/utils/random.py:31-156)."""

from __future__ import annotations

import random
from typing import Optional, Union

import numpy as np
import torch

from ._lazy_state import PartialState
from .dataclasses import DistributedType, RNGType


def set_seed(seed: int, device_specific: bool = False, deterministic: bool = False):
    """Seed python / numpy / torch (and CUDA). ``device_specific`` offsets the seed by the rank."""
    if device_specific:
        seed += PartialState().process_index
    random.seed(seed)
    np.random.seed(seed % (2**32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    if deterministic:
        torch.use_deterministic_algorithms(True)


def synchronize_rng_state(rng_type: Optional[RNGType] = None, generator: Optional[torch.Generator] = None):
    """Make every rank's RNG of the given kind equal to rank 0's (one small broadcast)."""
    state = PartialState()
    rng_type = RNGType(rng_type) if rng_type is not None else None
    if rng_type == RNGType.TORCH:
        rng_state = torch.get_rng_state()
    elif rng_type == RNGType.CUDA:
        if not torch.cuda.is_available():
            return
        rng_state = torch.cuda.get_rng_state()
    elif rng_type == RNGType.GENERATOR:
        if generator is None:
            raise ValueError("Need a generator to synchronize its seed.")
        rng_state = generator.get_state()
    else:
        return  # RNG kinds of other accelerators do not exist on this build
    if state.distributed_type != DistributedType.NO and state.num_processes > 1:
        rng_state = rng_state.to(state.device)
        torch.distributed.broadcast(rng_state, 0)
        rng_state = rng_state.cpu()
    if rng_type == RNGType.TORCH:
        torch.set_rng_state(rng_state)
    elif rng_type == RNGType.CUDA:
        torch.cuda.set_rng_state(rng_state)
    elif rng_type == RNGType.GENERATOR:
        generator.set_state(rng_state)


def synchronize_rng_states(rng_types: list[Union[str, RNGType]], generator: Optional[torch.Generator] = None):
    for rng_type in rng_types:
        synchronize_rng_state(RNGType(rng_type), generator=generator)
