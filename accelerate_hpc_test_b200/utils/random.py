"""Seeding, and making the ranks' random number generators agree.

``set_seed`` seeds every generator a training script can draw from.  ``synchronize_rng_states`` is what the prepared dataloaders call
at the start of an epoch: rank 0's generator state is broadcast so that a shuffling sampler produces the same permutation on every
rank (each rank then keeps its own slice of it).  Behavioural parity: reference ``src/accelerate/utils/random.py:31-156``; only the
generator kinds that exist on a CUDA build are implemented, the others are ignored.
"""

from __future__ import annotations

import random
from typing import Callable, Iterable, Optional, Union

import numpy as np
import torch

from ._lazy_state import PartialState
from .dataclasses import DistributedType, RNGType


def set_seed(seed: int, device_specific: bool = False, deterministic: bool = False) -> None:
    """Seed ``random``, ``numpy`` and ``torch`` (host and every CUDA device).  ``device_specific`` gives each rank its own stream
    (seed + rank); ``deterministic`` additionally forbids non-deterministic kernels."""
    if device_specific:
        seed += PartialState().process_index
    random.seed(seed)
    np.random.seed(seed % 2**32)  # numpy only takes 32 bits
    torch.manual_seed(seed)  # also seeds the default CUDA generators lazily
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    if deterministic:
        torch.use_deterministic_algorithms(True)


def _accessors(kind: RNGType, generator: Optional[torch.Generator]) -> Optional[tuple[Callable, Callable]]:
    """(read state, write state) of the generator ``kind`` names; None when there is nothing to synchronise on this build."""
    if kind == RNGType.TORCH:
        return torch.get_rng_state, torch.set_rng_state
    if kind == RNGType.CUDA:
        return (torch.cuda.get_rng_state, torch.cuda.set_rng_state) if torch.cuda.is_available() else None
    if kind == RNGType.GENERATOR:
        if generator is None:
            raise ValueError("Need a generator to synchronize its seed.")
        return generator.get_state, generator.set_state
    return None


def synchronize_rng_state(rng_type: Optional[RNGType] = None, generator: Optional[torch.Generator] = None) -> None:
    """Overwrite this rank's generator of kind ``rng_type`` with rank 0's (a broadcast of a few KB; nothing in one process)."""
    if rng_type is None:
        return
    access = _accessors(RNGType(rng_type), generator)
    if access is None:
        return
    read, write = access
    state = PartialState()
    snapshot = read()
    if state.num_processes > 1 and state.distributed_type != DistributedType.NO:
        on_wire = snapshot.to(state.device)  # NCCL moves device tensors; gloo takes the CPU tensor as is
        torch.distributed.broadcast(on_wire, src=0)
        snapshot = on_wire.cpu()
    write(snapshot)


def synchronize_rng_states(rng_types: Iterable[Union[str, RNGType]], generator: Optional[torch.Generator] = None) -> None:
    for kind in rng_types:
        synchronize_rng_state(RNGType(kind), generator=generator)
