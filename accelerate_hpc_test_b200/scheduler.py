"""``AcceleratedScheduler`` — keeps a learning-rate schedule in lock-step with *real* optimizer updates.

Behaviour (parity target: reference ``src/accelerate/scheduler.py:54-82``):

* micro-steps of a gradient-accumulation window do not advance the schedule (with
  ``GradientState.adjust_scheduler`` only the wrapped scheduler's internal call counter is bumped, so
  its "step() before optimizer.step()" bookkeeping stays quiet);
* an update that a GradScaler skipped (inf/nan gradients) does not advance it either;
* schedules are written for the single-process number of updates, and every rank consumes
  ``1/num_processes`` of the data, so one call advances ``num_processes`` ticks — unless the dataloader
  splits one global batch across ranks (``split_batches=True``), where one call is one tick;
* schedulers that know their horizon (``total_steps``, e.g. OneCycleLR) are never pushed past it.
"""

from __future__ import annotations

import warnings
from typing import Iterable

from .state import AcceleratorState, GradientState

# torch warns about scheduler/optimizer call order whenever a skipped step breaks its assumption
warnings.filterwarnings("ignore", category=UserWarning, module="torch.optim.lr_scheduler")

_FORWARDED = ("get_last_lr", "get_lr", "state_dict", "load_state_dict", "print_lr")


def _as_list(x) -> list:
    if isinstance(x, (list, tuple)):
        return list(x)
    return [x]


class AcceleratedScheduler:
    """Wrapper returned by ``Accelerator.prepare`` for every LR scheduler."""

    def __init__(self, scheduler, optimizers, step_with_optimizer: bool = True, split_batches: bool = False):
        self.scheduler = scheduler
        self.optimizers = _as_list(optimizers)
        self.step_with_optimizer = bool(step_with_optimizer)
        self.split_batches = bool(split_batches)
        self.gradient_state = GradientState()

    # -- how many ticks does one user call stand for --------------------------------------------
    def _ticks(self) -> int:
        if self.split_batches:
            return 1
        return max(1, int(AcceleratorState().num_processes))

    def _room_left(self) -> bool:
        horizon = getattr(self.scheduler, "total_steps", None)
        if horizon is None:
            return True
        return self.scheduler._step_count <= horizon

    def _any_update_skipped(self, optimizers: Iterable) -> bool:
        return any(getattr(o, "step_was_skipped", False) for o in optimizers)

    def step(self, *args, **kwargs):
        inner = self.scheduler
        if not self.step_with_optimizer:
            # the user drives the schedule by hand: pass straight through
            return inner.step(*args, **kwargs)
        gs = self.gradient_state
        if not gs.sync_gradients:
            if gs.adjust_scheduler:
                inner._step_count += 1
            return None
        if self._any_update_skipped(self.optimizers):
            return None
        for _ in range(self._ticks()):
            if not self._room_left():
                break
            inner.step(*args, **kwargs)
        return None

    def __getattr__(self, name):
        # plain forwarding of the read/persist API; anything else is an honest AttributeError
        if name in _FORWARDED:
            return getattr(self.__dict__["scheduler"], name)
        raise AttributeError(f"{type(self).__name__!s} has no attribute {name!r}")
