"""``AcceleratedOptimizer`` — what ``Accelerator.prepare`` returns for an optimizer.

It lets a training loop call ``zero_grad()`` / ``step()`` on every micro-batch and does the right thing:

* while gradients are being accumulated (``GradientState.sync_gradients`` is False) both calls do nothing;
* with an fp16 ``GradScaler`` the step goes through ``scaler.step`` / ``scaler.update`` and a step the scaler skipped (inf / nan
  gradients) is reported by ``step_was_skipped`` — the scheduler wrapper uses that to hold the learning-rate schedule;
* optimizer state created before ``prepare`` follows the parameters to the accelerator device;
* a native engine that owns the parameter update (sharded AdamW fused with the reduce-scatter, see ``parallel/fsdp.py``) can
  be attached: ``step`` / ``zero_grad`` / ``state_dict`` / ``load_state_dict`` are then routed to it.

Behavioural parity: reference ``src/accelerate/optimizer.py:38-213``.
"""

from __future__ import annotations

import inspect

import torch

from .state import AcceleratorState, GradientState
from .utils.diagnostics import fault_injector
from .utils.operations import honor_type


def move_to_device(state, device):
    """Recursively move every tensor of a (nested) optimizer state to ``device``; containers keep their type."""
    if isinstance(state, torch.Tensor):
        return state.to(device)
    if isinstance(state, dict):
        return type(state)((key, move_to_device(value, device)) for key, value in state.items())
    if isinstance(state, (list, tuple)):
        return honor_type(state, (move_to_device(item, device) for item in state))
    return state


def patch_optimizer_step(accelerated_optimizer: "AcceleratedOptimizer", method):
    """Reference-compatible helper: a ``step`` that records on ``accelerated_optimizer`` that it ran."""

    def stepped(*args, **kwargs):
        accelerated_optimizer._accelerate_step_called = True
        return method(*args, **kwargs)

    return stepped


class AcceleratedOptimizer(torch.optim.Optimizer):
    """Args: ``optimizer`` (any torch optimizer), ``device_placement`` (move existing state to the device), ``scaler``."""

    _TRANSIENT = ("_accelerate_step_called", "_optimizer_original_step_method", "_optimizer_patched_step_method", "_engine")

    def __init__(self, optimizer, device_placement: bool = True, scaler=None):
        self.optimizer = optimizer
        self.scaler = scaler
        self.accelerator_state = AcceleratorState()
        self.gradient_state = GradientState()
        self.device_placement = device_placement
        self._is_overflow = False
        self._engine = None
        self._steps_taken = 0
        self._arm_scaler_detection()
        if device_placement:
            on_device = move_to_device(self.optimizer.state_dict(), self.accelerator_state.device)
            self.optimizer.load_state_dict(on_device)

    def _arm_scaler_detection(self) -> None:
        self._accelerate_step_called = False
        self._optimizer_original_step_method = None
        self._optimizer_patched_step_method = None
        if self.scaler is not None:
            self._optimizer_original_step_method = self.optimizer.step
            self._optimizer_patched_step_method = patch_optimizer_step(self, self.optimizer.step)

    # ---- transparent views of the wrapped optimizer ------------------------------------------------
    state = property(lambda self: self.optimizer.state, lambda self, value: setattr(self.optimizer, "state", value))
    param_groups = property(lambda self: self.optimizer.param_groups, lambda self, value: setattr(self.optimizer, "param_groups", value))
    defaults = property(lambda self: self.optimizer.defaults, lambda self, value: setattr(self.optimizer, "defaults", value))

    def add_param_group(self, param_group):
        self.optimizer.add_param_group(param_group)

    def train(self):
        """Schedule-free optimizers have train / eval modes; for the others this is a no-op."""
        mode = getattr(self.optimizer, "train", None)
        if callable(mode):
            mode()

    def eval(self):
        mode = getattr(self.optimizer, "eval", None)
        if callable(mode):
            mode()

    # ---- engine hand-off ------------------------------------------------------------------------------
    def attach_engine(self, engine) -> None:
        """Let a native engine perform the update (e.g. reduce-scatter fused with AdamW on sharded state)."""
        self._engine = engine

    def _engine_hook(self, name: str):
        return getattr(self._engine, name, None) if self._engine is not None else None

    def state_dict(self):
        hook = self._engine_hook("optimizer_state_dict")
        return hook(self.optimizer) if hook else self.optimizer.state_dict()

    def load_state_dict(self, state_dict):
        hook = self._engine_hook("load_optimizer_state_dict")
        return hook(self.optimizer, state_dict) if hook else self.optimizer.load_state_dict(state_dict)

    # ---- the two calls of the training loop ---------------------------------------------------------------
    def zero_grad(self, set_to_none=None):
        if not self.gradient_state.sync_gradients:
            return
        hook = self._engine_hook("zero_grad")
        if hook:
            hook()
        if "set_to_none" in inspect.signature(self.optimizer.zero_grad).parameters:
            self.optimizer.zero_grad(set_to_none=True if set_to_none is None else set_to_none)
        elif set_to_none is not None:
            raise ValueError("`set_to_none` for Optimizer.zero_grad` is not supported by this optimizer.")
        else:
            self.optimizer.zero_grad()

    def step(self, closure=None):
        if not self.gradient_state.sync_gradients:
            return
        self._steps_taken += 1
        injector = fault_injector()  # test hook: ACCELERATE_B200_FAULT_INJECT kills a chosen rank at a chosen step
        if injector.armed:
            injector.maybe_fail(self._steps_taken)
        hook = self._engine_hook("optimizer_step")
        if hook:
            self._is_overflow = bool(hook(self.optimizer, self.scaler))
        elif self.scaler is None:
            self.optimizer.step(closure)
        else:
            self._scaled_step(closure)

    def _scaled_step(self, closure) -> None:
        # scaler.step() calls optimizer.step() only when every gradient is finite: watch whether it did
        self.optimizer.step = self._optimizer_patched_step_method
        try:
            self.scaler.step(self.optimizer, closure)
            self.scaler.update()
        finally:
            self.optimizer.step = self._optimizer_original_step_method
        self._is_overflow = not self._accelerate_step_called
        self._accelerate_step_called = False

    @property
    def step_was_skipped(self) -> bool:
        """True when the last ``step`` left the parameters untouched because the gradient scaler found inf / nan."""
        return self._is_overflow

    def _switch_parameters(self, parameters_map) -> None:
        for group in self.optimizer.param_groups:
            group["params"] = [parameters_map.get(p, p) for p in group["params"]]

    # ---- pickling (save_state stores the wrapped object) -----------------------------------------------------
    def __getstate__(self):
        return {key: value for key, value in self.__dict__.items() if key not in self._TRANSIENT}

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._engine = None
        self._arm_scaler_detection()
