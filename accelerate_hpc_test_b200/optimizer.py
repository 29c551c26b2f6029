"""``AcceleratedOptimizer``: the optimizer wrapper ``Accelerator.prepare`` hands back.

Behaviour parity with reference src/accelerate/optimizer.py:38-213 — ``step`` / ``zero_grad`` do
nothing while gradients are being accumulated, a GradScaler-skipped step is detected and exposed as
``step_was_skipped``, optimizer state follows the parameters to the device.  Added for this build:
the native engines (sharded / replicated data parallel) can take over the update entirely
(``engine_step``), which is how reduce-scatter + AdamW run as one kernel.
"""

from __future__ import annotations

import inspect

import torch

from .state import AcceleratorState, GradientState
from .utils.operations import honor_type


def move_to_device(state, device):
    if isinstance(state, (list, tuple)):
        return honor_type(state, (move_to_device(t, device) for t in state))
    if isinstance(state, dict):
        return type(state)({k: move_to_device(v, device) for k, v in state.items()})
    if isinstance(state, torch.Tensor):
        return state.to(device)
    return state


from .utils.diagnostics import fault_injector


class AcceleratedOptimizer(torch.optim.Optimizer):
    """Wraps a torch optimizer so user code can call ``step`` / ``zero_grad`` on every micro-batch.

    Args:
        optimizer: the optimizer to wrap.
        device_placement: move existing optimizer state to the accelerator device.
        scaler: an (fp16) ``GradScaler`` to step through, if any.
    """

    def __init__(self, optimizer, device_placement: bool = True, scaler=None):
        self.optimizer = optimizer
        self.scaler = scaler
        self.accelerator_state = AcceleratorState()
        self.gradient_state = GradientState()
        self.device_placement = device_placement
        self._is_overflow = False
        self._accelerate_step_called = False
        self._optimizer_original_step_method = None
        self._optimizer_patched_step_method = None
        self._engine = None  # a native engine that owns the parameter update (see parallel/fsdp.py)
        if self.scaler is not None:
            self._optimizer_original_step_method = self.optimizer.step
            self._optimizer_patched_step_method = patch_optimizer_step(self, self.optimizer.step)
        if device_placement:
            state_dict = self.optimizer.state_dict()
            state_dict = move_to_device(state_dict, self.accelerator_state.device)
            self.optimizer.load_state_dict(state_dict)

    @property
    def state(self):
        return self.optimizer.state

    @state.setter
    def state(self, state):
        self.optimizer.state = state

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @param_groups.setter
    def param_groups(self, param_groups):
        self.optimizer.param_groups = param_groups

    @property
    def defaults(self):
        return self.optimizer.defaults

    @defaults.setter
    def defaults(self, defaults):
        self.optimizer.defaults = defaults

    def add_param_group(self, param_group):
        self.optimizer.add_param_group(param_group)

    def load_state_dict(self, state_dict):
        if self._engine is not None and hasattr(self._engine, "load_optimizer_state_dict"):
            return self._engine.load_optimizer_state_dict(self.optimizer, state_dict)
        self.optimizer.load_state_dict(state_dict)

    def state_dict(self):
        if self._engine is not None and hasattr(self._engine, "optimizer_state_dict"):
            return self._engine.optimizer_state_dict(self.optimizer)
        return self.optimizer.state_dict()

    def attach_engine(self, engine):
        """Let a native engine perform the update (fused reduce-scatter + AdamW on sharded state)."""
        self._engine = engine

    def zero_grad(self, set_to_none=None):
        if not self.gradient_state.sync_gradients:
            return
        if self._engine is not None and hasattr(self._engine, "zero_grad"):
            self._engine.zero_grad()
        accept_arg = "set_to_none" in inspect.signature(self.optimizer.zero_grad).parameters
        if accept_arg:
            if set_to_none is None:
                set_to_none = True
            self.optimizer.zero_grad(set_to_none=set_to_none)
        else:
            if set_to_none is not None:
                raise ValueError("`set_to_none` for Optimizer.zero_grad` is not supported by this optimizer.")
            self.optimizer.zero_grad()

    def train(self):
        """Switch a schedule-free style optimizer to train mode (no-op otherwise)."""
        if hasattr(self.optimizer, "train") and callable(self.optimizer.train):
            self.optimizer.train()

    def eval(self):
        if hasattr(self.optimizer, "eval") and callable(self.optimizer.eval):
            self.optimizer.eval()

    def step(self, closure=None):
        if not self.gradient_state.sync_gradients:
            return
        # test hook: ACCELERATE_B200_FAULT_INJECT makes a chosen rank die at a chosen optimizer step (utils/diagnostics.py)
        self._steps_taken = getattr(self, "_steps_taken", 0) + 1
        inj = fault_injector()
        if inj.armed:
            inj.maybe_fail(self._steps_taken)
        if self._engine is not None and hasattr(self._engine, "optimizer_step"):
            self._is_overflow = bool(self._engine.optimizer_step(self.optimizer, self.scaler))
            return
        if self.scaler is not None:
            self.optimizer.step = self._optimizer_patched_step_method
            self.scaler.step(self.optimizer, closure)
            self.scaler.update()
            # the scaler skips optimizer.step when it found inf/nan gradients
            self._is_overflow = not self._accelerate_step_called
            self.optimizer.step = self._optimizer_original_step_method
            self._accelerate_step_called = False
        else:
            self.optimizer.step(closure)

    def _switch_parameters(self, parameters_map):
        for param_group in self.optimizer.param_groups:
            param_group["params"] = [parameters_map.get(p, p) for p in param_group["params"]]

    @property
    def step_was_skipped(self):
        """Whether the last ``step`` was skipped by the gradient scaler."""
        return self._is_overflow

    def __getstate__(self):
        _ignored = ["_accelerate_step_called", "_optimizer_original_step_method", "_optimizer_patched_step_method", "_engine"]
        return {k: v for k, v in self.__dict__.items() if k not in _ignored}

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._engine = None
        if self.scaler is not None:
            self._accelerate_step_called = False
            self._optimizer_original_step_method = self.optimizer.step
            self._optimizer_patched_step_method = patch_optimizer_step(self, self.optimizer.step)


def patch_optimizer_step(accelerated_optimizer: AcceleratedOptimizer, method):
    def patched_step(*args, **kwargs):
        accelerated_optimizer._accelerate_step_called = True
        return method(*args, **kwargs)

    return patched_step
