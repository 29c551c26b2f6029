"""DeepSpeed compatibility objects (reference src/accelerate/utils/deepspeed.py:119-385).

DeepSpeed is not a dependency of this build: a DeepSpeed-style configuration is executed by the native
engines (ZeRO-3 == full shard, ZeRO-2 == shard gradients + optimizer, Ulysses == :mod:`..parallel.ulysses`).
These classes keep user scripts written against the reference importable: ``DummyOptim`` /
``DummyScheduler`` placeholders are turned into real AdamW / LambdaLR objects at ``prepare`` time.
"""

from __future__ import annotations

import json
import os
from copy import deepcopy

import torch

from .dataclasses import DeepSpeedPlugin


def get_active_deepspeed_plugin(state):
    if state.distributed_type.value != "DEEPSPEED":
        raise ValueError("Couldn't retrieve the active `DeepSpeedPlugin` as none were enabled. Please make sure that either `Accelerator` is configured for `deepspeed` or make sure that the desired `DeepSpeedPlugin` has been enabled (`AcceleratorState().select_deepspeed_plugin(name)`) before calling this function.")
    if not isinstance(state.deepspeed_plugins, dict):
        return state.deepspeed_plugins
    return next(plugin for plugin in state.deepspeed_plugins.values() if plugin.selected)


class HfDeepSpeedConfig:
    """Dictionary-path access to a DeepSpeed config (dict, JSON path or base64 JSON)."""

    def __init__(self, config_file_or_dict):
        if isinstance(config_file_or_dict, dict):
            config = deepcopy(config_file_or_dict)
        elif os.path.exists(config_file_or_dict):
            with open(config_file_or_dict, encoding="utf-8") as f:
                config = json.load(f)
        else:
            try:
                import base64

                config = json.loads(base64.urlsafe_b64decode(config_file_or_dict).decode("utf-8"))
            except (UnicodeDecodeError, AttributeError, ValueError):
                raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary, or a base64 encoded string. Received: {config_file_or_dict}")
        self.config = config
        self.set_stage_and_offload()

    def set_stage_and_offload(self):
        self._stage = self.get_value("zero_optimization.stage", -1)
        self._offload = False
        if self.is_zero2() or self.is_zero3():
            offload_devices_valid = {"cpu", "nvme"}
            offload_devices = {self.get_value("zero_optimization.offload_optimizer.device"), self.get_value("zero_optimization.offload_param.device")}
            if len(offload_devices & offload_devices_valid) > 0:
                self._offload = True

    def find_config_node(self, ds_key_long):
        config = self.config
        nodes = ds_key_long.split(".")
        ds_key = nodes.pop()
        for node in nodes:
            config = config.get(node)
            if config is None:
                return None, ds_key
        return config, ds_key

    def get_value(self, ds_key_long, default=None):
        config, ds_key = self.find_config_node(ds_key_long)
        if config is None:
            return default
        return config.get(ds_key, default)

    def del_config_sub_tree(self, ds_key_long, must_exist=False):
        config = self.config
        nodes = ds_key_long.split(".")
        for node in nodes:
            parent_config = config
            config = config.get(node)
            if config is None:
                if must_exist:
                    raise ValueError(f"Can't find {ds_key_long} entry in the config: {self.config}")
                return
        if parent_config is not None:
            parent_config.pop(node)

    def is_true(self, ds_key_long):
        value = self.get_value(ds_key_long)
        return False if value is None else bool(value)

    def is_false(self, ds_key_long):
        value = self.get_value(ds_key_long)
        return False if value is None else not bool(value)

    def is_zero2(self):
        return self._stage == 2

    def is_zero3(self):
        return self._stage == 3

    def is_offload(self):
        return self._offload


class DummyOptim:
    """Placeholder optimizer for scripts whose optimizer is described in the DeepSpeed config."""

    def __init__(self, params, lr=0.001, weight_decay=0, **kwargs):
        self.params = params
        self.lr = lr
        self.weight_decay = weight_decay
        self.kwargs = kwargs

    def materialize(self):
        from ..optim import FusedAdamW

        return FusedAdamW(self.params, lr=self.lr, weight_decay=self.weight_decay,
                          **{k: v for k, v in self.kwargs.items() if k in ("betas", "eps")})


class DummyScheduler:
    """Placeholder scheduler (``total_num_steps`` / ``warmup_num_steps`` or an ``lr_scheduler_callable``)."""

    def __init__(self, optimizer, total_num_steps=None, warmup_num_steps=0, lr_scheduler_callable=None, **kwargs):
        self.optimizer = optimizer
        self.total_num_steps = total_num_steps
        self.warmup_num_steps = warmup_num_steps
        self.lr_scheduler_callable = lr_scheduler_callable
        self.kwargs = kwargs

    def materialize(self, optimizer):
        if self.lr_scheduler_callable is not None:
            return self.lr_scheduler_callable(optimizer)
        total, warm = self.total_num_steps, self.warmup_num_steps

        def lr_lambda(step):
            if warm and step < warm:
                return float(step) / float(max(1, warm))
            if total:
                return max(0.0, float(total - step) / float(max(1, total - warm)))
            return 1.0

        return torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda)


class DeepSpeedEngineWrapper:
    """Kept for API parity: ``backward`` on the wrapped engine-like object."""

    def __init__(self, engine):
        self.engine = engine

    def backward(self, loss, sync_gradients=True, **kwargs):
        loss.backward(**kwargs)

    def get_global_grad_norm(self):
        return getattr(self.engine, "_last_grad_norm", None)


class DeepSpeedOptimizerWrapper:
    def __init__(self, optimizer):
        self.optimizer = optimizer

    def zero_grad(self, set_to_none=None):
        self.optimizer.zero_grad()

    def step(self):
        self.optimizer.step()

    @property
    def step_was_skipped(self):
        return getattr(self.optimizer, "step_was_skipped", False)


class DeepSpeedSchedulerWrapper:
    def __init__(self, scheduler, optimizers):
        self.scheduler = scheduler
        self.optimizers = optimizers

    def step(self):
        self.scheduler.step()


def map_pytorch_optim_to_deepspeed(optimizer):
    """The reference maps Adam-family optimizers to DeepSpeed's fused versions; here the fused AdamW is ours."""
    from ..optim import FusedAdamW

    defaults = {k: v for k, v in optimizer.defaults.items() if k in ["lr", "weight_decay", "betas", "eps"]}
    return FusedAdamW(optimizer.param_groups, **defaults)
