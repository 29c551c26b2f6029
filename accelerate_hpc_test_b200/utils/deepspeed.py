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
    """The plugin in force: the only one, or — when several were registered by name — the one marked ``selected``."""
    if state.distributed_type.value != "DEEPSPEED":
        raise ValueError("Couldn't retrieve the active `DeepSpeedPlugin` as none were enabled. Please make sure that either `Accelerator` "
                         "is configured for `deepspeed` or make sure that the desired `DeepSpeedPlugin` has been enabled "
                         "(`AcceleratorState().select_deepspeed_plugin(name)`) before calling this function.")
    plugins = state.deepspeed_plugins
    if isinstance(plugins, dict):
        return next(plugin for plugin in plugins.values() if plugin.selected)
    return plugins


_MISSING = object()


class HfDeepSpeedConfig:
    """A DeepSpeed configuration with dotted-path access (``cfg.get_value("zero_optimization.stage")``).  Accepts a dict (copied), the
    path of a JSON file, or the JSON text base64-encoded (how launchers pass it through an environment variable)."""

    def __init__(self, config_file_or_dict):
        self.config = self._read(config_file_or_dict)
        self.set_stage_and_offload()

    @staticmethod
    def _read(source) -> dict:
        if isinstance(source, dict):
            return deepcopy(source)
        if os.path.exists(source):
            with open(source, encoding="utf-8") as handle:
                return json.load(handle)
        try:
            import base64

            return json.loads(base64.urlsafe_b64decode(source).decode("utf-8"))
        except (UnicodeDecodeError, AttributeError, ValueError):
            raise ValueError("Expected a string path to an existing deepspeed config, or a dictionary, or a base64 encoded string. "
                             f"Received: {source}") from None

    def _walk(self, dotted: str):
        """``(the dict that holds the last key, the last key)`` — the dict is None when the path breaks off earlier."""
        *parents, leaf = dotted.split(".")
        node = self.config
        for key in parents:
            node = node.get(key) if isinstance(node, dict) else None
            if node is None:
                break
        return node, leaf

    find_config_node = _walk  # reference name

    def get_value(self, ds_key_long, default=None):
        holder, leaf = self._walk(ds_key_long)
        return default if holder is None else holder.get(leaf, default)

    def del_config_sub_tree(self, ds_key_long, must_exist=False):
        """Remove the entry at ``ds_key_long`` (and everything below it)."""
        holder, leaf = self._walk(ds_key_long)
        if holder is None or holder.get(leaf) is None:
            if must_exist:
                raise ValueError(f"Can't find {ds_key_long} entry in the config: {self.config}")
            return
        del holder[leaf]

    def is_true(self, ds_key_long) -> bool:
        """Present and truthy (an absent key is neither true nor false)."""
        value = self.get_value(ds_key_long, _MISSING)
        return value is not _MISSING and value is not None and bool(value)

    def is_false(self, ds_key_long) -> bool:
        """Present and falsy."""
        value = self.get_value(ds_key_long, _MISSING)
        return value is not _MISSING and value is not None and not bool(value)

    def set_stage_and_offload(self) -> None:
        self._stage = self.get_value("zero_optimization.stage", -1)
        targets = {self.get_value(f"zero_optimization.{what}.device") for what in ("offload_optimizer", "offload_param")}
        self._offload = self._stage in (2, 3) and bool(targets & {"cpu", "nvme"})

    def is_zero2(self) -> bool:
        return self._stage == 2

    def is_zero3(self) -> bool:
        return self._stage == 3

    def is_offload(self) -> bool:
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
    """``backward`` through an engine-like object.  An object with its own ``backward`` / ``step`` (DeepSpeed's contract: the engine
    scales, accumulates and, at an accumulation boundary, steps) is driven that way; a native engine or plain module has neither,
    and the loss is simply differentiated — the optimizer step stays with the prepared optimizer."""

    def __init__(self, engine):
        self.engine = engine

    def backward(self, loss, sync_gradients=True, **kwargs):
        engine_backward = getattr(self.engine, "backward", None)
        if not callable(engine_backward):
            loss.backward(**kwargs)
            return
        boundary = getattr(self.engine, "set_gradient_accumulation_boundary", None)
        if callable(boundary):
            boundary(is_boundary=sync_gradients)
        engine_backward(loss, **kwargs)
        if sync_gradients and callable(getattr(self.engine, "step", None)):
            self.engine.step()

    def get_global_grad_norm(self):
        """The gradient norm the engine computed at its last step (a float), or None when it has none to report."""
        report = getattr(self.engine, "get_global_grad_norm", None)
        norm = report() if callable(report) else getattr(self.engine, "_last_grad_norm", None)
        return norm.item() if hasattr(norm, "item") else norm


class DeepSpeedOptimizerWrapper:
    """Under DeepSpeed the engine owns the update and this wrapper's ``step`` / ``zero_grad`` do nothing; on the native engines the
    optimizer is real, so they call through.  ``step_was_skipped`` reports an fp16 overflow either way."""

    def __init__(self, optimizer):
        self.optimizer = optimizer

    def zero_grad(self, set_to_none=None):
        self.optimizer.zero_grad()

    def step(self):
        self.optimizer.step()

    @property
    def step_was_skipped(self):
        return bool(getattr(self.optimizer, "step_was_skipped", getattr(self.optimizer, "overflow", False)))


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
