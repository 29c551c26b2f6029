"""The ``Accelerator``: one object that adapts a user's model / optimizer / dataloaders / scheduler to
whatever distributed regime the environment describes.

Public surface parity with the reference (src/accelerate/accelerator.py:184-4324; inventory in
SURVEY.md §2.2): same constructor arguments, properties, ``prepare`` / ``backward`` / ``accumulate``
/ ``no_sync`` / ``clip_grad_norm_`` / ``gather*`` / ``reduce`` / ``save_state`` / ``load_state`` /
tracker methods.  What is different is what ``prepare`` hands back on CUDA:

* ``MULTI_GPU``  -> :class:`~.parallel.ddp.B200DataParallel`, a bucketed reducer whose buckets live in
  peer-mapped memory and are reduced by a fused cast+scale+all-reduce kernel overlapped with backward
  (the reference wraps ``torch.nn.parallel.DistributedDataParallel``, :1865-1869);
* ``FSDP``       -> :class:`~.parallel.fsdp.B200ShardedModule`, flat per-unit shards, peer-pull
  all-gather with prefetch, reduce-scatter fused with the sharded AdamW update
  (the reference calls torch's ``fully_shard``, utils/fsdp_utils.py:621-737);
* tp / cp / sp axes of ``ParallelismConfig`` -> the engines in :mod:`.parallel`.

On CPU (gloo) the same classes run with c10d collectives, which is the oracle the tests use.
"""

from __future__ import annotations

import contextlib
import functools
import json
import os
import re
import shutil
import sys
import warnings
from collections import OrderedDict
from contextlib import contextmanager
from functools import partial
from types import MethodType
from typing import Any, Callable, Optional, Union

import torch

from .checkpointing import load_accelerator_state, load_custom_state, save_accelerator_state, save_custom_state
from .data_loader import DataLoaderDispatcher, prepare_data_loader, skip_first_batches
from .logging import get_logger
from .optimizer import AcceleratedOptimizer
from .parallelism_config import ParallelismConfig
from .scheduler import AcceleratedScheduler
from .state import AcceleratorState, GradientState, PartialState
from .utils.constants import (
    PROFILE_PATTERN_NAME,
    SAFE_WEIGHTS_INDEX_NAME,
    SAFE_WEIGHTS_NAME,
    SAFE_WEIGHTS_PATTERN_NAME,
    WEIGHTS_INDEX_NAME,
    WEIGHTS_NAME,
    WEIGHTS_PATTERN_NAME,
)
from .utils.dataclasses import (
    AORecipeKwargs,
    AutocastKwargs,
    DataLoaderConfiguration,
    DeepSpeedPlugin,
    DistributedDataParallelKwargs,
    DistributedType,
    DynamoBackend,
    FP8BackendType,
    FP8RecipeKwargs,
    FullyShardedDataParallelPlugin,
    GradientAccumulationPlugin,
    GradScalerKwargs,
    InitProcessGroupKwargs,
    KwargsHandler,
    MegatronLMPlugin,
    MSAMPRecipeKwargs,
    PrecisionType,
    ProfileKwargs,
    ProjectConfiguration,
    TERecipeKwargs,
    TorchDynamoPlugin,
    TorchTensorParallelPlugin,
)
from .utils.environment import parse_choice_from_env, parse_flag_from_env
from .utils.memory import release_memory
from .utils.operations import convert_outputs_to_fp32, gather, gather_object, pad_across_processes, recursively_apply, reduce
from .utils.other import check_os_kernel, clean_state_dict_for_safetensors, compile_regions, extract_model_from_parallel, is_compiled_module, save, wait_for_everyone

logger = get_logger(__name__)

_split_batches = object()
_dispatch_batches = object()
_even_batches = object()
_use_seedable_sampler = object()


def _remap_after_sharding(before: dict, after: dict, model) -> dict:
    """``id(old parameter) -> new parameter`` by name, for re-pointing optimizers after ``fully_shard`` replaced the parameters.
    A name that is gone afterwards cannot be mapped: with tied embeddings that is the expected casualty (the tie was broken or
    collapsed by the wrapping) and the message says what to do; otherwise it is a bug worth a loud ``KeyError``
    (reference accelerator.py:1690-1744)."""
    missing = [name for name in before if name not in after]
    if missing:
        tied = bool(getattr(getattr(model, "config", None), "tie_word_embeddings", False))
        if tied:
            raise ValueError(
                f"FSDP2 mapping failed: {missing} cannot be found after wrapping. The model uses tied embeddings "
                "(`tie_word_embeddings = True`), which the optimizer re-pointing cannot follow. Create the optimizer after "
                "`prepare(model)`, or load the model with `tie_word_embeddings = False`.")
        raise KeyError(f"Parameters missing after FSDP2 wrapping: {missing}")
    return {id(parameter): after[name] for name, parameter in before.items()}


def _is_module(x):
    return isinstance(x, torch.nn.Module)


def _is_optimizer(x):
    return isinstance(x, torch.optim.Optimizer)


def _is_dataloader(x):
    return isinstance(x, torch.utils.data.DataLoader)


def _is_scheduler(x):
    try:
        from torch.optim.lr_scheduler import LRScheduler
    except ImportError:  # pragma: no cover
        from torch.optim.lr_scheduler import _LRScheduler as LRScheduler
    return isinstance(x, LRScheduler)


_UNSET = object()


def _forwarded(owner: str, attribute: str, settable: bool = False, default=_UNSET):
    """Property that reads (and optionally writes) ``self.<owner>.<attribute>``."""

    def read(self):
        target = getattr(self, owner)
        return getattr(target, attribute) if default is _UNSET else getattr(target, attribute, default)

    def write(self, value):
        setattr(getattr(self, owner), attribute, value)

    return property(read, write if settable else None)


class Accelerator:
    """Creates the distributed state and wraps training objects.

    Args mirror the reference: ``device_placement, split_batches, mixed_precision,
    gradient_accumulation_steps, cpu, dataloader_config, deepspeed_plugin(s), fsdp_plugin,
    torch_tp_plugin (deprecated), megatron_lm_plugin, rng_types, log_with, project_dir, project_config,
    gradient_accumulation_plugin, step_scheduler_with_optimizer, kwargs_handlers, dynamo_backend,
    dynamo_plugin, parallelism_config``.
    """

    def __init__(
        self,
        device_placement: bool = True,
        split_batches: bool = _split_batches,
        mixed_precision: Optional[str] = None,
        gradient_accumulation_steps: int = 1,
        cpu: bool = False,
        dataloader_config: Optional[DataLoaderConfiguration] = None,
        deepspeed_plugin: Union[DeepSpeedPlugin, dict, None] = None,
        fsdp_plugin: Optional[FullyShardedDataParallelPlugin] = None,
        torch_tp_plugin: Optional[TorchTensorParallelPlugin] = None,
        megatron_lm_plugin: Optional[MegatronLMPlugin] = None,
        rng_types: Optional[list] = None,
        log_with=None,
        project_dir: Optional[Union[str, os.PathLike]] = None,
        project_config: Optional[ProjectConfiguration] = None,
        gradient_accumulation_plugin: Optional[GradientAccumulationPlugin] = None,
        step_scheduler_with_optimizer: bool = True,
        kwargs_handlers: Optional[list[KwargsHandler]] = None,
        dynamo_backend=None,
        dynamo_plugin: Optional[TorchDynamoPlugin] = None,
        deepspeed_plugins: Union[DeepSpeedPlugin, dict, None] = None,
        parallelism_config: Optional[ParallelismConfig] = None,
    ):
        self.trackers = []
        if project_config is not None:
            self.project_configuration = project_config
        else:
            self.project_configuration = ProjectConfiguration(project_dir=project_dir)
        if project_dir is not None and self.project_dir is None:
            self.project_configuration.set_directories(project_dir)
        if mixed_precision is not None:
            mixed_precision = str(mixed_precision)
            if mixed_precision not in PrecisionType:
                raise ValueError(f"Unknown mixed_precision mode: {mixed_precision}. Choose between {PrecisionType.list()}")
        if dynamo_plugin is not None and dynamo_backend is not None:
            raise ValueError("You cannot pass in both `dynamo_plugin` and `dynamo_backend`, please only pass in one.")
        if dynamo_backend is not None:
            dynamo_plugin = TorchDynamoPlugin(backend=dynamo_backend)
        elif dynamo_plugin is None:
            dynamo_plugin = TorchDynamoPlugin()

        if deepspeed_plugins is not None and deepspeed_plugin is not None:
            raise ValueError("You cannot pass in both `deepspeed_plugins` and `deepspeed_plugin`.")
        elif deepspeed_plugin is not None:
            deepspeed_plugins = deepspeed_plugin
        if deepspeed_plugins is None and PartialState._shared_state == {} and parse_flag_from_env("ACCELERATE_USE_DEEPSPEED"):
            deepspeed_plugins = DeepSpeedPlugin()
        if deepspeed_plugins is not None and not isinstance(deepspeed_plugins, dict):
            if not isinstance(deepspeed_plugins, DeepSpeedPlugin):
                raise TypeError("`deepspeed_plugin` must be a DeepSpeedPlugin or a dict of them")
            deepspeed_plugins = {"default": deepspeed_plugins}
        # kept even when the job does not run in DeepSpeed mode (one process, CPU): `prepare` still reads optimizer / scheduler
        # settings for `DummyOptim` / `DummyScheduler` from it
        self._requested_deepspeed_plugins = deepspeed_plugins
        if deepspeed_plugins is not None:
            for plugin in deepspeed_plugins.values():
                if not isinstance(plugin, DeepSpeedPlugin):
                    raise TypeError("`deepspeed_plugin` must be a DeepSpeedPlugin or a dict of them")
            os.environ["ACCELERATE_USE_DEEPSPEED"] = "true"

        if parse_flag_from_env("ACCELERATE_USE_FSDP") or isinstance(fsdp_plugin, FullyShardedDataParallelPlugin):
            pass
        if fsdp_plugin is None:
            if parse_flag_from_env("ACCELERATE_USE_FSDP"):
                fsdp_plugin = FullyShardedDataParallelPlugin()
        elif not isinstance(fsdp_plugin, FullyShardedDataParallelPlugin):
            raise TypeError("`fsdp_plugin` must be a FullyShardedDataParallelPlugin object.")
        if megatron_lm_plugin is None:
            if parse_flag_from_env("ACCELERATE_USE_MEGATRON_LM"):
                megatron_lm_plugin = MegatronLMPlugin()
        elif not isinstance(megatron_lm_plugin, MegatronLMPlugin):
            raise TypeError("`megatron_lm_plugin` must be a MegatronLMPlugin object.")

        # ---- kwargs handlers
        self.ddp_handler = None
        self.scaler_handler = None
        self.init_handler = None
        self.fp8_recipe_handler = None
        self.ao_recipe_handler = None
        self.te_recipe_handler = None
        self.msamp_recipe_handler = None
        self.autocast_handler = None
        self.profile_handler = None
        self.has_lomo_optimizer = False
        found_handlers = set()
        handler_slots = {
            DistributedDataParallelKwargs: "ddp_handler", GradScalerKwargs: "scaler_handler", InitProcessGroupKwargs: "init_handler",
            AORecipeKwargs: "ao_recipe_handler", TERecipeKwargs: "te_recipe_handler", MSAMPRecipeKwargs: "msamp_recipe_handler",
            AutocastKwargs: "autocast_handler", ProfileKwargs: "profile_handler", FP8RecipeKwargs: "fp8_recipe_handler",
        }
        for handler in kwargs_handlers or []:
            if not isinstance(handler, KwargsHandler):
                raise TypeError(f"Unsupported kwargs handler passed: {handler}, must be one that inherits `accelerate.utils.KwargsHandler`.")
            # exact class first so FP8RecipeKwargs (a TE+MSAMP subclass) lands in its own slot
            slot = handler_slots.get(type(handler))
            if slot is None:
                for cls, name in handler_slots.items():
                    if isinstance(handler, cls):
                        slot = name
                        break
            if slot is None:
                raise ValueError(f"Unsupported kwargs handler {handler.__class__}")
            if type(handler) in found_handlers:
                raise ValueError(f"You can only pass one {handler.__class__} in `kwargs_handlers`.")
            found_handlers.add(type(handler))
            setattr(self, slot, handler)

        if parallelism_config is None and parse_flag_from_env("ACCELERATE_USE_PARALLELISM_CONFIG"):
            parallelism_config = ParallelismConfig()
        if torch_tp_plugin is not None and parallelism_config is None:
            parallelism_config = ParallelismConfig(tp_size=torch_tp_plugin.tp_size)

        kwargs = self.init_handler.to_kwargs() if self.init_handler is not None else {}
        self.state = AcceleratorState(
            mixed_precision=mixed_precision, cpu=cpu, dynamo_plugin=dynamo_plugin, deepspeed_plugin=deepspeed_plugins,
            fsdp_plugin=fsdp_plugin, torch_tp_plugin=torch_tp_plugin, megatron_lm_plugin=megatron_lm_plugin,
            parallelism_config=parallelism_config, _from_accelerator=True, **kwargs,
        )
        if self.parallelism_config is not None:
            self.parallelism_config._validate_accelerator(self)
            if self.parallelism_config.total_size > 1 and self.state.device_mesh is None:
                self.state.device_mesh = self.parallelism_config.get_device_mesh(self.device.type)

        # ---- fp8 back-end selection: native MXFP8 first, then library back-ends
        self.fp8_enabled = self.state.mixed_precision == "fp8" or mixed_precision == "fp8"
        self.has_fp8_handler = any(h is not None for h in (self.fp8_recipe_handler, self.ao_recipe_handler, self.te_recipe_handler, self.msamp_recipe_handler))
        if self.fp8_enabled:
            self._fp8_backend = self._select_fp8_backend()
        else:
            self._fp8_backend = None
        self.delayed_fp8_autocast = False

        trackers = log_with
        from .tracking import filter_trackers

        self.log_with = filter_trackers(trackers, self.logging_dir)

        if (mixed_precision != "bf16") and getattr(self.state, "downcast_bfloat", False) and (self.state.distributed_type != DistributedType.XLA):
            raise ValueError("Can only use `downcast_bf16` when using `mixed_precision='bf16'` and on a TPU")

        # ---- gradient accumulation
        if gradient_accumulation_plugin is not None:
            if gradient_accumulation_steps != 1:
                raise ValueError("You can only pass one of `gradient_accumulation_steps` and `gradient_accumulation_plugin`. Please only pass in the created `GradientAccumulationPlugin` object.")
        else:
            gradient_accumulation_steps = int(parse_choice_from_env("ACCELERATE_GRADIENT_ACCUMULATION_STEPS", gradient_accumulation_steps))
            gradient_accumulation_plugin = GradientAccumulationPlugin(num_steps=gradient_accumulation_steps)
        self.gradient_state = GradientState(gradient_accumulation_plugin=gradient_accumulation_plugin)

        self.device_placement = device_placement
        if dataloader_config is None:
            dataloader_config = DataLoaderConfiguration()
        self.dataloader_config = dataloader_config
        if split_batches is not _split_batches:
            warnings.warn("`split_batches` as an Accelerator argument is deprecated; pass `dataloader_config=DataLoaderConfiguration(split_batches=...)`.", FutureWarning)
            self.dataloader_config.split_batches = split_batches
        self.step_scheduler_with_optimizer = step_scheduler_with_optimizer

        # ---- mixed precision: only fp16 needs a loss scaler
        self.scaler = None
        self.native_amp = False
        if self.state.mixed_precision == "fp16" and self.device.type != "cpu":
            self.native_amp = True
            kwargs = self.scaler_handler.to_kwargs() if self.scaler_handler is not None else {}
            self.scaler = torch.amp.GradScaler("cuda", **kwargs)
        elif self.state.mixed_precision == "fp16":
            self.native_amp = True
            kwargs = self.scaler_handler.to_kwargs() if self.scaler_handler is not None else {}
            self.scaler = torch.amp.GradScaler("cpu", **kwargs)
        elif self.state.mixed_precision == "bf16":
            self.native_amp = True
        elif self.state.mixed_precision == "fp8":
            self.native_amp = True

        self.step = 0
        self._optimizers = []
        self._models = []
        self._schedulers = []
        self._dataloaders = []
        self._custom_objects = []
        self._load_model_state_pre_hook = OrderedDict()
        self._save_model_state_pre_hook = OrderedDict()
        self.rng_types = rng_types
        if self.rng_types is None:
            self.rng_types = ["generator"]
        self.flag_tensor = None
        self._cp_context = None
        check_os_kernel()

    def _select_fp8_backend(self) -> FP8BackendType:
        from .utils.imports import is_b200_kernels_available, is_msamp_available, is_torchao_available, is_transformer_engine_available

        env = os.environ.get("ACCELERATE_FP8_BACKEND", None)
        if env is not None:
            return FP8BackendType(env.upper())
        if self.ao_recipe_handler is not None:
            return FP8BackendType.AO
        if self.msamp_recipe_handler is not None:
            return FP8BackendType.MSAMP
        if self.te_recipe_handler is not None and is_transformer_engine_available():
            return FP8BackendType.TE
        if self.device.type == "cuda" and is_b200_kernels_available():
            return FP8BackendType.NATIVE
        if is_torchao_available():
            return FP8BackendType.AO
        if is_transformer_engine_available():
            return FP8BackendType.TE
        if is_msamp_available():
            return FP8BackendType.MSAMP
        raise ImportError("fp8 mixed precision needs the native kernel library on a CUDA device, or torchao / transformer-engine / MS-AMP.")

    # ================================================================== properties
    @property
    def deepspeed_plugin(self):
        return self.state.deepspeed_plugin

    # read-only views of the process state, the dataloader configuration and the project configuration
    use_distributed = _forwarded("state", "use_distributed")
    distributed_type = _forwarded("state", "distributed_type")
    num_processes = _forwarded("state", "num_processes")
    process_index = _forwarded("state", "process_index")
    local_process_index = _forwarded("state", "local_process_index")
    device = _forwarded("state", "device")
    is_main_process = _forwarded("state", "is_main_process")
    is_local_main_process = _forwarded("state", "is_local_main_process")
    mixed_precision = _forwarded("state", "mixed_precision")
    is_fsdp2 = _forwarded("state", "is_fsdp2")
    split_batches = _forwarded("dataloader_config", "split_batches")
    dispatch_batches = _forwarded("dataloader_config", "dispatch_batches")
    even_batches = _forwarded("dataloader_config", "even_batches", settable=True)
    use_seedable_sampler = _forwarded("dataloader_config", "use_seedable_sampler")
    non_blocking = _forwarded("dataloader_config", "non_blocking")
    use_stateful_dataloader = _forwarded("dataloader_config", "use_stateful_dataloader", default=False)
    project_dir = _forwarded("project_configuration", "project_dir")
    logging_dir = _forwarded("project_configuration", "logging_dir")
    save_iteration = _forwarded("project_configuration", "iteration")

    @property
    def multi_device(self):
        return self.use_distributed and self.distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU)

    @property
    def is_last_process(self):
        return self.process_index == self.num_processes - 1

    @property
    def is_composable_parallelism_enabled(self):
        return self.is_fsdp2

    @property
    def parallelism_config(self) -> Optional[ParallelismConfig]:
        return self.state.parallelism_config

    @property
    def torch_device_mesh(self):
        return self.state.device_mesh

    @property
    def should_save_model(self):
        """False on ranks that hold a replica another rank already saves (non-zero tp rank etc.)."""
        if (pc := self.parallelism_config) is None or pc.total_size == 1:
            return True
        mesh = self.torch_device_mesh
        names = tuple(mesh.mesh_dim_names or ()) if mesh is not None else ()
        for axis in ("dp_replicate", "cp", "sp"):
            if axis in names and mesh.get_local_rank(axis) != 0:
                return False
        return True

    def _mesh_rank(self, axis: str, what: str) -> int:
        mesh = self.torch_device_mesh
        if mesh is None or axis not in (mesh.mesh_dim_names or ()):
            raise RuntimeError(f"{what} is not enabled. Please check your configuration.")
        return mesh.get_local_rank(axis)

    @property
    def tensor_parallel_rank(self) -> int:
        return self._mesh_rank("tp", "Tensor parallelism")

    @property
    def pipeline_parallel_rank(self) -> int:
        raise NotImplementedError("Pipeline parallelism is currently not supported in Accelerate.")

    @property
    def context_parallel_rank(self) -> int:
        """This process's position in its context-parallel group (the reference raises here: its context parallelism is torch's
        and gives no rank; the native one has a mesh axis like the others)."""
        return self._mesh_rank("cp", "Context parallelism")

    @property
    def sequence_parallel_rank(self) -> int:
        """This process's position in its Ulysses sequence-parallel group."""
        return self._mesh_rank("sp", "Sequence parallelism")

    @property
    def data_parallel_rank(self) -> int:
        return self._mesh_rank("dp_replicate", "Data parallelism")

    @property
    def data_parallel_shard_rank(self) -> int:
        return self._mesh_rank("dp_shard", "Data parallelism")

    @property
    def sync_gradients(self):
        return self.gradient_state.sync_gradients

    @sync_gradients.setter
    def sync_gradients(self, sync_gradients):
        self.gradient_state.sync_gradients = sync_gradients

    @property
    def gradient_accumulation_steps(self):
        return self.gradient_state.num_steps

    @gradient_accumulation_steps.setter
    def gradient_accumulation_steps(self, gradient_accumulation_steps):
        self.gradient_state.plugin_kwargs.update({"num_steps": gradient_accumulation_steps})

    @property
    def optimizer_step_was_skipped(self):
        """True if the last optimizer step was skipped (fp16 overflow)."""
        return any(opt.step_was_skipped for opt in self._optimizers)

    @property
    def fp8_backend(self):
        """The fp8 back-end in use (``NATIVE`` = own MXFP8 kernels), or None."""
        return self._fp8_backend

    # ================================================================== process control passthroughs
    @contextmanager
    def split_between_processes(self, inputs, apply_padding: bool = False):
        with PartialState().split_between_processes(inputs, apply_padding=apply_padding) as inputs:
            yield inputs

    def _proc_decorator(self, state_method, function, **kw):
        if function is None:
            if "Accelerator." in getattr(self, "__qualname__", ""):
                function = self
            else:
                raise ValueError("The decorator must wrap a function.")

        def _inner(*args, **kwargs):
            return state_method(function, **kw)(*args, **kwargs)

        return functools.wraps(function)(_inner)

    def on_main_process(self, function: Callable[..., Any] = None):
        return self._proc_decorator(PartialState().on_main_process, function)

    def on_local_main_process(self, function: Callable[..., Any] = None):
        return self._proc_decorator(PartialState().on_local_main_process, function)

    def on_last_process(self, function: Callable[..., Any]):
        return self._proc_decorator(PartialState().on_last_process, function)

    def on_process(self, function: Callable[..., Any] = None, process_index: int = None):
        if (self is not None) and (process_index is not None) and (function is None):
            return partial(self.on_process, process_index=process_index)
        return self._proc_decorator(PartialState().on_process, function, process_index=process_index)

    def on_local_process(self, function: Callable[..., Any] = None, local_process_index: int = None):
        if (self is not None) and (local_process_index is not None) and (function is None):
            return partial(self.on_local_process, local_process_index=local_process_index)
        return self._proc_decorator(PartialState().on_local_process, function, local_process_index=local_process_index)

    @contextmanager
    def main_process_first(self):
        with self.state.main_process_first():
            yield

    @contextmanager
    def local_main_process_first(self):
        with self.state.local_main_process_first():
            yield

    def wait_for_everyone(self):
        wait_for_everyone()

    def print(self, *args, **kwargs):
        self.state.print(*args, **kwargs)

    # ================================================================== gradient synchronisation
    @contextmanager
    def no_sync(self, model):
        """Skip the gradient reduction of the wrapped ``model`` for the steps inside the block."""
        if self.is_fsdp2 and hasattr(model, "set_requires_gradient_sync"):
            model.set_requires_gradient_sync(False)
            try:
                yield
            finally:
                model.set_requires_gradient_sync(True)
            return
        context = contextlib.nullcontext
        if self.use_distributed:
            context = getattr(model, "no_sync", context)
        with context():
            yield

    @staticmethod
    @contextmanager
    def trigger_sync_in_backward(model):
        """Re-enable the reduction for a backward whose forward ran under ``no_sync``."""
        if hasattr(model, "trigger_sync_in_backward"):
            with model.trigger_sync_in_backward():
                yield
            return
        if not isinstance(model, torch.nn.parallel.DistributedDataParallel):
            yield
            return
        old_require_backward_grad_sync = model.require_backward_grad_sync
        old_require_forward_param_sync = model.require_forward_param_sync
        model.require_backward_grad_sync = True
        model.require_forward_param_sync = True
        model.reducer.prepare_for_backward([])
        try:
            yield
        finally:
            model.require_backward_grad_sync = old_require_backward_grad_sync
            model.require_forward_param_sync = old_require_forward_param_sync

    def _do_sync(self):
        """Decide whether the current micro-step ends an accumulation window."""
        if self.gradient_state.sync_with_dataloader and self.gradient_state.end_of_dataloader:
            self.step = 0
            self.gradient_state._set_sync_gradients(True)
        else:
            self.step += 1
            self.gradient_state._set_sync_gradients((self.step % self.gradient_state.num_steps) == 0)

    @contextmanager
    def accumulate(self, *models):
        """Gradient accumulation: inside the block ``sync_gradients`` says whether this is a real step."""
        self._do_sync()
        allow_gradient_sync = self.sync_gradients or (
            self.use_distributed and self.gradient_state.plugin_kwargs.get("sync_each_batch", False)
        )
        with contextlib.ExitStack() as cm_stack:
            for m in models:
                cm_stack.enter_context(contextlib.nullcontext() if allow_gradient_sync else self.no_sync(m))
            yield

    @contextmanager
    def join_uneven_inputs(self, joinables, even_batches=None):
        """Train on uneven per-rank inputs (torch ``Join``); optionally override ``even_batches`` inside."""
        if self.multi_device or self.distributed_type == DistributedType.FSDP:
            from torch.distributed.algorithms.join import Join

            dl_even_batches_values = []
            if even_batches is not None:
                iterable_dl_seen = False
                for dl_idx, dl in enumerate(self._dataloaders):
                    if isinstance(dl, DataLoaderDispatcher) or type(dl) is DataLoaderDispatcher:
                        iterable_dl_seen = True
                        continue
                    if getattr(dl, "batch_sampler", None) is not None and hasattr(dl.batch_sampler, "even_batches"):
                        dl_even_batches_values.append((dl_idx, dl.batch_sampler.even_batches))
                        dl.batch_sampler.even_batches = even_batches
                if iterable_dl_seen:
                    warnings.warn("Overridding even_batches is only supported for map-style datasets, yet some dataloaders given were iterable")
            joinables = [j for j in joinables if hasattr(j, "join_hook")]
            try:
                if joinables:
                    with Join(joinables):
                        yield
                else:
                    yield
            finally:
                for dl_idx, value in dl_even_batches_values:
                    self._dataloaders[dl_idx].batch_sampler.even_batches = value
        else:
            if self.distributed_type != DistributedType.NO:
                warnings.warn("Joining uneven inputs is only supported for multi-GPU training, as a result `join_uneven_inputs` will have no effect.")
            with contextlib.nullcontext(joinables):
                yield

    # ================================================================== prepare
    def _prepare_one(self, obj, first_pass: bool = False, device_placement=None):
        if first_pass:
            if _is_dataloader(obj):
                return self.prepare_data_loader(obj, device_placement=device_placement)
            if _is_module(obj):
                return self.prepare_model(obj, device_placement=device_placement)
            if _is_optimizer(obj):
                optimizer = self.prepare_optimizer(obj, device_placement=device_placement)
                return optimizer
        elif _is_scheduler(obj):
            return self.prepare_scheduler(obj)
        return obj

    def prepare(self, *args, device_placement=None):
        """Prepare models, optimizers, dataloaders and schedulers; returns them in the order given."""
        if device_placement is None:
            device_placement = [None for _ in args]
        elif self.distributed_type in (DistributedType.DEEPSPEED, DistributedType.MEGATRON_LM):
            raise ValueError("You can't customize device placements with DeepSpeed or Megatron-LM.")
        elif len(device_placement) != len(args):
            raise ValueError(f"`device_placement` should be a list with {len(args)} elements (the number of objects passed).")

        for obj in args:
            if _is_module(obj) and self.verify_device_map(obj) and self.distributed_type != DistributedType.NO and os.environ.get("ACCELERATE_BYPASS_DEVICE_MAP", "false") != "true":
                raise ValueError(
                    "You can't train a model that has been loaded with `device_map='auto'` in any distributed mode."
                    " Please rerun your script specifying `--num_processes=1` or by launching with `python {{myscript.py}}`."
                )
        if self.distributed_type in (DistributedType.FSDP,) or self.is_fsdp2:
            n_models = sum(1 for o in args if _is_module(o))
            n_opts = sum(1 for o in args if _is_optimizer(o))
            if self.is_fsdp2 and n_opts > 0 and n_models == 0:
                raise ValueError("When using FSDP2, a model and optimizer must be passed together to `Accelerator.prepare()` as the optimizer needs to have its parameters modified after the model is converted.")
            if self.is_fsdp2 and n_models > 1:
                raise ValueError("Only one model is supported when using FSDP2")

        if self.distributed_type == DistributedType.DEEPSPEED and self.deepspeed_plugin is not None:
            self._resolve_deepspeed_config(args)
        if self.distributed_type == DistributedType.MEGATRON_LM and self.state.megatron_lm_plugin is not None:
            self._record_megatron_arguments(args)
        args = self._materialize_placeholders(args)
        # already-prepared objects pass through untouched (idempotence)
        result = []
        models_this_call = []
        optimizers_this_call = []
        for obj, placement in zip(args, device_placement):
            if getattr(obj, "_is_accelerate_prepared", False):
                result.append(obj)
                continue
            if _is_module(obj):
                new = self.prepare_model(obj, device_placement=placement)
                models_this_call.append((obj, new))
                result.append(new)
            elif _is_dataloader(obj):
                result.append(self.prepare_data_loader(obj, device_placement=placement))
            elif _is_optimizer(obj):
                new = self.prepare_optimizer(obj, device_placement=placement)
                optimizers_this_call.append(new)
                result.append(new)
            else:
                result.append(obj)
        # optimizers learn about the engines that own their parameters (sharded update, fused clip)
        for _, engine_model in models_this_call:
            remap = getattr(engine_model, "_b200_param_remap", None)
            if remap:
                # the model's parameters were replaced by `prepare` (torch FSDP2 fallback: same repair the reference makes after
                # fully_shard, accelerator.py:1690-1744; a transformers Llama adopted as the native model): re-point the optimizers.
                # A replacement of None means "folded into another parameter"; nothing is listed twice.
                for opt in optimizers_this_call or self._optimizers:
                    listed = set()
                    for group in opt.optimizer.param_groups:
                        fresh = []
                        for q in group["params"]:
                            new_q = remap.get(id(q), q)
                            if new_q is not None and id(new_q) not in listed:
                                listed.add(id(new_q))
                                fresh.append(new_q)
                        group["params"] = fresh
                    opt.optimizer.state.clear()
            engine = engine_model if hasattr(engine_model, "bind_optimizer") else None
            if engine is not None:
                for opt in optimizers_this_call or self._optimizers:
                    engine.bind_optimizer(opt)
        # schedulers last: they need the prepared optimizers
        result = [self.prepare_scheduler(o) if (_is_scheduler(o) and not getattr(o, "_is_accelerate_prepared", False)) else o for o in result]
        for item in result:
            if any(item in container for container in (self._dataloaders, self._models, self._optimizers, self._schedulers)):
                try:
                    item._is_accelerate_prepared = True
                except Exception:
                    pass
        return tuple(result) if len(result) > 1 else result[0]

    def _record_megatron_arguments(self, args) -> None:
        """Under a ``MegatronLMPlugin`` the objects handed to ``prepare`` are described in Megatron's argument names on the plugin
        (``megatron_lm_default_args``): network size from the model's config, batch geometry from the dataloader, the optimizer's
        rule and the dummy scheduler's step counts — what the reference's ``_prepare_megatron_lm`` (accelerator.py:2478-2579)
        collects before it initialises Megatron.  The native engines do not read them; scripts and ``MegatronEngine`` users do."""
        from .utils.megatron_lm import MegatronLMDummyScheduler

        plugin = self.state.megatron_lm_plugin
        model = next((obj for obj in args if _is_module(obj)), None)
        loader = next((obj for obj in args if _is_dataloader(obj)), None)
        optimizer = next((obj for obj in args if _is_optimizer(obj)), None)
        scheduler = next((obj for obj in args if isinstance(obj, MegatronLMDummyScheduler)), None)
        pc = self.parallelism_config
        model_parallel = pc.non_data_parallel_size if pc is not None else max(1, (plugin.tp_degree or 1) * (plugin.context_parallel_size or 1))
        micro_batch = None
        if loader is not None:
            micro_batch = getattr(loader, "batch_size", None) or getattr(getattr(loader, "batch_sampler", None), "batch_size", None)
        if micro_batch is not None:
            plugin.set_training_args(micro_batch, max(1, self.num_processes // model_parallel))
        if model is not None and getattr(getattr(model, "config", None), "model_type", None):
            try:
                plugin.set_network_size_args(model, None)
            except ValueError as error:
                logger.warning(f"{error} (the model trains all the same; only the Megatron-style description of it is missing)")
        plugin.set_mixed_precision(self.mixed_precision)
        if optimizer is not None:
            try:
                plugin.set_optimizer_type(optimizer)
            except ValueError as error:
                logger.warning(str(error))
        if scheduler is not None and "data_parallel_size" in plugin.megatron_lm_default_args:
            plugin.set_scheduler_args(scheduler)

    def _resolve_deepspeed_config(self, args) -> None:
        """Fill the ``"auto"`` entries of the active DeepSpeed-style configuration from what ``prepare`` was handed, and refuse
        combinations that say the same thing twice.  What the reference does at the top of ``_prepare_deepspeed``
        (accelerator.py:2132-2330) before building DeepSpeed's engine; here the resolved configuration drives the native engines:

        * micro batch size from the dataloaders (smallest or largest, ``is_train_batch_min``), ``train_batch_size`` from it;
        * gradient accumulation: the configuration's value wins over the ``Accelerator`` argument (with a warning);
        * bucket sizes that depend on the model's hidden size; optimizer / scheduler parameters from ``DummyOptim`` / ``DummyScheduler``;
        * an optimizer (scheduler) described in the configuration *and* built in code, or a placeholder without a description, is an error;
        * a numeric ``gradient_clipping`` is applied at every optimizer step (``backward`` does it on synchronising steps).
        """
        from .utils.deepspeed import DummyOptim, DummyScheduler

        plugin = self.deepspeed_plugin
        config = plugin.deepspeed_config
        loaders = [obj for obj in args if _is_dataloader(obj)]
        if plugin.is_auto("train_micro_batch_size_per_gpu"):
            if not loaders:
                raise ValueError(
                    "When using DeepSpeed, `accelerate.prepare()` requires you to pass at least one of training or evaluation dataloaders "
                    "with `batch_size` attribute returning an integer value "
                    "or alternatively set an integer value in `train_micro_batch_size_per_gpu` in the deepspeed config file "
                    "or assign integer value to `AcceleratorState().deepspeed_plugin.deepspeed_config['train_micro_batch_size_per_gpu']`.")
            sizes = [getattr(dl, "batch_size", None) if getattr(dl, "batch_size", None) is not None
                     else getattr(getattr(dl, "batch_sampler", None), "batch_size", None) for dl in loaders]
            if any(size is None for size in sizes):
                raise ValueError(
                    "At least one of the dataloaders passed to `accelerate.prepare()` has `None` as batch size. "
                    "Please set an integer value in `train_micro_batch_size_per_gpu` in the deepspeed config file "
                    "or assign integer value to `AcceleratorState().deepspeed_plugin.deepspeed_config['train_micro_batch_size_per_gpu']`.")
            if self.split_batches:
                sizes = [size // self.num_processes for size in sizes]
            micro_batch = min(sizes) if plugin.is_train_batch_min else max(sizes)
            if len(sizes) > 1:
                logger.info(f"Since you passed both train and evaluation dataloader, `is_train_batch_min` (here {plugin.is_train_batch_min} "
                            f"will decide the `train_batch_size` ({micro_batch}).")
        else:
            micro_batch = plugin.get_value("train_micro_batch_size_per_gpu")

        plugin.fill_match("gradient_accumulation_steps", must_match=False, gradient_accumulation_steps=self.gradient_accumulation_steps)
        stated_steps = plugin.get_value("gradient_accumulation_steps")
        if stated_steps != self.gradient_accumulation_steps:
            logger.warning(f"Gradient accumulation steps mismatch: GradientAccumulationPlugin has {self.gradient_accumulation_steps}, "
                           f"DeepSpeed config has {stated_steps}. Using DeepSpeed's value.")
            self.gradient_accumulation_steps = stated_steps

        known = {"gradient_clipping": 1.0, "zero_optimization.stage3_gather_16bit_weights_on_model_save": False}
        if micro_batch is not None:
            pc = self.parallelism_config
            sequence_ranks = pc.sp_size if (pc is not None and pc.sp_enabled) else 1
            known["train_micro_batch_size_per_gpu"] = micro_batch
            known["train_batch_size"] = micro_batch * stated_steps * self.num_processes // sequence_ranks

        model = next((obj for obj in args if _is_module(obj)), None)
        optimizer = next((obj for obj in args if isinstance(obj, DummyOptim) or _is_optimizer(obj)), None)
        scheduler = next((obj for obj in args if isinstance(obj, DummyScheduler) or _is_scheduler(obj)), None)
        if optimizer is not None:
            placeholder = isinstance(optimizer, DummyOptim)
            if "optimizer" in config and not placeholder:
                raise ValueError("You cannot specify an optimizer in the config file and in the code at the same time. "
                                 "Please remove the optimizer from the config file or create `accelerate.utils.DummyOptim` in the code.")
            if "optimizer" not in config and placeholder:
                raise ValueError("You cannot create a `DummyOptim` without specifying an optimizer in the config file.")
            if not placeholder:
                config["zero_allow_untested_optimizer"] = True
        if scheduler is not None:
            placeholder = isinstance(scheduler, DummyScheduler)
            if "scheduler" in config and not placeholder:
                raise ValueError("You cannot specify a scheduler in the config file and in the code at the same time. "
                                 "Please remove the scheduler from the config file or create `accelerate.utils.DummyScheduler` in the code.")
            if "scheduler" not in config and placeholder and scheduler.lr_scheduler_callable is None:
                raise ValueError("Either specify a scheduler in the config file or "
                                 "pass in the `lr_scheduler_callable` parameter when using `accelerate.utils.DummyScheduler`.")
        if isinstance(optimizer, DummyOptim) and scheduler is not None and not isinstance(scheduler, DummyScheduler):
            raise ValueError("You can only specify `accelerate.utils.DummyScheduler` in the code when using `accelerate.utils.DummyOptim`.")

        if model is not None:
            plugin.set_moe_leaf_modules(model)
            by_hidden_size = ("zero_optimization.reduce_bucket_size", "zero_optimization.stage3_prefetch_bucket_size",
                              "zero_optimization.stage3_param_persistence_threshold")
            open_entries = [path for path in by_hidden_size if plugin.is_auto(path)]
            if open_entries:
                why = ("therefore it's not possible to automatically fill out the following `auto` entries "
                       f"in the DeepSpeed config file: {open_entries}. You can fix that by replacing "
                       "`auto` values for these keys with an integer value of your choice.")
                model_config = getattr(model, "config", None)
                if model_config is None:
                    raise ValueError("Can't find `model.config` entry, " + why)
                if hasattr(model_config, "hidden_size"):
                    hidden = model_config.hidden_size
                elif hasattr(model_config, "hidden_sizes"):
                    hidden = max(model_config.hidden_sizes)
                else:
                    raise ValueError("Can find neither `model.config.hidden_size` nor `model.config.hidden_sizes`, " + why)
                known.update({by_hidden_size[0]: hidden * hidden, by_hidden_size[1]: int(0.9 * hidden * hidden), by_hidden_size[2]: 10 * hidden})
            if isinstance(optimizer, DummyOptim):
                known.update({"optimizer.params.lr": optimizer.lr, "optimizer.params.weight_decay": optimizer.weight_decay})
            if isinstance(scheduler, DummyScheduler) and scheduler.lr_scheduler_callable is None:
                peak = getattr(scheduler.optimizer, "lr", None)
                if peak is None:
                    peak = getattr(scheduler.optimizer, "defaults", {}).get("lr")
                known.update({"scheduler.params.warmup_min_lr": 0, "scheduler.params.warmup_max_lr": peak,
                              "scheduler.params.warmup_num_steps": scheduler.warmup_num_steps})
                if scheduler.total_num_steps is not None:
                    ranks = self.num_processes if not self.split_batches else 1
                    known["scheduler.params.total_num_steps"] = -(-scheduler.total_num_steps // ranks)
            plugin.deepspeed_config_process(must_match=False, **known)
        else:
            # no model in this call (dataloaders prepared on their own): fill what can be filled, leave the rest for the call that has it
            for path, value in known.items():
                if plugin.is_auto(path):
                    plugin.fill_match(path, must_match=False, **{path: value})
        clip = plugin.get_value("gradient_clipping")
        self._config_clip_norm = float(clip) if isinstance(clip, (int, float)) and not isinstance(clip, bool) and clip > 0 else None

    def _materialize_placeholders(self, args) -> tuple:
        """Scripts written for a DeepSpeed config pass ``DummyOptim`` / ``DummyScheduler`` where the config file describes the
        optimizer and the schedule.  Here such a configuration runs on the native engines, so the placeholders become real
        objects: the optimizer from the placeholder's own arguments overlaid with the config's ``optimizer.params``, the
        scheduler bound to that optimizer (``scheduler.params`` of the config: ``total_num_steps`` / ``warmup_num_steps``)."""
        from .utils.deepspeed import DummyOptim, DummyScheduler

        if not any(isinstance(obj, (DummyOptim, DummyScheduler)) for obj in args):
            return args
        plugin = self.deepspeed_plugin  # None unless the job runs in DeepSpeed mode ...
        if plugin is None and self._requested_deepspeed_plugins:
            plugin = next(iter(self._requested_deepspeed_plugins.values()))  # ... but a requested configuration still applies

        def setting(path, fallback):
            value = plugin.get_value(path) if plugin is not None else None
            return fallback if value in (None, "auto") else value

        optimizer = None
        out = list(args)
        for i, obj in enumerate(out):
            if isinstance(obj, DummyOptim):
                obj.lr = setting("optimizer.params.lr", obj.lr)
                obj.weight_decay = setting("optimizer.params.weight_decay", obj.weight_decay)
                for key in ("betas", "eps"):
                    if setting(f"optimizer.params.{key}", None) is not None:
                        obj.kwargs[key] = tuple(setting(f"optimizer.params.{key}", None)) if key == "betas" else setting(f"optimizer.params.{key}", None)
                out[i] = optimizer = obj.materialize()
        for i, obj in enumerate(out):
            if isinstance(obj, DummyScheduler):
                target = optimizer if optimizer is not None else obj.optimizer
                if isinstance(target, DummyOptim):
                    raise ValueError("A `DummyScheduler` needs its optimizer (a real one, or a `DummyOptim` passed to the same `prepare` call).")
                obj.total_num_steps = setting("scheduler.params.total_num_steps", obj.total_num_steps)
                obj.warmup_num_steps = setting("scheduler.params.warmup_num_steps", obj.warmup_num_steps)
                out[i] = obj.materialize(target)
        return tuple(out)

    def _mixed_precision_dtype(self):
        return {"fp16": torch.float16, "bf16": torch.bfloat16, "fp8": torch.bfloat16}.get(self.mixed_precision)

    def prepare_model(self, model: torch.nn.Module, device_placement: Optional[bool] = None, evaluation_mode: bool = False):
        """Place ``model`` on the device, add mixed precision and wrap it in the engine of the active strategy."""
        if device_placement is None:
            device_placement = self.device_placement and self.distributed_type != DistributedType.FSDP
        self._models.append(model)

        if self.verify_device_map(model) and self.distributed_type != DistributedType.NO and os.environ.get("ACCELERATE_BYPASS_DEVICE_MAP", "false") != "true":
            raise ValueError("You can't train a model that has been loaded with `device_map='auto'` in any distributed mode.")

        # ---- mixed precision on the forward
        if self.native_amp:
            model._original_forward = model.forward
            autocast_context = self._get_autocast_context()
            model_forward_func = model.forward.__func__ if hasattr(model.forward, "__func__") else model.forward
            new_forward = autocast_context(model_forward_func)
            if hasattr(model.forward, "__func__"):
                model.forward = MethodType(new_forward, model)
                model.forward = MethodType(convert_outputs_to_fp32(model.forward.__func__), model)
            else:
                model.forward = convert_outputs_to_fp32(new_forward)

        if self.fp8_enabled and self._fp8_backend == FP8BackendType.NATIVE:
            from .parallel.precision import convert_model_to_mxfp8

            recipe = self.te_recipe_handler
            model = convert_model_to_mxfp8(model, recipe)

        if (getattr(model, "is_loaded_in_8bit", False) or getattr(model, "is_loaded_in_4bit", False)) and getattr(model, "hf_device_map", False):
            model_devices = set(model.hf_device_map.values())
            if len(model_devices) > 1 and self.distributed_type != DistributedType.NO:
                raise ValueError("You can't train a model that has been loaded in 8-bit or 4-bit precision on multiple devices in any distributed mode.")
        elif device_placement and not self.verify_device_map(model):
            model = model.to(self.device)

        if not evaluation_mode:
            pc = self.parallelism_config
            model_parallel = pc is not None and (pc.tp_enabled or pc.sp_enabled or pc.cp_enabled)
            if model_parallel or self.distributed_type in (DistributedType.FSDP, DistributedType.DEEPSPEED, DistributedType.MEGATRON_LM):
                # tensor / sequence / context parallelism always runs inside the sharded engine, also without a sharding plugin
                # (dp_shard = 1: a shard group of one, no gather or reduce kernels): ranks of one model-parallel group hold
                # different weight shards, so a data-parallel all-reduce over all ranks would be wrong
                model = self._wrap_sharded(model)
            elif self.distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
                if any(p.requires_grad for p in model.parameters()):
                    model = self._wrap_data_parallel(model)
        # activation checkpointing can be requested through the FSDP plugin even on one process
        plugin = self.state.fsdp_plugin
        if plugin is not None and plugin.activation_checkpointing and self.distributed_type == DistributedType.NO:
            from .parallel.fsdp import apply_activation_checkpointing

            plugin.set_auto_wrap_policy(model)
            apply_activation_checkpointing(model, plugin)

        if self.state.dynamo_plugin.backend != DynamoBackend.NO and not is_compiled_module(model):
            if self.state.dynamo_plugin.use_regional_compilation:
                model = compile_regions(model, **self.state.dynamo_plugin.to_kwargs())
            else:
                model = torch.compile(model, **self.state.dynamo_plugin.to_kwargs())
        self._models[-1] = model
        return model

    def _wrap_data_parallel(self, model):
        kwargs = self.ddp_handler.to_kwargs() if self.ddp_handler is not None else {}
        pc = self.parallelism_config
        if pc is not None and pc.tp_enabled and self.device.type == "cuda":
            from .parallel.tensor_parallel import apply_tensor_parallel

            model = apply_tensor_parallel(model, self.torch_device_mesh["tp"], pc.tp_handler)
        if self.device.type == "cuda" and os.environ.get("ACCELERATE_B200_DDP", "1") != "0":
            from .parallel.ddp import B200DataParallel
            from .parallel.symm import peer_domain
            from .utils.dataclasses import DDPCommunicationHookType

            group = None
            if self.torch_device_mesh is not None and "dp_replicate" in (self.torch_device_mesh.mesh_dim_names or ()):
                group = self.torch_device_mesh["dp_replicate"].get_group()
            ok, why = peer_domain(group)
            hook = getattr(self.ddp_handler, "comm_hook", None)
            if hook in (DDPCommunicationHookType.POWER_SGD, DDPCommunicationHookType.BATCHED_POWER_SGD):
                ok, why = False, f"comm_hook={hook} is a torch DDP communication hook"
            if ok:
                return B200DataParallel(model, process_group=group, handler=self.ddp_handler, gradient_state=self.gradient_state)
            # groups that span hosts / more than 8 ranks, or torch-only comm hooks: the reference's own path (torch DDP over NCCL)
            logger.warning(f"native data-parallel engine not used ({why}); falling back to torch DistributedDataParallel", main_process_only=True)
        device_ids, output_device = (None, None)
        if self.device.type == "cuda":
            device_ids, output_device = [self.local_process_index], self.local_process_index
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=device_ids, output_device=output_device, **kwargs)
        if self.ddp_handler is not None:
            self.ddp_handler.register_comm_hook(model)
        return model

    def _adopt_for_native_parallelism(self, model):
        """Tensor / sequence / context parallelism are implemented on the native Llama.  A ``transformers`` Llama (recognised by
        its config) is therefore rebuilt as the native model with the same weights — where the reference accepts a
        ``tp_plan="auto"`` transformers model.  The optimizer, created on the transformers parameters, is re-pointed in
        ``prepare`` through ``_b200_param_remap`` (fused projections take the group of their first part)."""
        from .models.llama import LlamaForCausalLM

        if isinstance(model, LlamaForCausalLM) or getattr(getattr(model, "config", None), "model_type", None) != "llama":
            return model
        on_gpu = self.device.type == "cuda"
        native = LlamaForCausalLM.from_transformers(model, dtype=torch.bfloat16 if on_gpu else None, device=self.device)
        theirs = dict(model.named_parameters())
        remap = {id(p): None for p in theirs.values()}  # None: folded into a fused native parameter, dropped from the optimizer
        representative = {}
        for name, _ in native.named_parameters():
            representative[name] = name.replace("qkv_proj", "q_proj").replace("gate_up_proj", "gate_proj")
        mine = dict(native.named_parameters())
        for name, param in mine.items():
            source = theirs.get(representative[name])
            if source is not None:
                remap[id(source)] = param
        native._b200_param_remap = remap
        logger.warning("a transformers Llama was rebuilt as the native Llama (same weights) so that tensor / sequence / context "
                       "parallelism can run on the fused kernels; use `accelerator.unwrap_model(model).hf_state_dict()` to export "
                       "weights under transformers' names", main_process_only=True)
        return native

    def _wrap_sharded(self, model):
        from .parallel.fsdp import B200ShardedModule

        plugin = self.state.fsdp_plugin
        if plugin is None and self.distributed_type == DistributedType.DEEPSPEED:
            plugin = self.deepspeed_plugin.to_fsdp_plugin()
        if plugin is None:
            plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="no_wrap")
        if self.mixed_precision != "no":
            plugin.set_mixed_precision(self.mixed_precision)
        pc = self.parallelism_config
        mesh = self.torch_device_mesh
        if pc is not None and (pc.tp_enabled or pc.sp_enabled or pc.cp_enabled):
            model = self._adopt_for_native_parallelism(model)
        if pc is not None and pc.tp_enabled:
            from .parallel.tensor_parallel import apply_tensor_parallel

            model = apply_tensor_parallel(model, mesh["tp"], pc.tp_handler)
        if pc is not None and pc.sp_enabled:
            from .parallel.ulysses import apply_ulysses

            model = apply_ulysses(model, mesh["sp"], pc.sp_handler)
        if pc is not None and pc.cp_enabled:
            from .parallel.context_parallel import apply_context_parallel

            model = apply_context_parallel(model, mesh["cp"], pc.cp_handler)
        shard_group = replicate_group = None
        if mesh is not None:
            names = mesh.mesh_dim_names or ()
            if pc is not None and ("dp_shard" in names or "cp" in names or "sp" in names):
                shard_group = mesh["dp_shard_cp"].get_group()
            elif "dp_replicate" not in names:
                # only model-parallel axes (tp) are active: every rank keeps its whole (already tp-sharded) parameters.
                # new_group is collective, so all ranks create all singleton groups in the same order.
                for r in range(self.num_processes):
                    g = torch.distributed.new_group([r])
                    if r == self.process_index:
                        shard_group = g
            if "dp_replicate" in names:
                replicate_group = mesh["dp_replicate"].get_group()
        plugin.set_auto_wrap_policy(model)
        dp_group = None
        if replicate_group is not None and mesh is not None:
            names = mesh.mesh_dim_names or ()
            flat = "dp_cp" if ("cp" in names or "sp" in names) else "dp"
            try:
                dp_group = mesh[flat].get_group()
            except Exception:  # noqa: BLE001 - no flattened mesh of that name: keep the two-collective path
                dp_group = None
        fallback = self._torch_fsdp2_reason(model, plugin, dp_group if dp_group is not None else shard_group)
        if fallback:
            logger.warning(f"native sharded engine not used ({fallback}); falling back to torch FSDP2 (fully_shard)", main_process_only=True)
            return self._wrap_torch_fsdp2(model, plugin, mesh)
        return B200ShardedModule(model, plugin=plugin, device=self.device, shard_group=shard_group, replicate_group=replicate_group,
                                 gradient_state=self.gradient_state, shard_group_is_world=(mesh is None), dp_group=dp_group)

    def _torch_fsdp2_reason(self, model, plugin, group) -> str:
        """Why the native sharded engine cannot take this job ('' if it can).  It needs one NVLink domain (<= 8 ranks on one
        host) and bf16 compute on CUDA; everything else runs the reference's path: torch FSDP2 over NCCL."""
        if os.environ.get("ACCELERATE_B200_FSDP", "1") == "0":
            return "ACCELERATE_B200_FSDP=0"
        if self.device.type != "cuda":
            return ""  # the gloo oracle path of the native engine
        from .parallel.symm import peer_domain

        ok, why = peer_domain(group)
        if not ok:
            return why
        first = next((p for p in model.parameters()), None)
        pol = plugin.mixed_precision_policy
        bf16 = first is not None and (first.dtype == torch.bfloat16 or (pol is not None and getattr(pol, "param_dtype", None) == torch.bfloat16))
        if not bf16:
            return f"{first.dtype if first is not None else 'no'} parameters without a bf16 mixed-precision policy"
        if getattr(plugin, "offloads_parameters", False):
            return "CPU offload of parameters"
        return ""

    def _wrap_torch_fsdp2(self, model, plugin, mesh):
        """The reference's FSDP2 path (utils/fsdp_utils.py:621-737): ``fully_shard`` per auto-wrapped module, then the root."""
        from torch.distributed.fsdp import CPUOffloadPolicy, MixedPrecisionPolicy, fully_shard

        kw = {"reshard_after_forward": plugin.effective_reshard_after_forward}
        pol = plugin.mixed_precision_policy
        if pol is not None:
            kw["mp_policy"] = pol if isinstance(pol, MixedPrecisionPolicy) else MixedPrecisionPolicy(
                param_dtype=getattr(pol, "param_dtype", None), reduce_dtype=getattr(pol, "reduce_dtype", None),
                output_dtype=getattr(pol, "output_dtype", None))
        if getattr(plugin, "offloads_parameters", False):
            kw["offload_policy"] = CPUOffloadPolicy()
        if mesh is not None:
            names = mesh.mesh_dim_names or ()
            dims = tuple(n for n in ("dp_replicate", "dp_shard_cp") if n in names or n == "dp_shard_cp" and ("dp_shard" in names or "cp" in names))
            if dims:
                kw["mesh"] = mesh[dims] if len(dims) > 1 else mesh[dims[0]]
        if plugin.activation_checkpointing:
            from .parallel.fsdp import apply_activation_checkpointing

            apply_activation_checkpointing(model, plugin)
        model = model.to(self.device)
        before = {name: p for name, p in model.named_parameters()}
        for sub in list(model.modules())[::-1]:
            if sub is not model and plugin.should_wrap(sub):
                fully_shard(sub, **kw)
        fully_shard(model, **kw)
        # fully_shard replaced every Parameter by a DTensor one: optimizers built on the old objects are re-pointed in prepare()
        model._b200_param_remap = _remap_after_sharding(before, dict(model.named_parameters()), model)
        model._b200_torch_fsdp2 = True
        return model

    def prepare_data_loader(self, data_loader: torch.utils.data.DataLoader, device_placement=None, slice_fn_for_dispatch=None):
        if getattr(data_loader, "_is_accelerate_prepared", False):
            if data_loader not in self._dataloaders:
                self._dataloaders.append(data_loader)
            return data_loader
        if device_placement is None:
            device_placement = self.device_placement
        device_mesh = None
        pc = self.parallelism_config
        if pc is not None and pc.total_size > 1 and (pc.tp_enabled or pc.cp_enabled or pc.sp_enabled):
            device_mesh = self.torch_device_mesh
        prepared = prepare_data_loader(
            data_loader, self.device, num_processes=self.num_processes, process_index=self.process_index,
            split_batches=self.split_batches, put_on_device=device_placement, rng_types=self.rng_types.copy(),
            dispatch_batches=self.dispatch_batches, even_batches=self.even_batches, slice_fn_for_dispatch=slice_fn_for_dispatch,
            use_seedable_sampler=self.use_seedable_sampler, data_seed=self.dataloader_config.data_seed, non_blocking=self.non_blocking,
            use_stateful_dataloader=self.use_stateful_dataloader, torch_device_mesh=device_mesh,
        )
        if pc is not None and pc.sp_enabled:
            from .parallel.ulysses import UlyssesDataLoaderAdapter

            prepared = UlyssesDataLoaderAdapter(prepared, self.torch_device_mesh["sp"], device=self.device)
        self._dataloaders.append(prepared)
        return prepared

    def prepare_optimizer(self, optimizer: torch.optim.Optimizer, device_placement=None):
        if getattr(optimizer, "_is_accelerate_prepared", False):
            if optimizer not in self._optimizers:
                self._optimizers.append(optimizer)
            return optimizer
        if device_placement is None:
            device_placement = self.device_placement
        scaler = None if self.is_fsdp2 and self.scaler is None else self.scaler
        if (hasattr(optimizer, "enable_update_in_backward") and self.distributed_type == DistributedType.NO
                and self.device.type == "cuda" and scaler is None):
            # single device: nothing has to be reduced first, so FusedAdamW may update each parameter inside backward
            gs = self.gradient_state
            optimizer.enable_update_in_backward(sync_gradients_fn=lambda: gs.sync_gradients)
        optimizer = AcceleratedOptimizer(optimizer, device_placement=device_placement, scaler=scaler)
        self._optimizers.append(optimizer)
        return optimizer

    def prepare_scheduler(self, scheduler):
        if getattr(scheduler, "_is_accelerate_prepared", False):
            if scheduler not in self._schedulers:
                self._schedulers.append(scheduler)
            return scheduler
        optimizer = self._optimizers
        for opt in self._optimizers:
            if getattr(scheduler, "optimizer", None) == opt.optimizer:
                optimizer = opt
                break
        scheduler = AcceleratedScheduler(scheduler, optimizer, step_with_optimizer=self.step_scheduler_with_optimizer, split_batches=self.split_batches)
        self._schedulers.append(scheduler)
        return scheduler

    def deepspeed_ulysses_dl_adapter(self, dl, model=None):
        """Wrap a dataloader so each batch is sharded along the sequence over the ``sp`` axis."""
        from .parallel.ulysses import UlyssesDataLoaderAdapter

        return UlyssesDataLoaderAdapter(dl, self.torch_device_mesh["sp"], device=self.device)

    # ================================================================== the training step
    def backward(self, loss, **kwargs):
        """``loss / gradient_accumulation_steps`` then backward (through the GradScaler for fp16)."""
        learning_rate = kwargs.pop("learning_rate", None)
        # (a DeepSpeed-style configuration runs on the native engines here, so the loss is divided for it too — the reference
        # leaves that to DeepSpeed's own engine)
        if self.gradient_accumulation_steps != 1:
            loss = loss / self.gradient_accumulation_steps
        if learning_rate is not None and self.has_lomo_optimizer:
            self.lomo_backward(loss, learning_rate)
        elif self.scaler is not None:
            self.scaler.scale(loss).backward(**kwargs)
        else:
            loss.backward(**kwargs)
        if getattr(self, "_config_clip_norm", None) is not None and self.sync_gradients:
            # `gradient_clipping` of a DeepSpeed-style configuration: DeepSpeed's engine clips at every step boundary on its own
            self._clip_from_config()

    def set_trigger(self):
        """Raise the cross-rank flag (e.g. "stop training") checked by :meth:`check_trigger`."""
        self.flag_tensor = torch.tensor(1, device=self.device)

    def check_trigger(self):
        """True on every rank if any rank called :meth:`set_trigger`; resets the flag."""
        if self.flag_tensor is None:
            self.flag_tensor = torch.tensor(0, device=self.device)
        flag_tensor = self.reduce(self.flag_tensor)
        if flag_tensor.item() >= 1:
            self.flag_tensor = torch.tensor(0, device=self.device)
            return True
        return False

    def unscale_gradients(self, optimizer=None):
        """Unscale fp16 gradients (no-op unless a GradScaler is active)."""
        if self.native_amp and self.mixed_precision == "fp16" and self.scaler is not None:
            if optimizer is None:
                optimizer = self._optimizers
            elif not isinstance(optimizer, (tuple, list)):
                optimizer = [optimizer]
            for opt in optimizer:
                while isinstance(opt, AcceleratedOptimizer):
                    opt = opt.optimizer
                self.scaler.unscale_(opt)

    def _clip_from_config(self):
        limit, self._config_clip_norm = self._config_clip_norm, None  # (so that the call below really clips)
        try:
            for model in self._models:
                params = [p for p in model.parameters() if p.requires_grad]
                if params:
                    self._last_config_clip_norm = self.clip_grad_norm_(params, limit)
        finally:
            self._config_clip_norm = limit

    def clip_grad_norm_(self, parameters, max_norm, norm_type=2):
        """Clip to ``max_norm`` and return the total norm (a device tensor; no host sync on the native path)."""
        if getattr(self, "_config_clip_norm", None) is not None:
            # the configuration already clipped this step's gradients (reference: under DeepSpeed this call only reports the norm)
            return getattr(self, "_last_config_clip_norm", None)
        parameters = [p for p in parameters]
        for model in self._models:
            if hasattr(model, "clip_grad_norm_") and not isinstance(model, torch.nn.parallel.DistributedDataParallel):
                mine = {id(p) for p in model.parameters()}
                if parameters and all(id(p) in mine for p in parameters):
                    self.unscale_gradients()
                    return model.clip_grad_norm_(max_norm, norm_type)
        self.unscale_gradients()
        self._notify_grad_consumer("gradient clipping")
        fused = [opt.optimizer for opt in self._optimizers if hasattr(opt.optimizer, "set_clip_coef")]
        if norm_type == 2 and self.device.type == "cuda" and len(fused) == len(self._optimizers) and len(fused) > 0 and self.scaler is None:
            from .optim import grad_global_norm

            norm, coef, _ = grad_global_norm(parameters, max_norm)
            for o in fused:
                o.set_clip_coef(coef)  # applied inside the AdamW kernel: no extra pass over the gradients
            return norm.reshape(())
        return torch.nn.utils.clip_grad_norm_(parameters, max_norm, norm_type=norm_type)

    def clip_grad_value_(self, parameters, clip_value):
        if self.distributed_type in (DistributedType.DEEPSPEED, DistributedType.FSDP):
            raise Exception("DeepSpeed and FSDP  do not support `clip_grad_value_`. Use `clip_grad_norm_` instead.")
        self.unscale_gradients()
        self._notify_grad_consumer("gradient clipping")
        torch.nn.utils.clip_grad_value_(parameters, clip_value)

    def _notify_grad_consumer(self, what: str):
        for opt in self._optimizers:
            inner = getattr(opt, "optimizer", opt)
            if hasattr(inner, "_note_grad_consumer"):
                inner._note_grad_consumer(what)

    # ================================================================== collectives for metrics
    def gather(self, tensor):
        return gather(tensor)

    def gather_for_metrics(self, input_data, use_gather_object: bool = False):
        """Gather and drop the samples the last batch duplicated to keep ranks even."""
        try:
            recursively_apply(lambda x: x, input_data, error_on_other_type=True)
            all_tensors = True
        except TypeError:
            all_tensors = False
        use_gather_object = use_gather_object or not all_tensors
        data = gather_object(input_data) if use_gather_object else self.gather(input_data)
        try:
            if self.gradient_state.end_of_dataloader:
                if self.gradient_state.remainder == -1:
                    logger.info("The used dataset had no length, returning gathered tensors. You should drop the remainder yourself.")
                    return data
                elif self.gradient_state.remainder > 0:
                    def _adjust_samples(tensor):
                        return tensor[: self.gradient_state.remainder]

                    if use_gather_object:
                        return _adjust_samples(data)
                    return recursively_apply(_adjust_samples, data)
                return data
            return data
        except Exception:
            return data

    def reduce(self, tensor, reduction="sum", scale=1.0):
        return reduce(tensor, reduction, scale)

    def pad_across_processes(self, tensor, dim=0, pad_index=0, pad_first=False):
        return pad_across_processes(tensor, dim=dim, pad_index=pad_index, pad_first=pad_first)

    # ================================================================== model utilities
    def unwrap_model(self, model, keep_fp32_wrapper: bool = True, keep_torch_compile: bool = True):
        return extract_model_from_parallel(model, keep_fp32_wrapper, keep_torch_compile)

    def verify_device_map(self, model: torch.nn.Module) -> bool:
        """True if ``model`` was dispatched over several devices with a ``device_map``."""
        for m in model.modules():
            if hasattr(m, "hf_device_map") and len(m.hf_device_map) > 1:
                return True
        return False

    def lomo_backward(self, loss: torch.Tensor, learning_rate: float) -> None:
        if learning_rate is None:
            raise ValueError("A learning rate must be passed in order to call backward pass with LOMO optimizers.")
        for optimizer in self._optimizers:
            if hasattr(optimizer.optimizer, "fused_backward"):
                optimizer.optimizer.fused_backward(loss, learning_rate)
                return
        raise ValueError("Backward pass not properly called on LOMO optimizers. Are you sure you passed a LOMO optimizer in `accelerator.prepare`?")

    def get_state_dict(self, model, unwrap: bool = True):
        """Full (un-sharded) state dict of ``model`` on every rank that needs it (CPU tensors under sharding)."""
        if hasattr(model, "full_state_dict"):
            return model.full_state_dict()
        if unwrap:
            model = self.unwrap_model(model)
        return model.state_dict()

    def save(self, obj, f, safe_serialization: bool = False):
        save(obj, f, save_on_each_node=self.project_configuration.save_on_each_node, safe_serialization=safe_serialization)

    def save_model(self, model: torch.nn.Module, save_directory: Union[str, os.PathLike], max_shard_size: Union[int, str] = "10GB",
                   safe_serialization: bool = True):
        """HF-style export: ``model.safetensors`` or ``model-0000X-of-0000Y.safetensors`` + index json."""
        if os.path.isfile(save_directory):
            logger.error(f"Provided path ({save_directory}) should be a directory, not a file")
            return
        os.makedirs(save_directory, exist_ok=True)
        if any(getattr(m, "_hf_hook", None) is not None and getattr(m._hf_hook, "offload", False) for m in model.modules()):
            from .utils.modeling import get_state_dict_offloaded_model

            state_dict = get_state_dict_offloaded_model(model)
        else:
            if any(p.device == torch.device("meta") for p in model.parameters()):
                raise RuntimeError("You can't save the model since some parameters are on the meta device.")
            state_dict = self.get_state_dict(model)
        if safe_serialization:
            state_dict = clean_state_dict_for_safetensors(state_dict)
        weights_name = SAFE_WEIGHTS_NAME if safe_serialization else WEIGHTS_NAME
        filename_pattern = SAFE_WEIGHTS_PATTERN_NAME if safe_serialization else WEIGHTS_PATTERN_NAME
        from .utils.modeling import shard_checkpoint

        shards, index = shard_checkpoint(state_dict, max_shard_size=max_shard_size, weights_name=weights_name)
        # remove stale shard files of an earlier, differently sharded save
        for filename in os.listdir(save_directory):
            full_filename = os.path.join(save_directory, filename)
            weights_no_suffix = weights_name.replace(".bin", "").replace(".safetensors", "")
            filename_no_suffix = filename.replace(".bin", "").replace(".safetensors", "")
            reg = re.compile(r"(.*?)-\d{5}-of-\d{5}")
            if filename.startswith(weights_no_suffix) and os.path.isfile(full_filename) and filename not in shards.keys() and reg.fullmatch(filename_no_suffix) is not None and PartialState().is_main_process:
                os.remove(full_filename)
        for shard_file, shard in shards.items():
            self.save(shard, os.path.join(save_directory, shard_file), safe_serialization=safe_serialization)
        if index is None:
            logger.info(f"Model weights saved in {os.path.join(save_directory, weights_name)}")
        else:
            save_index_file = SAFE_WEIGHTS_INDEX_NAME if safe_serialization else WEIGHTS_INDEX_NAME
            save_index_file = os.path.join(save_directory, save_index_file)
            if self.is_main_process:
                with open(save_index_file, "w", encoding="utf-8") as f:
                    f.write(json.dumps(index, indent=2, sort_keys=True) + "\n")
            logger.info(f"The model is bigger than the maximum size per checkpoint ({max_shard_size}) and is going to be split in {len(shards)} checkpoint shards. You can find where each parameters has been saved in the index located at {save_index_file}.")

    # ================================================================== checkpoint hooks
    def register_save_state_pre_hook(self, hook: Callable[..., None]):
        """``hook(models, weights, output_dir)`` runs inside ``save_state`` before the files are written."""
        handle = torch.utils.hooks.RemovableHandle(self._save_model_state_pre_hook)
        self._save_model_state_pre_hook[handle.id] = hook
        return handle

    def register_load_state_pre_hook(self, hook: Callable[..., None]):
        """``hook(models, input_dir)`` runs inside ``load_state`` before the files are read."""
        handle = torch.utils.hooks.RemovableHandle(self._load_model_state_pre_hook)
        self._load_model_state_pre_hook[handle.id] = hook
        return handle

    def register_for_checkpointing(self, *objects):
        """Track extra objects (with ``state_dict`` / ``load_state_dict``) in ``save_state`` / ``load_state``."""
        rejected = [obj for obj in objects if not (hasattr(obj, "state_dict") and hasattr(obj, "load_state_dict"))]
        if rejected:
            listing = "".join(f"\n\t- Item at index {i}, `{get_pretty_name(obj)}`" for i, obj in enumerate(rejected))
            raise ValueError("All `objects` must include a `state_dict` and `load_state_dict` function to be stored. The following inputs "
                             f"are invalid:{listing}")
        self._custom_objects.extend(objects)

    # ================================================================== save_state / load_state
    @staticmethod
    def _checkpoint_number(folder: str) -> int:
        """``.../checkpoint_12`` -> 12 (the last run of digits of the last path component)."""
        return int(re.findall(r"(\d+)", os.path.basename(os.path.normpath(folder)))[-1])

    def _numbered_checkpoints(self) -> list:
        """Existing ``<project_dir>/checkpoints/*`` folders, oldest first."""
        root = os.path.join(self.project_dir, "checkpoints")
        return sorted((os.path.join(root, name) for name in os.listdir(root)), key=self._checkpoint_number)

    def _claim_checkpoint_dir(self, output_dir: Optional[str]) -> str:
        """Where this ``save_state`` writes.  Under automatic naming: ``checkpoints/checkpoint_<iteration>``, after the oldest
        checkpoints were removed to respect ``total_limit``."""
        cfg = self.project_configuration
        if not cfg.automatic_checkpoint_naming:
            os.makedirs(output_dir, exist_ok=True)
            return output_dir
        os.makedirs(os.path.join(self.project_dir, "checkpoints"), exist_ok=True)
        existing = self._numbered_checkpoints()
        surplus = 0 if cfg.total_limit is None else len(existing) + 1 - cfg.total_limit
        if surplus > 0 and self.is_main_process:
            logger.warning(f"Deleting {surplus} checkpoints to make room for new checkpoint.")
            for folder in existing[:surplus]:
                shutil.rmtree(folder, ignore_errors=True)
        target = os.path.join(self.project_dir, "checkpoints", f"checkpoint_{self.save_iteration}")
        if os.path.exists(target):
            raise ValueError(f"Checkpoint directory {target} ({self.save_iteration}) already exists. Please manually override "
                             "`self.save_iteration` with what iteration to start with.")
        self.wait_for_everyone()
        os.makedirs(target, exist_ok=True)
        return target

    def _sharded_optimizer_io(self, opt, index: int, folder: str, save: bool) -> bool:
        """Optimizers whose state is sharded go through their engine (native) or DCP (torch FSDP2).  True when handled."""
        from .utils.fsdp_utils import load_fsdp_optimizer, save_fsdp_optimizer

        engine = getattr(opt, "_engine", None)
        method = "save_sharded_optimizer" if save else "load_sharded_optimizer"
        if engine is not None and hasattr(engine, method):
            getattr(engine, method)(opt, folder, index, self.state.fsdp_plugin)
            return True
        torch_fsdp2 = [m for m in self._models if getattr(m, "_b200_torch_fsdp2", False)]
        if torch_fsdp2:
            (save_fsdp_optimizer if save else load_fsdp_optimizer)(self.state.fsdp_plugin, self, opt, torch_fsdp2[0], folder, index)
            return True
        return False

    @staticmethod
    def _is_sharded_model(model) -> bool:
        return hasattr(model, "save_sharded_state") or getattr(model, "_b200_torch_fsdp2", False)

    def save_state(self, output_dir: Optional[str] = None, safe_serialization: bool = True, **save_model_func_kwargs):
        """Checkpoint everything ``prepare`` saw (plus registered objects and RNG state) under ``output_dir``."""
        from .utils.fsdp_utils import save_fsdp_model

        output_dir = self._claim_checkpoint_dir(output_dir)
        logger.info(f"Saving current state to {output_dir}")
        weights = []
        for i, model in enumerate(self._models):
            if self._is_sharded_model(model):
                save_fsdp_model(self.state.fsdp_plugin, self, model, output_dir, i)
                logger.info(f"Sharded model saved to output dir {output_dir}")
            else:
                weights.append(self.get_state_dict(model, unwrap=False))
        plain_optimizers = [opt for i, opt in enumerate(self._optimizers) if not self._sharded_optimizer_io(opt, i, output_dir, save=True)]
        for hook in self._save_model_state_pre_hook.values():
            hook(self._models, weights, output_dir)
        per_node = self.project_configuration.save_on_each_node
        save_location = save_accelerator_state(output_dir, weights, plain_optimizers, list(self._schedulers), self._dataloaders,
                                               self.state.process_index, self.step, self.scaler, save_on_each_node=per_node,
                                               safe_serialization=safe_serialization)
        for i, obj in enumerate(self._custom_objects):
            save_custom_state(obj, output_dir, i, save_on_each_node=per_node)
        self.project_configuration.iteration += 1
        return save_location

    def _find_checkpoint_dir(self, input_dir: Optional[str]) -> str:
        if input_dir is not None:
            input_dir = os.path.expanduser(input_dir)
            if not os.path.isdir(input_dir):
                raise ValueError(f"Tried to find {input_dir} but folder does not exist")
            return input_dir
        if not self.project_configuration.automatic_checkpoint_naming:
            raise ValueError("No input_dir provided and automatic checkpoint naming is disabled.")
        return self._numbered_checkpoints()[-1]

    def load_state(self, input_dir: Optional[str] = None, load_kwargs: Optional[dict] = None, **load_model_func_kwargs):
        """Restore what :meth:`save_state` wrote (latest ``checkpoint_N`` under automatic naming)."""
        from .utils.fsdp_utils import load_fsdp_model

        input_dir = self._find_checkpoint_dir(input_dir)
        logger.info(f"Loading states from {input_dir}")
        plain_models = []
        for i, model in enumerate(self._models):
            if self._is_sharded_model(model) or hasattr(model, "load_sharded_state"):
                load_fsdp_model(self.state.fsdp_plugin, self, model, input_dir, i)
                logger.info(f"Sharded model loaded from input dir {input_dir}")
            else:
                plain_models.append(model)
        plain_optimizers = [opt for i, opt in enumerate(self._optimizers) if not self._sharded_optimizer_io(opt, i, input_dir, save=False)]
        for hook in self._load_model_state_pre_hook.values():
            hook(plain_models, input_dir)
        map_location = load_model_func_kwargs.pop("map_location", None)
        if map_location is None:
            on_gpus = self.num_processes > 1 and self.multi_device and self.distributed_type != DistributedType.MULTI_CPU
            map_location = "on_device" if on_gpus else "cpu"
        restored = load_accelerator_state(input_dir, plain_models, plain_optimizers, list(self._schedulers), self._dataloaders,
                                          self.state.process_index, self.scaler, map_location, load_kwargs, **load_model_func_kwargs)
        self.step = restored.get("step", self.step)
        found = sum(1 for name in os.listdir(input_dir) if re.fullmatch(r"custom_checkpoint_\d+\.pkl", name))
        if found != len(self._custom_objects):
            raise RuntimeError(
                f"Number of custom checkpoints in folder {input_dir} does not match the number of registered objects:"
                f"\n\tFound checkpoints: {found}\n\tRegistered objects: {len(self._custom_objects)}\n"
                "Please make sure to only load checkpoints from folders that were created with the same set of registered objects,"
                "or avoid using `custom_checkpoint` in the filename for files in that same directory and load them in manually."
            )
        logger.info(f"Loading in {found} custom states")
        for index, obj in enumerate(self._custom_objects):
            load_custom_state(obj, input_dir, index)

    def free_memory(self, *objects):
        """Drop every reference the accelerator keeps and empty the device cache."""
        for model in self._models:
            if hasattr(model, "release"):
                model.release()
        if hasattr(self, "deepspeed_engine_wrapped"):
            self.deepspeed_engine_wrapped = None
        objects = release_memory(*objects)
        self._schedulers = []
        self._optimizers = []
        self._models = []
        self._dataloaders = []
        self.step = 0
        return objects

    def clear(self, *objects):
        return self.free_memory(*objects)

    def _get_named_parameters(self, *args, drop_refs=False):
        named_parameters = {}
        for obj in args:
            if isinstance(obj, torch.nn.Module):
                obj = extract_model_from_parallel(obj)
                named_parameters.update({n: p for n, p in obj.named_parameters()})
        return named_parameters

    def _get_devices(self, *args):
        model_device = optimizer_device = None
        for obj in args:
            if isinstance(obj, torch.nn.Module):
                for param in obj.parameters():
                    model_device = param.device
                    break
            if isinstance(obj, torch.optim.Optimizer):
                for param_group in obj.param_groups:
                    if len(param_group["params"]) > 0:
                        optimizer_device = param_group["params"][0].device
                        break
        return (model_device, optimizer_device)

    # ================================================================== contexts
    def _get_autocast_context(self, autocast_handler: AutocastKwargs = None):
        if autocast_handler is None:
            autocast_handler = self.autocast_handler
        kwargs = autocast_handler.to_kwargs() if autocast_handler is not None else {}
        dtype = self._mixed_precision_dtype()
        if not self.native_amp or dtype is None:
            return contextlib.nullcontext()
        device_type = "cuda" if self.device.type == "cuda" else "cpu"
        if device_type == "cpu" and dtype == torch.float16:
            dtype = torch.bfloat16
        return torch.autocast(device_type=device_type, dtype=dtype, **kwargs)

    @contextmanager
    def autocast(self, autocast_handler: AutocastKwargs = None):
        """Run the block under the accelerator's mixed-precision autocast."""
        ctx = self._get_autocast_context(autocast_handler)
        ctx.__enter__()
        try:
            yield
        finally:
            ctx.__exit__(*sys.exc_info())

    @contextmanager
    def maybe_context_parallel(self, buffers=None, buffer_seq_dims=None, no_restore_buffers=None):
        """Shard ``buffers`` along their sequence dims over the ``cp`` axis for the duration of the step."""
        pc = self.parallelism_config
        if pc is None or not pc.cp_enabled:
            logger.warning_once("Context parallelism is not enabled: `maybe_context_parallel` is a no-op.")
            yield
            return
        from .parallel.context_parallel import context_parallel_buffers

        with context_parallel_buffers(self.torch_device_mesh["cp"], buffers or [], buffer_seq_dims or [], no_restore_buffers or set()):
            yield

    @contextmanager
    def profile(self, profile_handler: Optional[ProfileKwargs] = None):
        """``torch.profiler`` context; on exit writes ``profile_{rank}.json`` under ``output_trace_dir``."""
        profile_handler = profile_handler or self.profile_handler or ProfileKwargs()
        with profile_handler.build() as profiler:
            yield profiler
        if profile_handler.output_trace_dir is None:
            return
        os.makedirs(profile_handler.output_trace_dir, exist_ok=True)
        profiler.export_chrome_trace(os.path.join(profile_handler.output_trace_dir, PROFILE_PATTERN_NAME.format(suffix=self.process_index)))
        self.wait_for_everyone()

    def skip_first_batches(self, dataloader, num_batches: int = 0):
        return skip_first_batches(dataloader, num_batches=num_batches)

    # ================================================================== trackers
    def init_trackers(self, project_name: str, config: Optional[dict] = None, init_kwargs: Optional[dict] = None):
        from .tracking import LOGGER_TYPE_TO_CLASS, GeneralTracker

        init_kwargs = init_kwargs or {}
        for tracker in self.log_with:
            if issubclass(type(tracker), GeneralTracker):
                self.trackers.append(tracker)
            else:
                tracker_init = LOGGER_TYPE_TO_CLASS[str(tracker)]
                if tracker_init.requires_logging_directory:
                    self.trackers.append(tracker_init(project_name, self.logging_dir, **init_kwargs.get(str(tracker), {})))
                else:
                    self.trackers.append(tracker_init(project_name, **init_kwargs.get(str(tracker), {})))
        for tracker in self.trackers:
            tracker.start()
        if config is not None:
            for tracker in self.trackers:
                tracker.store_init_configuration(config)

    def get_tracker(self, name: str, unwrap: bool = False):
        from .tracking import GeneralTracker

        if len(self.trackers) > 0:
            for tracker in self.trackers:
                if tracker.name == name:
                    return tracker.tracker if unwrap else tracker
            raise ValueError(f"{name} is not an available tracker stored inside the `Accelerator`.")
        return GeneralTracker(_blank=True)

    def log(self, values: dict, step: Optional[int] = None, log_kwargs: Optional[dict] = None):
        log_kwargs = log_kwargs or {}
        for tracker in self.trackers:
            tracker.log(values, step=step, **log_kwargs.get(tracker.name, {}))

    def end_training(self):
        """Finish trackers and tear down the process group."""
        for tracker in self.trackers:
            tracker.finish()
        self.state.destroy_process_group()

    def __deepcopy__(self, memo):
        logger.info("Deep copying the `Accelerator` object, note that this will point to the same original object.")
        return self


def get_pretty_name(obj):
    from .utils.other import get_pretty_name as _g

    return _g(obj)
