"""Process / device state singletons: ``PartialState``, ``AcceleratorState``, ``GradientState``.

Same contract as the reference (src/accelerate/state.py): three Borg classes whose instances share
one ``__dict__`` (:161,:177), so every part of the framework sees the same rank / device / plugin
facts; ``_reset_state`` empties them for tests (:338-346).  What is different here:

* one process per GPU, NCCL for the control plane, gloo for CPU tests — the vendor ladder of
  ``_prepare_backend`` (reference :753-812) collapses to ``nccl | gloo | mpi``;
* ``PartialState`` also owns the :class:`~.parallel.symm.SymmetricWorkspace` (lazily), the
  peer-mapped memory the fused collectives run on.
"""

from __future__ import annotations

import logging
import os
import threading
import weakref
from contextlib import contextmanager
from functools import partial
from typing import Any, Callable, Optional

import torch

from .utils.dataclasses import (
    DistributedType,
    DynamoBackend,
    GradientAccumulationPlugin,
    PrecisionType,
    TorchDynamoPlugin,
)
from .utils.environment import (
    check_cuda_fp8_capability,
    check_cuda_p2p_ib_support,
    get_cpu_distributed_information,
    get_int_from_env,
    parse_choice_from_env,
    parse_flag_from_env,
    set_numa_affinity,
)

logger = logging.getLogger(__name__)


def is_initialized() -> bool:
    """True once an ``AcceleratorState`` has been constructed (and not reset)."""
    return AcceleratorState._shared_state != {}


def do_nothing(*args, **kwargs):
    return None


class ThreadLocalSharedDict(threading.local):
    """Descriptor giving each thread its own shared-state dict (kept for API parity; reference :90-118)."""

    def __init__(self, thread_local: bool = False):
        self._storage = {}

    def __get__(self, obj, objtype=None):
        return self._storage

    def __set__(self, obj, value):
        self._storage = value


SharedDict = dict


class PartialState:
    """Rank / world / device facts plus process-control helpers.

    Attributes: ``device``, ``distributed_type``, ``num_processes``, ``process_index``,
    ``local_process_index``, ``backend``, ``debug``, ``fork_launched``, ``is_main_process``,
    ``is_local_main_process``, ``is_last_process``.
    """

    _shared_state = SharedDict()
    _known_attrs = [
        "_cpu", "_mixed_precision", "_shared_state", "backend", "debug", "device", "distributed_type",
        "fork_launched", "local_process_index", "num_processes", "process_index",
    ]

    def __init__(self, cpu: bool = False, **kwargs):
        self.__dict__ = self._shared_state
        if self.initialized:
            return
        self._cpu = cpu
        self.backend = None
        env_device = os.environ.get("ACCELERATE_TORCH_DEVICE", None)
        self.device = torch.device(env_device) if env_device is not None else None
        self.debug = parse_flag_from_env("ACCELERATE_DEBUG_MODE")
        use_cpu = cpu or parse_flag_from_env("ACCELERATE_USE_CPU") or not torch.cuda.is_available()
        self._cpu = use_cpu
        original_backend = kwargs.pop("backend", None)
        backend, distributed_type = self._prepare_backend(use_cpu, original_backend)
        if original_backend is not None and backend != original_backend:
            raise ValueError(f"Your assigned backend {original_backend} is not available, please use {backend}")
        self.backend = backend
        self.distributed_type = distributed_type
        self._launch_distributed_type = distributed_type  # what the launch gave us, before any engine plugin promotes it
        self._symm_workspace = None

        if distributed_type == DistributedType.NO:
            self.num_processes = 1
            self.process_index = 0
            self.local_process_index = 0
        else:
            if use_cpu:
                info = get_cpu_distributed_information()
                os.environ.setdefault("RANK", str(info.rank))
                os.environ.setdefault("WORLD_SIZE", str(info.world_size))
                os.environ.setdefault("LOCAL_RANK", str(info.local_rank))
                os.environ.setdefault("LOCAL_WORLD_SIZE", str(info.local_world_size))
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29500")
            if not torch.distributed.is_initialized():
                self._init_process_group(use_cpu, kwargs)
            self.num_processes = torch.distributed.get_world_size()
            self.process_index = torch.distributed.get_rank()
            self.local_process_index = int(os.environ.get("LOCAL_RANK", -1)) if "LOCAL_RANK" in os.environ else self.process_index
        self.set_device()
        if parse_flag_from_env("ACCELERATE_CPU_AFFINITY", False):
            set_numa_affinity(self.local_process_index)
        if self.device.type == "cuda" and not check_cuda_p2p_ib_support():
            if "NCCL_P2P_DISABLE" not in os.environ or "NCCL_IB_DISABLE" not in os.environ:
                raise NotImplementedError(
                    "This GPU/driver pair does not support peer-to-peer or InfiniBand transfers; set "
                    "NCCL_P2P_DISABLE=1 NCCL_IB_DISABLE=1 (the `launch` command does it automatically)."
                )
        self.fork_launched = parse_flag_from_env("FORK_LAUNCHED", 0)

    # ------------------------------------------------------------------ bootstrap
    def _prepare_backend(self, cpu: bool, backend: Optional[str]):
        """(backend, DistributedType) from the environment: nccl on CUDA, gloo (or mpi) on CPU."""
        distributed_type = None
        if int(os.environ.get("LOCAL_RANK", -1)) != -1 and not cpu and torch.cuda.is_available():
            distributed_type = DistributedType.MULTI_GPU
            if backend is None:
                backend = "nccl"
        if distributed_type is None and (
            int(os.environ.get("LOCAL_RANK", -1)) != -1
            or get_int_from_env(["PMI_SIZE", "OMPI_COMM_WORLD_SIZE", "MV2_COMM_WORLD_SIZE", "WORLD_SIZE"], 1) > 1
        ):
            if not cpu and torch.cuda.is_available():
                distributed_type = DistributedType.MULTI_GPU
                backend = backend or "nccl"
            else:
                distributed_type = DistributedType.MULTI_CPU
                if backend in (None, "nccl"):
                    backend = "mpi" if (torch.distributed.is_mpi_available() and "OMPI_COMM_WORLD_SIZE" in os.environ) else "gloo"
        if distributed_type is None:
            distributed_type = DistributedType.NO
        return backend, distributed_type

    def _init_process_group(self, use_cpu: bool, kwargs: dict):
        kwargs = dict(kwargs)
        backend = self.backend
        # sharded state dicts with CPU offload need CPU collectives next to the CUDA ones
        if backend == "nccl" and (
            (parse_flag_from_env("ACCELERATE_USE_FSDP") and (
                parse_flag_from_env("FSDP_OFFLOAD_PARAMS") or os.environ.get("FSDP_STATE_DICT_TYPE", "") == "FULL_STATE_DICT"))
        ):
            backend = "cuda:nccl,cpu:gloo"
        rdv_file = os.environ.get("ACCELERATE_DEBUG_RDV_FILE", None)
        if rdv_file is not None and "init_method" not in kwargs and "store" not in kwargs:
            # debug_launcher: rendezvous through a file, no TCP port needed
            kwargs["store"] = torch.distributed.FileStore(rdv_file, int(os.environ["WORLD_SIZE"]))
            kwargs["rank"] = int(os.environ["RANK"])
            kwargs["world_size"] = int(os.environ["WORLD_SIZE"])
        attempt = int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") or 0)
        if (attempt > 0 and "gloo" in str(backend) and "init_method" not in kwargs and "store" not in kwargs
                and os.environ.get("LOCAL_WORLD_SIZE") == os.environ.get("WORLD_SIZE")):
            # Restarted worker group (torchelastic --max_restarts) on one node with gloo: in this torch build gloo's
            # connectFullMesh finds the addresses the dead attempt left in the agent's store ("Connection refused").  A
            # fresh file store per attempt has no such leftovers.  (NCCL groups do not go through these keys.)
            import tempfile

            run_id = os.environ.get("TORCHELASTIC_RUN_ID", "none")
            path = os.path.join(tempfile.gettempdir(), f"b200_rdv_{run_id}_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}_{attempt}")
            kwargs["store"] = torch.distributed.FileStore(path, int(os.environ["WORLD_SIZE"]))
            kwargs["rank"] = int(os.environ["RANK"])
            kwargs["world_size"] = int(os.environ["WORLD_SIZE"])
        if not use_cpu and torch.cuda.is_available():
            local_rank = int(os.environ.get("LOCAL_RANK", 0))
            torch.cuda.set_device(local_rank % max(torch.cuda.device_count(), 1))
            kwargs.setdefault("device_id", torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1)))
        kwargs.pop("backend", None)
        try:
            torch.distributed.init_process_group(backend=backend, **kwargs)
        except TypeError:
            kwargs.pop("device_id", None)
            torch.distributed.init_process_group(backend=backend, **kwargs)
        self.backend = backend

    def set_device(self):
        """Bind this process to ``cuda:{local_rank % device_count}`` (or cpu)."""
        if self.device is not None:
            return
        if self._cpu or not torch.cuda.is_available():
            self.device = torch.device("cpu")
            return
        n = torch.cuda.device_count()
        idx = self.local_process_index % n if self.distributed_type != DistributedType.NO else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        torch.cuda.set_device(self.device)

    # ------------------------------------------------------------------ basic facts
    def __repr__(self) -> str:
        return (
            f"Distributed environment: {self.distributed_type}{('  Backend: ' + self.backend) if self.backend else ''}\n"
            f"Num processes: {self.num_processes}\n"
            f"Process index: {self.process_index}\n"
            f"Local process index: {self.local_process_index}\n"
            f"Device: {self.device}\n"
        )

    @staticmethod
    def _reset_state():
        PartialState._shared_state.clear()

    @property
    def initialized(self) -> bool:
        return self._shared_state != {}

    @property
    def use_distributed(self):
        return self.distributed_type != DistributedType.NO and self.num_processes > 1

    @property
    def is_last_process(self) -> bool:
        return self.process_index == self.num_processes - 1

    @property
    def is_main_process(self) -> bool:
        return self.process_index == 0

    @property
    def is_local_main_process(self) -> bool:
        return self.local_process_index == 0

    @property
    def default_device(self) -> torch.device:
        if torch.cuda.is_available() and not self._cpu:
            return torch.device("cuda")
        return torch.device("cpu")

    @property
    def symmetric_workspace(self):
        """Peer-mapped workspace of the fused collectives, created on first use (CUDA multi-process only)."""
        if self._symm_workspace is None and self.device.type == "cuda" and self.num_processes > 1:
            from .parallel.symm import SymmetricWorkspace

            self._symm_workspace = SymmetricWorkspace.create_default(self)
        return self._symm_workspace

    # ------------------------------------------------------------------ process control
    def wait_for_everyone(self):
        """Barrier over all processes; a no-op on a single process."""
        if self.distributed_type != DistributedType.NO and torch.distributed.is_initialized():
            if self.device.type == "cuda" and "nccl" in str(self.backend):
                torch.distributed.barrier(device_ids=[self.device.index])
            else:
                torch.distributed.barrier()

    def _goes_first(self, is_main: bool):
        if not is_main:
            self.wait_for_everyone()
        yield
        if is_main:
            self.wait_for_everyone()

    @contextmanager
    def split_between_processes(self, inputs, apply_padding: bool = False):
        """Yield this rank's contiguous slice of ``inputs`` (list / tuple / dict of equal-length values / tensor).

        The first ``len % world`` ranks get one extra element; with ``apply_padding`` every rank's
        slice is padded with the last element to the longest slice so the results can be gathered.
        """
        if self.num_processes == 1:
            yield inputs
            return
        length = len(inputs)
        if isinstance(inputs, dict):
            length = len(inputs[list(inputs.keys())[0]])
            if not all(len(v) == length for v in inputs.values()):
                raise ValueError("All values in the dictionary must have the same length")
        per, extras = divmod(length, self.num_processes)
        start = self.process_index * per + min(self.process_index, extras)
        end = start + per + (1 if self.process_index < extras else 0)

        def _slice(obj):
            if isinstance(obj, dict):
                return {k: _slice(v) for k, v in obj.items()}
            if isinstance(obj, (list, tuple, torch.Tensor)):
                if start >= len(obj):
                    result = obj[-1:]
                else:
                    result = obj[start:end]
                if apply_padding:
                    target = per + (1 if extras > 0 else 0)
                    missing = target - len(result)
                    if missing > 0:
                        if isinstance(result, torch.Tensor):
                            from .utils.operations import pad_input_tensors

                            result = torch.cat([result, result[-1:].expand(missing, *result.shape[1:])], dim=0) if len(result) else result
                        else:
                            result = list(result) + [result[-1]] * missing
                return result
            return obj

        yield _slice(inputs)

    @contextmanager
    def main_process_first(self):
        yield from self._goes_first(self.is_main_process)

    @contextmanager
    def local_main_process_first(self):
        yield from self._goes_first(self.is_local_main_process)

    def on_main_process(self, function: Callable[..., Any] = None):
        if not self.initialized:
            raise ValueError("The `PartialState` or `Accelerator` must be initialized before calling this function.")
        if self.is_main_process or not self.use_distributed:
            return function
        return do_nothing

    def on_local_main_process(self, function: Callable[..., Any] = None):
        if self.is_local_main_process or not self.use_distributed:
            return function
        return do_nothing

    def on_last_process(self, function: Callable[..., Any]):
        if self.is_last_process or not self.use_distributed:
            return function
        return do_nothing

    def on_process(self, function: Callable[..., Any] = None, process_index: int = None):
        if function is None:
            return partial(self.on_process, process_index=process_index)
        if self.process_index == process_index or not self.use_distributed:
            return function
        return do_nothing

    def on_local_process(self, function: Callable[..., Any] = None, local_process_index: int = None):
        if function is None:
            return partial(self.on_local_process, local_process_index=local_process_index)
        if self.local_process_index == local_process_index or not self.use_distributed:
            return function
        return do_nothing

    def print(self, *args, **kwargs):
        if self.is_local_main_process:
            print(*args, **kwargs)

    def destroy_process_group(self, group=None):
        """Tear down the process group (and the symmetric workspace that depends on it)."""
        ws = self.__dict__.get("_symm_workspace", None)
        if ws is not None:
            try:
                ws.close()
            except Exception:  # pragma: no cover
                pass
            self._symm_workspace = None
        if self.fork_launched and group is None:
            return
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.destroy_process_group(group)

    def __getattr__(self, name: str):
        if name in self._known_attrs:
            raise AttributeError(
                f"`PartialState` object has no attribute `{name}`. This happens if `PartialState._reset_state()` "
                "was called and an `Accelerator` or `PartialState` was not reinitialized."
            )
        raise AttributeError(f"'PartialState' object has no attribute '{name}'")


def _from_partial_state(name: str, call: bool = False):
    """Class attribute of ``AcceleratorState`` that defers to the ``PartialState`` singleton: a read-only property, or (``call``)
    a method with the same signature."""
    if not call:
        return property(lambda self: getattr(PartialState(), name))

    def method(self, *args, **kwargs):
        return getattr(PartialState(), name)(*args, **kwargs)

    method.__name__ = name
    return method


class AcceleratorState:
    """Everything in ``PartialState`` plus the training configuration: mixed precision, plugins,
    ``ParallelismConfig`` and the device mesh (reference src/accelerate/state.py:863-1240)."""

    _shared_state = SharedDict()
    _known_attrs = PartialState._known_attrs + [
        "deepspeed_plugin", "use_ipex", "fsdp_plugin", "megatron_lm_plugin", "dynamo_plugin", "parallelism_config", "device_mesh",
    ]

    def __init__(self, mixed_precision: str = None, cpu: bool = False, dynamo_plugin=None, deepspeed_plugin=None,
                 fsdp_plugin=None, torch_tp_plugin=None, megatron_lm_plugin=None, parallelism_config=None,
                 _from_accelerator: bool = False, **kwargs):
        self.__dict__ = self._shared_state
        if parse_flag_from_env("ACCELERATE_USE_CPU"):
            cpu = True
        if PartialState._shared_state == {}:
            PartialState(cpu, **kwargs)
        self.__dict__.update(PartialState._shared_state)
        self._check_initialized(mixed_precision, cpu)
        if self.initialized:
            return
        self.deepspeed_plugins = None
        self.use_ipex = None
        self.torch_tp_plugin = torch_tp_plugin
        self.parallelism_config = parallelism_config
        self.device_mesh = None
        mixed_precision = parse_choice_from_env("ACCELERATE_MIXED_PRECISION", "no") if mixed_precision is None else mixed_precision.lower()
        if mixed_precision not in PrecisionType.list():
            raise ValueError(f"mixed_precision must be one of {PrecisionType.list()}, got {mixed_precision}")
        if mixed_precision == "fp8" and self.device.type == "cuda" and not check_cuda_fp8_capability():
            logger.warning("fp8 needs compute capability >= 8.9; falling back to fp16 mixed precision.")
            mixed_precision = "fp16"
        self.dynamo_plugin = dynamo_plugin if dynamo_plugin is not None else TorchDynamoPlugin()
        if not _from_accelerator:
            raise ValueError("Please make sure to properly initialize your accelerator via `accelerator = Accelerator()` "
                             "before using any functionality from the `accelerate` library.")
        self._mixed_precision = mixed_precision
        self.fsdp_plugin = None
        self.megatron_lm_plugin = None

        cp_or_sp = parallelism_config is not None and (parallelism_config.cp_enabled or parallelism_config.sp_enabled)
        # start from the launch's own type: an earlier AcceleratorState (reset since) may have left its promotion on PartialState
        self.distributed_type = PartialState._shared_state.get("_launch_distributed_type", self.distributed_type)
        if self.distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
            # (the reference refuses DeepSpeed on CPU; here a ZeRO configuration runs on the native sharded engine, whose gloo
            # path works on CPU too — that is how the DeepSpeed-style scripts are tested without GPUs)
            if parse_flag_from_env("ACCELERATE_USE_DEEPSPEED") or deepspeed_plugin is not None:
                self.distributed_type = DistributedType.DEEPSPEED
                if not isinstance(deepspeed_plugin, dict) and deepspeed_plugin is not None:
                    deepspeed_plugin = {"default": deepspeed_plugin}
                self.deepspeed_plugins = deepspeed_plugin
                if deepspeed_plugin:
                    chosen = next(iter(deepspeed_plugin.values()))
                    # a configuration that states its own precision decides when none was asked for (the native engines need
                    # to know it; the reference leaves precision to DeepSpeed's engine and reads it back from the config)
                    if self._mixed_precision == "no" and chosen.stated_precision() is not None:
                        self._mixed_precision = chosen.stated_precision()
                    for plugin in deepspeed_plugin.values():
                        plugin.set_mixed_precision(self._mixed_precision)
                    chosen.select(_from_accelerator_state=True)
            elif parse_flag_from_env("ACCELERATE_USE_FSDP") or fsdp_plugin is not None or (parallelism_config is not None and parallelism_config.cp_enabled):
                if parallelism_config is not None and parallelism_config.cp_enabled and fsdp_plugin is None:
                    from .utils.dataclasses import FullyShardedDataParallelPlugin

                    fsdp_plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                                 state_dict_type="SHARDED_STATE_DICT")
                    logger.warning("Context parallelism shards parameters with FSDP2; an FSDP2 plugin was created for you.")
                if fsdp_plugin is not None and parallelism_config is not None and parallelism_config.cp_enabled and fsdp_plugin.fsdp_version != 2:
                    raise ValueError("context parallelism requires fsdp_version=2")
                if fsdp_plugin is not None and not getattr(fsdp_plugin, "is_no_shard", False):
                    self.distributed_type = DistributedType.FSDP
                    if self._mixed_precision != "no":
                        fsdp_plugin.set_mixed_precision(self._mixed_precision)
                self.fsdp_plugin = fsdp_plugin
            elif parse_flag_from_env("ACCELERATE_USE_MEGATRON_LM") or megatron_lm_plugin is not None:
                self.distributed_type = DistributedType.MEGATRON_LM
                self.megatron_lm_plugin = megatron_lm_plugin
        elif fsdp_plugin is not None:
            # single process: keep the plugin (checkpoint naming, activation checkpointing) but do not shard
            self.fsdp_plugin = fsdp_plugin
        if self.dynamo_plugin.backend != DynamoBackend.NO and self._mixed_precision == "no" and self.device.type == "cuda":
            torch.backends.cuda.matmul.allow_tf32 = True
        PartialState._shared_state["distributed_type"] = self.distributed_type

    @property
    def initialized(self) -> bool:
        return self._shared_state != PartialState._shared_state

    def __repr__(self):
        repr_ = PartialState().__repr__() + f"\nMixed precision type: {self.mixed_precision}\n"
        if self.distributed_type == DistributedType.DEEPSPEED:
            repr_ += f"ds_config: {self.deepspeed_plugin.deepspeed_config}\n"
        return repr_

    def _check_initialized(self, mixed_precision=None, cpu=None):
        """Raise if an already-built state is asked for a conflicting configuration."""
        if self.initialized:
            err = "AcceleratorState has already been initialized and cannot be changed, restart your runtime completely and pass `{flag}` to `Accelerator()`."
            if cpu and self.device.type != "cpu":
                raise ValueError(err.format(flag="cpu=True"))
            if mixed_precision is not None and mixed_precision != self._mixed_precision and self.distributed_type != DistributedType.DEEPSPEED:
                raise ValueError(err.format(flag=f"mixed_precision='{mixed_precision}'"))

    @property
    def mixed_precision(self):
        return self._mixed_precision

    @staticmethod
    def _reset_state(reset_partial_state: bool = False):
        AcceleratorState._shared_state.clear()
        if reset_partial_state:
            PartialState._reset_state()

    # everything about the process group itself is answered by PartialState ---------------------------------
    fork_launched = _from_partial_state("fork_launched")
    use_distributed = _from_partial_state("use_distributed")
    is_last_process = _from_partial_state("is_last_process")
    is_main_process = _from_partial_state("is_main_process")
    is_local_main_process = _from_partial_state("is_local_main_process")
    destroy_process_group = _from_partial_state("destroy_process_group", call=True)
    wait_for_everyone = _from_partial_state("wait_for_everyone", call=True)
    split_between_processes = _from_partial_state("split_between_processes", call=True)  # these three return context managers
    main_process_first = _from_partial_state("main_process_first", call=True)
    local_main_process_first = _from_partial_state("local_main_process_first", call=True)
    print = _from_partial_state("print", call=True)

    @property
    def is_fsdp2(self) -> bool:
        return self.distributed_type == DistributedType.FSDP and self.fsdp_plugin is not None and self.fsdp_plugin.fsdp_version == 2

    # DeepSpeed plugin registry ----------------------------------------------------------------
    @property
    def deepspeed_plugin(self):
        if self.distributed_type != DistributedType.DEEPSPEED or not self.deepspeed_plugins:
            return None
        return next(plugin for plugin in self.deepspeed_plugins.values() if plugin.selected)

    def get_deepspeed_plugin(self, name: str):
        return self.deepspeed_plugins[name]

    def select_deepspeed_plugin(self, name: str = None):
        for key, plugin in self.deepspeed_plugins.items():
            if key != name:
                plugin._unselect()
        self.deepspeed_plugins[name].select(_from_accelerator_state=True)

    def __getattr__(self, name: str):
        if name in self._known_attrs:
            raise AttributeError(
                f"`AcceleratorState` object has no attribute `{name}`. This happens if `AcceleratorState._reset_state()` "
                "was called and an `Accelerator` or `PartialState` was not reinitialized."
            )
        raise AttributeError(f"'AcceleratorState' object has no attribute '{name}'")


def _plugin_setting(key: str, default):
    return property(lambda self: self.plugin_kwargs.get(key, default))


class GradientState:
    """Gradient-accumulation bookkeeping shared by the Accelerator, optimizers, schedulers and dataloaders.

    ``sync_gradients`` — whether this micro-step ends in a gradient reduction + optimizer step;
    ``end_of_dataloader`` / ``remainder`` — facts of the innermost active dataloader;
    ``num_steps``, ``adjust_scheduler``, ``sync_with_dataloader``, ``sync_each_batch`` — the plugin.
    """

    _shared_state = SharedDict()

    def __init__(self, gradient_accumulation_plugin: Optional[GradientAccumulationPlugin] = None):
        self.__dict__ = self._shared_state
        if not self.initialized:
            self.sync_gradients = True
            self.active_dataloader = None
            self.dataloader_references = [None]
            self.plugin_kwargs = gradient_accumulation_plugin.to_kwargs() if gradient_accumulation_plugin is not None else {}
            self._is_xla_gradients_synced = False
        # a later Accelerator with a different plugin overrides the stored kwargs
        if gradient_accumulation_plugin is not None and self.plugin_kwargs != gradient_accumulation_plugin.to_kwargs():
            self.plugin_kwargs = gradient_accumulation_plugin.to_kwargs()

    # the plugin's settings, with the defaults that apply when no plugin was given
    num_steps = _plugin_setting("num_steps", 1)
    adjust_scheduler = _plugin_setting("adjust_scheduler", False)
    sync_with_dataloader = _plugin_setting("sync_with_dataloader", True)
    sync_each_batch = _plugin_setting("sync_each_batch", False)

    @property
    def initialized(self) -> bool:
        return bool(GradientState._shared_state)

    @property
    def in_dataloader(self) -> bool:
        return self.active_dataloader is not None

    @property
    def end_of_dataloader(self) -> bool:
        """Is the innermost active dataloader handing out its last batch?  False outside of any dataloader."""
        return self.in_dataloader and self.active_dataloader.end_of_dataloader

    @property
    def remainder(self) -> int:
        """Real samples in the last global batch of the active dataloader (-1: unknown or not iterating)."""
        return self.active_dataloader.remainder if self.in_dataloader else -1

    def __repr__(self):
        facts = {"Sync Gradients": self.sync_gradients, "At end of current dataloader": self.end_of_dataloader,
                 "Extra samples added": self.remainder, "Gradient accumulation plugin": self.plugin_kwargs}
        return "".join(f"{label}: {value}\n" for label, value in facts.items())

    @property
    def is_xla_gradients_synced(self):
        return True

    @is_xla_gradients_synced.setter
    def is_xla_gradients_synced(self, value):
        self._is_xla_gradients_synced = value

    def _set_sync_gradients(self, sync_gradients: bool):
        self.sync_gradients = sync_gradients

    def _add_dataloader(self, dataloader):
        self.active_dataloader = dataloader
        self.dataloader_references.append(weakref.ref(dataloader))

    def _remove_dataloader(self, dataloader):
        refs = self.dataloader_references
        for i in range(len(refs) - 1, 0, -1):
            obj = refs[i]() if refs[i] is not None else None
            if obj is dataloader or obj is None:
                del refs[i]
                if obj is dataloader:
                    break
        last = refs[-1]
        self.active_dataloader = last() if last is not None else None

    @staticmethod
    def _reset_state():
        GradientState._shared_state.clear()
