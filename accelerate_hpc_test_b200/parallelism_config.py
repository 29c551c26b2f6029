"""``ParallelismConfig``: how the processes of a job are arranged — ``dp_replicate x dp_shard x cp x sp x tp``.

One table (``_AXES``) names the axes in mesh order, outermost first; everything else — which axes are enabled, the mesh shape, the
flattened sub-meshes the engines ask for (``dp``, ``dp_shard_cp``, ``dp_cp``), the ``PARALLELISM_CONFIG_*`` environment defaults the
launcher exports — is derived from it.  On a B200 node the inner axes (tp, sp, cp) land on NVSwitch neighbours first, which is what
the fused GEMM+collective kernels of those axes want.

API parity: reference ``src/accelerate/parallelism_config.py`` (fields :70-81, env defaults :274-289, mesh order :267, flattened
sub-meshes :237-242, validation :328-377).  One deliberate difference: the ``sp`` axis is always part of the mesh — the reference builds
no mesh when the DeepSpeed SP backend is active (:218-221) although its own ALST example indexes ``torch_device_mesh["sp"]``.
"""

from __future__ import annotations

import copy
import os
import warnings
from dataclasses import dataclass
from typing import TYPE_CHECKING, Optional, Union

from .utils.dataclasses import DeepSpeedSequenceParallelConfig, TorchContextParallelConfig, TorchTensorParallelConfig

if TYPE_CHECKING:
    from .accelerator import Accelerator

# mesh order, outermost first
_AXES = ("dp_replicate", "dp_shard", "cp", "sp", "tp")
_ENV_PREFIX = "PARALLELISM_CONFIG_"
_BACKENDS = {"cp": ("torch", ("torch", "b200")), "sp": ("deepspeed", ("deepspeed", "b200"))}  # axis -> (default, accepted)
_HANDLER_TYPES = {"tp": TorchTensorParallelConfig, "cp": TorchContextParallelConfig, "sp": DeepSpeedSequenceParallelConfig}


@dataclass
class ParallelismConfig:
    """Size of every parallel axis; the product must equal the number of processes.  Unset sizes come from ``PARALLELISM_CONFIG_*``
    (what ``accelerate launch`` exports) and default to 1.

    * ``dp_replicate_size`` — replicas that all-reduce their gradients (DDP-style);
    * ``dp_shard_size`` — shards of parameters, gradients and optimizer state (FSDP2-style);
    * ``tp_size`` — tensor parallel: column / row-parallel linears with all-gather→GEMM and GEMM→reduce-scatter fused;
    * ``cp_size`` — context parallel: attention over sequence shards, K/V exchanged (``cp_backend`` ``torch`` or ``b200``);
    * ``sp_size`` — Ulysses sequence parallel: all-to-all between sequence- and head-sharding (``sp_backend`` ``deepspeed`` or ``b200``).
    """

    dp_replicate_size: Optional[int] = None
    dp_shard_size: Optional[int] = None
    tp_size: Optional[int] = None
    cp_size: Optional[int] = None
    cp_backend: Optional[str] = None
    sp_size: Optional[int] = None
    sp_backend: Optional[str] = None
    tp_handler: Union[None, TorchTensorParallelConfig] = None
    cp_handler: Union[None, TorchContextParallelConfig] = None
    sp_handler: Union[None, DeepSpeedSequenceParallelConfig] = None
    device_mesh = None

    # ------------------------------------------------------------------ construction
    def __post_init__(self):
        for axis in _AXES:
            if self._size(axis) is None:
                setattr(self, f"{axis}_size", int(os.environ.get(f"{_ENV_PREFIX}{axis.upper()}_SIZE", "1")))
            if self._size(axis) < 1:
                raise ValueError(f"{axis}_size must be at least 1, but got {self._size(axis)}")
        for axis, (default, accepted) in _BACKENDS.items():
            field_name = f"{axis}_backend"
            if getattr(self, field_name) is None:
                setattr(self, field_name, os.environ.get(f"{_ENV_PREFIX}{axis.upper()}_BACKEND", default))
            if getattr(self, field_name) not in accepted:
                raise ValueError(f"Invalid {field_name}: {getattr(self, field_name)}. Valid backends are: {list(accepted)}")
        for axis, handler_type in _HANDLER_TYPES.items():
            if self._enabled(axis) and getattr(self, f"{axis}_handler") is None:
                setattr(self, f"{axis}_handler", handler_type())
        if self.cp_enabled and self.sp_enabled:
            raise ValueError("Context parallelism (cp_size > 1) and sequence parallelism (sp_size > 1) are mutually exclusive; pick one.")
        if (self.tp_enabled or self.cp_enabled) and self.dp_replicate_enabled and not self.dp_shard_enabled:
            raise ValueError(
                "Tensor/Context parallelism (tp/cp_size > 1) cannot be used with pure data parallelism (dp_replicate_size > 1 and "
                "dp_shard_size == 1). Please set dp_shard_size > 1 and dp_replicate_size == 1 to compose FSDP + TP/CP for 2D parallel, "
                "or set dp_replicate_size == 1 and dp_shard_size > 1 to compose HSDP + TP/CP for 3D parallel."
            )

    def _size(self, axis: str) -> Optional[int]:
        return getattr(self, f"{axis}_size")

    def _enabled(self, axis: str) -> bool:
        return self._size(axis) > 1

    def _names(self, *axes: str) -> list[str]:
        """Those of ``axes`` that are enabled, in the order given."""
        return [axis for axis in axes if self._enabled(axis)]

    def _product(self, *axes: str) -> int:
        total = 1
        for axis in axes:
            total *= self._size(axis)
        return total

    def __repr__(self):
        shown = [f"{axis}_size={self._size(axis)}" for axis in ("dp_replicate", "dp_shard", "tp", "cp")]
        shown += [f"cp_backend={self.cp_backend}", f"sp_size={self.sp_size}", f"sp_backend={self.sp_backend}", f"total_size={self.total_size}",
                  f"tp_handler={self.tp_handler}", f"cp_handler={self.cp_handler}"]
        return "ParallelismConfig(\n\t" + ",\n\t".join(shown) + ")\n"

    def to_json(self) -> dict:
        """Plain-dict form for logging / trackers (handlers expanded, the mesh left out)."""
        fields = {key: value for key, value in vars(self).items() if key != "device_mesh"}
        return {key: (copy.deepcopy(vars(value)) if hasattr(value, "__dict__") else value) for key, value in fields.items()}

    # ------------------------------------------------------------------ which axes are on
    dp_replicate_enabled = property(lambda self: self._enabled("dp_replicate"))
    dp_shard_enabled = property(lambda self: self._enabled("dp_shard"))
    tp_enabled = property(lambda self: self._enabled("tp"))
    cp_enabled = property(lambda self: self._enabled("cp"))
    sp_enabled = property(lambda self: self._enabled("sp"))

    # ------------------------------------------------------------------ groups of axes (mesh dimension names)
    @property
    def dp_dim_names(self) -> list[str]:
        """Axes whose ranks see different batches."""
        return self._names("dp_replicate", "dp_shard")

    @property
    def non_dp_dim_names(self) -> list[str]:
        """Axes whose ranks cooperate on the same batch."""
        return self._names("tp", "cp", "sp")

    @property
    def dp_shard_cp_dim_names(self) -> list[str]:
        """Axes the sharded engine shards over: sequence shards (cp, or sp — they are exclusive) count as data shards for it."""
        return self._names("dp_shard", "cp", "sp")

    @property
    def dp_cp_dim_names(self) -> list[str]:
        """Axes the loss has to be averaged over."""
        return self._names("dp_replicate", "dp_shard", "cp")

    @property
    def fsdp_dim_names(self) -> list[str]:
        """Sub-meshes handed to the sharded engine: ``[dp_replicate,] dp_shard_cp`` (two entries mean hybrid sharding)."""
        return self._names("dp_replicate") + ["dp_shard_cp"]

    @property
    def active_mesh_dims(self) -> list[str]:
        return self.dp_dim_names + self.non_dp_dim_names

    # ------------------------------------------------------------------ sizes
    @property
    def total_size(self) -> int:
        return self._product(*_AXES)

    @property
    def non_data_parallel_size(self) -> int:
        return self._product("tp", "cp", "sp")

    @property
    def data_parallel_size(self) -> int:
        return self._product("dp_replicate", "dp_shard")

    # ------------------------------------------------------------------ mesh
    def _get_mesh(self) -> tuple:
        """``(names, sizes)`` of the enabled axes in mesh order — an empty tuple when every axis is 1."""
        enabled = self._names(*_AXES)
        if not enabled:
            return ()
        return tuple(enabled), tuple(self._size(axis) for axis in enabled)

    def build_device_mesh(self, device_type: str):
        """A ``DeviceMesh`` over the enabled axes, with the flattened views ``dp``, ``dp_shard_cp`` and ``dp_cp`` registered on it."""
        from torch.distributed.device_mesh import init_device_mesh

        layout = self._get_mesh()
        if not layout:
            return None
        names, shape = layout
        mesh = init_device_mesh(device_type, shape, mesh_dim_names=names)
        for flat_name, members in (("dp", self.dp_dim_names), ("dp_shard_cp", self.dp_shard_cp_dim_names), ("dp_cp", self.dp_cp_dim_names)):
            if members:
                mesh[tuple(members)]._flatten(flat_name)
        return mesh

    def get_device_mesh(self, device_type: Optional[str] = None):
        """The mesh, built on first use."""
        if self.device_mesh is None:
            if device_type is None:
                raise ValueError("You need to pass a device_type e.g cuda to build the device mesh")
            self.device_mesh = self.build_device_mesh(device_type)
        elif device_type is not None and device_type != self.device_mesh.device_type:
            raise ValueError(
                f"The device_mesh is already created with device type {self.device_mesh.device_type}. However, you are trying to get a "
                f"device mesh with device_type {device_type}. Please check if you correctly initialized your device_mesh"
            )
        return self.device_mesh

    # ------------------------------------------------------------------ fitting the config to the launched job
    def _set_size(self, parallelism: str, size: int) -> None:
        assert parallelism in _AXES, f"Parallelism must be one of {', '.join(_AXES)}, got {parallelism}"
        assert size == 1 or self.device_mesh is None, "Cannot set the size after the device mesh was built"
        setattr(self, f"{parallelism}_size", size)

    def _validate_accelerator(self, accelerator: "Accelerator") -> None:
        """Called by ``Accelerator``: the axes must multiply to the number of processes.  A config that asks for nothing on a
        multi-process job means plain data parallelism."""
        if self.total_size == 1:
            if not accelerator.multi_device:
                return
            if accelerator.num_processes > 1:
                self._set_size("dp_replicate", accelerator.num_processes)
        if self.total_size != accelerator.num_processes:
            raise ValueError(
                f"ParallelismConfig total_size ({self.total_size}) does not match num_processes ({accelerator.num_processes}). "
                "Please adjust dp_replicate_size/ dp_shard_size/tp_size/cp_size/sp_size."
            )
        if self.total_size > 1 and not (accelerator.is_fsdp2 or accelerator.multi_device):
            raise ValueError(
                "ParallelismConfig is only compatible DistributedType.FSDP (version 2) or DistributedType.Multi{Device}, but got "
                f"{accelerator.distributed_type}."
            )
        idle = [f"ParallelismConfig.{axis}_handler is set, but {axis}_size is set to 1. This handler will be ignored."
                for axis in _HANDLER_TYPES if not self._enabled(axis) and getattr(self, f"{axis}_handler") is not None]
        if idle and accelerator.is_main_process:
            warnings.warn("ParallelismConfig has the following warnings:\n" + "\n".join(idle), UserWarning)
