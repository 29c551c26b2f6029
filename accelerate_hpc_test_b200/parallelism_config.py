"""``ParallelismConfig``: the N-D layout ``dp_replicate x dp_shard x cp x sp x tp`` of a job.

API parity with reference src/accelerate/parallelism_config.py (fields :70-81, env defaults
``PARALLELISM_CONFIG_*`` :274-289, mesh order :267, flattened sub-meshes ``dp`` / ``dp_shard_cp`` /
``dp_cp`` :237-242, validation :328-377).  One deliberate difference: the ``sp`` axis is always part
of the mesh — the reference builds no mesh when the DeepSpeed SP backend is active
(:218-221) although its own ALST example indexes ``torch_device_mesh["sp"]``.
"""

from __future__ import annotations

import os
import warnings
from dataclasses import dataclass
from typing import TYPE_CHECKING, Optional, Union

from .utils.dataclasses import DeepSpeedSequenceParallelConfig, DistributedType, TorchContextParallelConfig, TorchTensorParallelConfig
from .utils.versions import is_torch_version

if TYPE_CHECKING:
    from .accelerator import Accelerator


@dataclass
class ParallelismConfig:
    """Sizes of every parallel axis.  The product must equal the number of processes.

    ``dp_replicate_size`` — replicas (DDP-style all-reduce between them);
    ``dp_shard_size`` — parameter/gradient/optimizer shards (FSDP2-style);
    ``tp_size`` — tensor parallel (column/row-parallel linears, fused AG->GEMM / GEMM->RS);
    ``cp_size`` — context parallel (ring / all-gather attention over sequence shards);
    ``sp_size`` — Ulysses sequence parallel (all-to-all between sequence- and head-sharding).
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

    def __repr__(self):
        return (
            "ParallelismConfig(\n "
            f"\tdp_replicate_size={self.dp_replicate_size},\n"
            f"\tdp_shard_size={self.dp_shard_size},\n"
            f"\ttp_size={self.tp_size},\n"
            f"\tcp_size={self.cp_size},\n"
            f"\tcp_backend={self.cp_backend},\n"
            f"\tsp_size={self.sp_size},\n"
            f"\tsp_backend={self.sp_backend},\n"
            f"\ttotal_size={self.total_size}\n"
            f"\ttp_handler={self.tp_handler},\n"
            f"\tcp_handler={self.cp_handler})\n"
        )

    def to_json(self):
        import copy

        d = copy.deepcopy(self.__dict__)
        d.pop("device_mesh", None)
        return {k: (v.__dict__ if hasattr(v, "__dict__") else v) for k, v in d.items()}

    # ------------------------------------------------------------------ axis names
    @property
    def dp_dim_names(self):
        dims = []
        if self.dp_replicate_enabled:
            dims.append("dp_replicate")
        if self.dp_shard_enabled:
            dims.append("dp_shard")
        return dims

    @property
    def non_dp_dim_names(self):
        dims = []
        if self.tp_enabled:
            dims.append("tp")
        if self.cp_enabled:
            dims.append("cp")
        if self.sp_enabled:
            dims.append("sp")
        return dims

    @property
    def dp_shard_cp_dim_names(self):
        dims = []
        if self.dp_shard_enabled:
            dims.append("dp_shard")
        if self.cp_enabled:
            dims.append("cp")
        if self.sp_enabled:
            dims.append("sp")  # sequence shards are data shards for the sharded engine (cp and sp are exclusive)
        return dims

    @property
    def dp_cp_dim_names(self):
        dims = list(self.dp_dim_names)
        if self.cp_enabled:
            dims.append("cp")
        return dims

    @property
    def fsdp_dim_names(self):
        """Mesh axes the sharded engine spans: ``[dp_replicate,] dp_shard_cp``."""
        dims = []
        if self.dp_replicate_enabled:
            dims.append("dp_replicate")
        dims.append("dp_shard_cp")
        return dims

    # ------------------------------------------------------------------ sizes / flags
    @property
    def total_size(self):
        return self.dp_replicate_size * self.dp_shard_size * self.tp_size * self.cp_size * self.sp_size

    @property
    def non_data_parallel_size(self):
        return self.tp_size * self.cp_size * self.sp_size

    @property
    def data_parallel_size(self):
        return self.dp_replicate_size * self.dp_shard_size

    @property
    def dp_replicate_enabled(self):
        return self.dp_replicate_size > 1

    @property
    def dp_shard_enabled(self):
        return self.dp_shard_size > 1

    @property
    def tp_enabled(self):
        return self.tp_size > 1

    @property
    def cp_enabled(self):
        return self.cp_size > 1

    @property
    def sp_enabled(self):
        return self.sp_size > 1

    @property
    def active_mesh_dims(self):
        return self.dp_dim_names + self.non_dp_dim_names

    # ------------------------------------------------------------------ mesh
    def build_device_mesh(self, device_type: str):
        """``torch.distributed.DeviceMesh`` over the enabled axes plus the flattened dp sub-meshes."""
        from torch.distributed.device_mesh import init_device_mesh

        mesh = self._get_mesh()
        if len(mesh) == 0:
            return None
        dim_names, shape = mesh
        device_mesh = init_device_mesh(device_type, shape, mesh_dim_names=dim_names)
        if self.dp_dim_names:
            device_mesh[tuple(self.dp_dim_names)]._flatten("dp")
        if self.dp_shard_cp_dim_names:
            device_mesh[tuple(self.dp_shard_cp_dim_names)]._flatten("dp_shard_cp")
        if self.dp_cp_dim_names:
            device_mesh[tuple(self.dp_cp_dim_names)]._flatten("dp_cp")
        return device_mesh

    def get_device_mesh(self, device_type: Optional[str] = None):
        if self.device_mesh is None:
            if device_type is None:
                raise ValueError("You need to pass a device_type e.g cuda to build the device mesh")
            self.device_mesh = self.build_device_mesh(device_type)
        elif device_type is not None and self.device_mesh.device_type != device_type:
            raise ValueError(f"The device_mesh is already created with device type {self.device_mesh.device_type}. However, you are trying to get a device mesh with device_type {device_type}. Please check if you correctly initialized your device_mesh")
        return self.device_mesh

    def _get_mesh(self) -> tuple:
        """(names, sizes) of the enabled axes in the canonical order dp_replicate, dp_shard, cp, sp, tp."""
        order = ["dp_replicate", "dp_shard", "cp", "sp", "tp"]
        sizes = {name: getattr(self, f"{name}_size") for name in order}
        enabled = [(n, s) for n, s in sizes.items() if s > 1]
        if not enabled:
            return tuple()
        return tuple(zip(*enabled))

    # ------------------------------------------------------------------ validation
    def __post_init__(self):
        p = "PARALLELISM_CONFIG_"
        if self.dp_replicate_size is None:
            self.dp_replicate_size = int(os.environ.get(p + "DP_REPLICATE_SIZE", "1"))
        if self.dp_shard_size is None:
            self.dp_shard_size = int(os.environ.get(p + "DP_SHARD_SIZE", "1"))
        if self.tp_size is None:
            self.tp_size = int(os.environ.get(p + "TP_SIZE", "1"))
        if self.cp_size is None:
            self.cp_size = int(os.environ.get(p + "CP_SIZE", "1"))
        if self.cp_backend is None:
            self.cp_backend = os.environ.get(p + "CP_BACKEND", "torch")
        if self.sp_size is None:
            self.sp_size = int(os.environ.get(p + "SP_SIZE", "1"))
        if self.sp_backend is None:
            self.sp_backend = os.environ.get(p + "SP_BACKEND", "deepspeed")
        if self.tp_size > 1 and self.tp_handler is None:
            self.tp_handler = TorchTensorParallelConfig()
        if self.cp_size > 1 and self.cp_handler is None:
            self.cp_handler = TorchContextParallelConfig()
        if self.sp_size > 1 and self.sp_handler is None:
            self.sp_handler = DeepSpeedSequenceParallelConfig()
        for name in ("dp_replicate_size", "dp_shard_size", "tp_size", "cp_size", "sp_size"):
            if getattr(self, name) < 1:
                raise ValueError(f"{name} must be at least 1, but got {getattr(self, name)}")
        if self.cp_backend not in ("torch", "b200"):
            raise ValueError(f"Invalid cp_backend: {self.cp_backend}. Valid backends are: ['torch', 'b200']")
        if self.sp_backend not in ("deepspeed", "b200"):
            raise ValueError(f"Invalid sp_backend: {self.sp_backend}. Valid backends are: ['deepspeed', 'b200']")
        if self.cp_size > 1 and self.sp_size > 1:
            raise ValueError("Context parallelism (cp_size > 1) and sequence parallelism (sp_size > 1) are mutually exclusive; pick one.")
        if (self.tp_size > 1 or self.cp_size > 1) and self.dp_replicate_size > 1 and self.dp_shard_size == 1:
            raise ValueError("Tensor/Context parallelism (tp/cp_size > 1) cannot be used with pure data parallelism (dp_replicate_size > 1 and dp_shard_size == 1). Please set dp_shard_size > 1 and dp_replicate_size == 1 to compose FSDP + TP/CP for 2D parallel, or set dp_replicate_size == 1 and dp_shard_size > 1 to compose HSDP + TP/CP for 3D parallel.")

    def _set_size(self, parallelism: str, size: int):
        assert parallelism in ("dp_replicate", "dp_shard", "tp", "cp", "sp"), f"Parallelism must be one of dp_replicate, dp_shard, tp, cp, sp, got {parallelism}"
        if size == 1:
            setattr(self, f"{parallelism}_size", 1)
            return
        assert self.device_mesh is None, "Cannot set the size after the device mesh was built"
        setattr(self, f"{parallelism}_size", size)

    def _validate_accelerator(self, accelerator: "Accelerator"):
        """Check that the configuration matches the launched job; demote to plain DDP when trivial."""
        _warnings = set()
        if not accelerator.multi_device and self.total_size == 1:
            return
        if accelerator.num_processes > 1 and self.total_size == 1:
            # nothing asked: everything is data parallel
            self._set_size("dp_replicate", accelerator.num_processes)
        if self.total_size != accelerator.num_processes:
            raise ValueError(
                f"ParallelismConfig total_size ({self.total_size}) does not match num_processes ({accelerator.num_processes}). "
                "Please adjust dp_replicate_size/ dp_shard_size/tp_size/cp_size/sp_size."
            )
        if self.total_size > 1 and not (accelerator.is_fsdp2 or accelerator.multi_device):
            raise ValueError(f"ParallelismConfig is only compatible DistributedType.FSDP (version 2) or DistributedType.Multi{{Device}}, but got {accelerator.distributed_type}.")
        for parallelism, size in (("tp", self.tp_size), ("cp", self.cp_size), ("sp", self.sp_size)):
            if size == 1 and getattr(self, f"{parallelism}_handler") is not None:
                _warnings.add(f"ParallelismConfig.{parallelism}_handler is set, but {parallelism}_size is set to 1. This handler will be ignored.")
        if _warnings and accelerator.is_main_process:
            warnings.warn("ParallelismConfig has the following warnings:\n" + "\n".join(_warnings), UserWarning)
