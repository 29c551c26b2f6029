"""Schema and file I/O of the launcher configuration (``default_config.yaml``).

* Location: ``$HF_HOME/accelerate`` or ``$XDG_CACHE_HOME/huggingface/accelerate`` (default
  ``~/.cache/huggingface/accelerate``) — the same place the reference uses, so existing files are found.
* ``.json`` files are read/written as JSON, everything else as YAML.
* Files written by older versions are upgraded on load (``fp16: true`` → ``mixed_precision: fp16``, a bare
  ``dynamo_backend`` → ``dynamo_config``, missing ``use_cpu`` / ``debug`` / ``enable_cpu_affinity``); keys the
  schema does not know are rejected with the full list, never silently dropped.

Parity target: reference ``src/accelerate/commands/config/config_args.py:29-252`` (field names are the file format).
"""

from __future__ import annotations

import dataclasses
import enum
import json
import os
from dataclasses import dataclass
from typing import Optional, Union

import yaml

from ...utils.dataclasses import ComputeEnvironment, DistributedType, SageMakerDistributedType


def _cache_root() -> str:
    xdg = os.environ.get("XDG_CACHE_HOME", "~/.cache")
    return os.path.expanduser(os.environ.get("HF_HOME", os.path.join(xdg, "huggingface")))


hf_cache_home = _cache_root()
cache_dir = os.path.join(hf_cache_home, "accelerate")
default_yaml_config_file = os.path.join(cache_dir, "default_config.yaml")
default_json_config_file = default_yaml_config_file  # historical alias: the default has been YAML for a long time
default_config_file = default_yaml_config_file


# --------------------------------------------------------------------------- raw file access
def _is_json(path) -> bool:
    return str(path).endswith(".json")


def _read_mapping(path) -> dict:
    with open(path, encoding="utf-8") as fh:
        data = json.load(fh) if _is_json(path) else yaml.safe_load(fh)
    return dict(data or {})


def _write_mapping(path, data: dict) -> None:
    with open(path, "w", encoding="utf-8") as fh:
        if _is_json(path):
            fh.write(json.dumps(data, indent=2, sort_keys=True) + "\n")
        else:
            yaml.safe_dump(data, fh)


def _plain(value):
    """Enum → its value, nested dicts likewise; an empty dict becomes ``None`` (and is then left out of the file)."""
    if isinstance(value, enum.Enum):
        return value.value
    if isinstance(value, dict):
        return {k: _plain(v) for k, v in value.items()} or None
    return value


def load_config_from_file(config_file=None):
    """Load ``config_file`` (or the default file) into a ``ClusterConfig`` / ``SageMakerConfig``."""
    if config_file is None:
        config_file = default_config_file
    elif not os.path.isfile(config_file):
        raise FileNotFoundError(
            f"The passed configuration file `{config_file}` does not exist. Please pass an existing file to `accelerate launch`, "
            "or use the default one created through `accelerate config` and run `accelerate launch` without the `--config_file` argument."
        )
    where = _read_mapping(config_file).get("compute_environment", ComputeEnvironment.LOCAL_MACHINE)
    schema = ClusterConfig if where == ComputeEnvironment.LOCAL_MACHINE else SageMakerConfig
    return schema.load(config_file)


# --------------------------------------------------------------------------- schemas
_BACKFILL = {"use_cpu": False, "debug": False, "enable_cpu_affinity": False}


@dataclass
class BaseConfig:
    compute_environment: ComputeEnvironment
    distributed_type: Union[DistributedType, SageMakerDistributedType]
    mixed_precision: str
    use_cpu: bool
    debug: bool

    # ---- normalisation ----
    def __post_init__(self):
        if isinstance(self.compute_environment, str):
            self.compute_environment = ComputeEnvironment(self.compute_environment)
        if isinstance(self.distributed_type, str):
            on_sagemaker = self.compute_environment == ComputeEnvironment.AMAZON_SAGEMAKER
            self.distributed_type = (SageMakerDistributedType if on_sagemaker else DistributedType)(self.distributed_type)
        if getattr(self, "dynamo_config", None) is None:
            self.dynamo_config = {}

    @staticmethod
    def process_config(config_dict: dict) -> dict:
        """Upgrade a mapping written by an older version to today's keys (in place; also returned)."""
        cfg = config_dict
        cfg.setdefault("compute_environment", ComputeEnvironment.LOCAL_MACHINE)
        if "distributed_type" not in cfg:
            raise ValueError("A `distributed_type` must be specified in the config file.")
        if cfg["distributed_type"] == DistributedType.NO:
            cfg.setdefault("num_processes", 1)
        legacy_fp16 = bool(cfg.pop("fp16", False))
        if "mixed_precision" not in cfg:
            cfg["mixed_precision"] = "fp16" if legacy_fp16 else None
        if "dynamo_backend" in cfg:
            backend = cfg.pop("dynamo_backend")
            cfg["dynamo_config"] = {} if backend == "NO" else {"dynamo_backend": backend}
        for key, value in _BACKFILL.items():
            cfg.setdefault(key, value)
        return cfg

    @classmethod
    def _reject_unknown(cls, mapping: dict, source) -> None:
        known = {f.name for f in dataclasses.fields(cls)}
        unknown = sorted(set(mapping) - known)
        if unknown:
            raise ValueError(
                f"The config file at {source} had unknown keys ({unknown}), please try upgrading your `accelerate` version "
                "or fix (and potentially remove) these keys from your config file."
            )

    # ---- load / save ----
    @classmethod
    def load(cls, path):
        mapping = cls.process_config(_read_mapping(path))
        cls._reject_unknown(mapping, path)
        return cls(**mapping)

    def save(self, path) -> None:
        _write_mapping(path, self.to_dict())

    def to_dict(self) -> dict:
        out = {}
        for key, value in vars(self).items():
            value = _plain(value)
            if value is not None:
                out[key] = value
        return out

    # reference-compatible spellings
    @classmethod
    def from_json_file(cls, json_file=None):
        return cls.load(default_json_config_file if json_file is None else json_file)

    @classmethod
    def from_yaml_file(cls, yaml_file=None):
        return cls.load(default_yaml_config_file if yaml_file is None else yaml_file)

    def to_json_file(self, json_file) -> None:
        with open(json_file, "w", encoding="utf-8") as fh:
            fh.write(json.dumps(self.to_dict(), indent=2, sort_keys=True) + "\n")

    def to_yaml_file(self, yaml_file) -> None:
        with open(yaml_file, "w", encoding="utf-8") as fh:
            yaml.safe_dump(self.to_dict(), fh)


_CLUSTER_SECTIONS = ("deepspeed_config", "fsdp_config", "parallelism_config", "megatron_lm_config", "ipex_config",
                     "mpirun_config", "fp8_config")


@dataclass
class ClusterConfig(BaseConfig):
    num_processes: int = -1  # -1: decide at launch time from the visible devices
    machine_rank: int = 0
    num_machines: int = 1
    gpu_ids: Optional[str] = None
    main_process_ip: Optional[str] = None
    main_process_port: Optional[int] = None
    rdzv_backend: Optional[str] = "static"
    same_network: Optional[bool] = False
    main_training_function: str = "main"
    enable_cpu_affinity: bool = False
    # one mapping per strategy; empty means "not used"
    deepspeed_config: dict = None
    fsdp_config: dict = None
    parallelism_config: dict = None
    megatron_lm_config: dict = None
    ipex_config: dict = None
    mpirun_config: dict = None
    downcast_bf16: bool = False
    # TPU keys: accepted so that old files load, ignored by this build
    tpu_name: str = None
    tpu_zone: str = None
    tpu_use_cluster: bool = False
    tpu_use_sudo: bool = False
    command_file: str = None
    commands: list = None
    tpu_vm: list = None
    tpu_env: list = None
    dynamo_config: dict = None
    fp8_config: dict = None

    def __post_init__(self):
        for name in _CLUSTER_SECTIONS:
            if getattr(self, name) is None:
                setattr(self, name, {})
        super().__post_init__()


@dataclass
class SageMakerConfig(BaseConfig):
    """Accepted for file compatibility only; the launcher of this build refuses SageMaker environments."""

    ec2_instance_type: str = None
    iam_role_name: str = None
    image_uri: Optional[str] = None
    profile: Optional[str] = None
    region: str = "us-east-1"
    num_machines: int = 1
    gpu_ids: str = "all"
    base_job_name: str = None
    pytorch_version: str = None
    transformers_version: str = None
    py_version: str = None
    sagemaker_inputs_file: str = None
    sagemaker_metrics_file: str = None
    additional_args: dict = None
    dynamo_config: dict = None
    enable_cpu_affinity: bool = False
