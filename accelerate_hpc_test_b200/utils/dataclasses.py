"""Option objects of the framework: enums, ``KwargsHandler`` s and strategy plugins.

The public names, fields and environment variables follow the reference so user scripts and YAML
configs keep working (reference src/accelerate/utils/dataclasses.py: KwargsHandler :68-86,
DDP kwargs :155-237, InitProcessGroupKwargs :273-299, AO/TE recipes :310-434, ProfileKwargs
:483-597, enums :600-805, DataLoaderConfiguration :813-905, ProjectConfiguration :908-968,
GradientAccumulationPlugin :971-1020, TorchDynamoPlugin :1023-1109, DeepSpeedPlugin :1113-1562,
FullyShardedDataParallelPlugin :1566-2167, TP/CP/SP configs :2171-2282, MegatronLMPlugin
:2286-2650, BnbQuantizationConfig :3025-3176).  What differs is where they lead: on this build the
FSDP/DDP/TP/SP plugins configure the native sm_100a engines in :mod:`..parallel`, and the
DeepSpeed / Megatron-LM plugins are configuration shells mapped onto those engines.
"""

from __future__ import annotations

import copy
import enum
import functools
import logging
import os
import warnings
from contextlib import contextmanager
from dataclasses import dataclass, field, fields
from datetime import timedelta
from typing import Any, Callable, Iterable, Literal, Optional, Union

import torch

from .constants import FSDP_AUTO_WRAP_POLICY, FSDP_BACKWARD_PREFETCH, FSDP_SHARDING_STRATEGY
from .environment import parse_flag_from_env, str_to_bool

logger = logging.getLogger(__name__)


# =============================================================================== base machinery
class KwargsHandler:
    """Dataclass mix-in: ``to_kwargs()`` returns only the fields whose value differs from the default."""

    def to_dict(self):
        return copy.deepcopy(self.__dict__)

    def to_kwargs(self):
        from .other import clear_environment_for_defaults

        with clear_environment_for_defaults():
            default = self.__class__().to_dict()
        mine = self.to_dict()
        return {k: v for k, v in mine.items() if default.get(k, object()) != v}


class EnumWithContains(enum.EnumMeta):
    """Enum metaclass so ``"bf16" in PrecisionType`` works on values."""

    def __contains__(cls, item):
        try:
            cls(item)
        except ValueError:
            return False
        return True


class BaseEnum(enum.Enum, metaclass=EnumWithContains):
    def __str__(self):
        return self.value

    @classmethod
    def list(cls):
        return list(map(str, cls))


# =============================================================================== enums
class DistributedType(str, enum.Enum):
    """How this job is distributed. Only the CUDA/CPU members are ever selected on this build."""

    NO = "NO"
    MULTI_CPU = "MULTI_CPU"
    MULTI_GPU = "MULTI_GPU"
    DEEPSPEED = "DEEPSPEED"
    FSDP = "FSDP"
    MEGATRON_LM = "MEGATRON_LM"
    # names kept so configs written for other back-ends still parse; never selected here
    MULTI_NPU = "MULTI_NPU"
    MULTI_MLU = "MULTI_MLU"
    MULTI_SDAA = "MULTI_SDAA"
    MULTI_MUSA = "MULTI_MUSA"
    MULTI_XPU = "MULTI_XPU"
    MULTI_HPU = "MULTI_HPU"
    XLA = "XLA"


class SageMakerDistributedType(str, enum.Enum):
    NO = "NO"
    DATA_PARALLEL = "DATA_PARALLEL"
    MODEL_PARALLEL = "MODEL_PARALLEL"


class FP8BackendType(str, enum.Enum):
    """fp8 implementation. ``NATIVE`` = this framework's own MXFP8 block-scaled path."""

    NO = "NO"
    NATIVE = "NATIVE"
    TE = "TE"
    MSAMP = "MSAMP"
    AO = "AO"


# value sets of the fp8 recipe fields, for annotations and validation
Backend = Literal["MSAMP", "TE"]
OptLevel = Literal["O1", "O2"]
FP8Format = Literal["HYBRID", "E4M3", "E5M2"]
AmaxComputeAlgorithm = Literal["max", "most_recent"]


class ComputeEnvironment(str, enum.Enum):
    LOCAL_MACHINE = "LOCAL_MACHINE"
    AMAZON_SAGEMAKER = "AMAZON_SAGEMAKER"


class DynamoBackend(str, BaseEnum):
    NO = "NO"
    EAGER = "EAGER"
    AOT_EAGER = "AOT_EAGER"
    INDUCTOR = "INDUCTOR"
    AOT_TS_NVFUSER = "AOT_TS_NVFUSER"
    NVPRIMS_NVFUSER = "NVPRIMS_NVFUSER"
    CUDAGRAPHS = "CUDAGRAPHS"
    OFI = "OFI"
    FX2TRT = "FX2TRT"
    ONNXRT = "ONNXRT"
    TENSORRT = "TENSORRT"
    AOT_TORCHXLA_TRACE_ONCE = "AOT_TORCHXLA_TRACE_ONCE"
    TORCHXLA_TRACE_ONCE = "TORCHXLA_TRACE_ONCE"
    IPEX = "IPEX"
    TVM = "TVM"
    HPU_BACKEND = "HPU_BACKEND"


class LoggerType(BaseEnum):
    ALL = "all"
    AIM = "aim"
    TENSORBOARD = "tensorboard"
    WANDB = "wandb"
    COMETML = "comet_ml"
    MLFLOW = "mlflow"
    CLEARML = "clearml"
    DVCLIVE = "dvclive"
    SWANLAB = "swanlab"
    TRACKIO = "trackio"
    JSONL = "jsonl"


class PrecisionType(str, BaseEnum):
    NO = "no"
    FP8 = "fp8"
    FP16 = "fp16"
    BF16 = "bf16"


class RNGType(BaseEnum):
    TORCH = "torch"
    CUDA = "cuda"
    MLU = "mlu"
    SDAA = "sdaa"
    MUSA = "musa"
    NPU = "npu"
    XLA = "xla"
    XPU = "xpu"
    HPU = "hpu"
    GENERATOR = "generator"


class CustomDtype(enum.Enum):
    """Storage dtypes torch has no name for (used by the device-map size arithmetic)."""

    FP8 = "fp8"
    INT4 = "int4"
    INT2 = "int2"


class DDPCommunicationHookType(BaseEnum):
    NO = "no"
    FP16 = "fp16"
    BF16 = "bf16"
    POWER_SGD = "power_sgd"
    BATCHED_POWER_SGD = "batched_power_sgd"


# =============================================================================== kwargs handlers
@dataclass
class AutocastKwargs(KwargsHandler):
    enabled: bool = True
    cache_enabled: bool = None


@dataclass
class DistributedDataParallelKwargs(KwargsHandler):
    """Knobs of the data-parallel gradient reducer.

    On CUDA they configure the native bucketed reducer (:class:`..parallel.ddp.B200DataParallel`):
    ``bucket_cap_mb`` sizes the fused cast+scale+all-reduce buckets, ``comm_hook`` FP16/BF16 selects
    the wire dtype, ``gradient_as_bucket_view`` makes ``.grad`` alias the flat bucket.  On CPU (gloo)
    they are forwarded to ``torch.nn.parallel.DistributedDataParallel``.
    """

    dim: int = 0
    broadcast_buffers: bool = True
    bucket_cap_mb: int = 25
    find_unused_parameters: bool = False
    check_reduction: bool = False
    gradient_as_bucket_view: bool = False
    static_graph: bool = False
    comm_hook: DDPCommunicationHookType = DDPCommunicationHookType.NO
    comm_wrapper: Literal[
        DDPCommunicationHookType.NO, DDPCommunicationHookType.FP16, DDPCommunicationHookType.BF16
    ] = DDPCommunicationHookType.NO
    comm_state_option: dict = field(default_factory=dict)

    def to_dict(self, ignore_keys=("comm_hook", "comm_wrapper", "comm_state_option")):
        return {k: v for k, v in super().to_dict().items() if k not in ignore_keys}

    def register_comm_hook(self, model):
        """Attach the configured torch DDP communication hook (CPU/gloo path and torch-DDP models)."""
        from torch.distributed.algorithms.ddp_comm_hooks import default_hooks, powerSGD_hook

        hooks = {
            DDPCommunicationHookType.FP16: default_hooks.fp16_compress_hook,
            DDPCommunicationHookType.BF16: default_hooks.bf16_compress_hook,
            DDPCommunicationHookType.POWER_SGD: powerSGD_hook.powerSGD_hook,
            DDPCommunicationHookType.BATCHED_POWER_SGD: powerSGD_hook.batched_powerSGD_hook,
        }
        wrappers = {
            DDPCommunicationHookType.FP16: default_hooks.fp16_compress_wrapper,
            DDPCommunicationHookType.BF16: default_hooks.bf16_compress_wrapper,
        }
        hook = hooks.get(self.comm_hook)
        wrapper = wrappers.get(self.comm_wrapper)
        if hook and wrapper:
            hook = wrapper(hook)
        if hook:
            state = (
                powerSGD_hook.PowerSGDState(None, **self.comm_state_option)
                if self.comm_hook in (DDPCommunicationHookType.POWER_SGD, DDPCommunicationHookType.BATCHED_POWER_SGD)
                else None
            )
            model.register_comm_hook(state=state, hook=hook)


@dataclass
class GradScalerKwargs(KwargsHandler):
    init_scale: float = 65536.0
    growth_factor: float = 2.0
    backoff_factor: float = 0.5
    growth_interval: int = 2000
    enabled: bool = True


@dataclass
class InitProcessGroupKwargs(KwargsHandler):
    backend: Optional[str] = "nccl"
    init_method: Optional[str] = None
    timeout: Optional[timedelta] = None

    def __post_init__(self):
        if self.timeout is None:
            seconds = 1800 if self.backend != "nccl" else 600
            self.timeout = timedelta(seconds=seconds)


@dataclass
class AORecipeKwargs(KwargsHandler):
    """torchao float8 recipe (library backend; the native MXFP8 path needs no recipe object)."""

    config: Optional[Any] = None
    module_filter_func: Optional[Callable] = None
    pad_inner_dim: Optional[bool] = None
    enable_fsdp_float8_all_gather: Optional[bool] = None

    def __post_init__(self):
        env_prefix = "ACCELERATE_FP8_"
        if not (self.config is None or self.config.__class__.__name__ == "Float8LinearConfig"):
            raise ValueError("`config` must be a torchao `Float8LinearConfig`")
        if self.pad_inner_dim is None:
            self.pad_inner_dim = parse_flag_from_env(env_prefix + "PAD_INNER_DIM", default=True)
        if self.enable_fsdp_float8_all_gather is None:
            self.enable_fsdp_float8_all_gather = parse_flag_from_env(env_prefix + "ENABLE_FSDP_FLOAT8_ALL_GATHER", default=True)


@dataclass
class TERecipeKwargs(KwargsHandler):
    """Scaling recipe. ``use_mxfp8_block_scaling`` selects 1x32 UE8M0 block scaling (native on B200)."""

    use_autocast_during_eval: bool = None
    margin: int = None
    interval: int = None
    fp8_format: Literal["HYBRID", "E4M3", "E5M2"] = None
    amax_history_len: int = None
    amax_compute_algo: Literal["max", "most_recent"] = None
    override_linear_precision: tuple[bool, bool, bool] = None
    use_mxfp8_block_scaling: bool = None

    def __post_init__(self):
        env_prefix = "ACCELERATE_FP8_"
        if self.use_autocast_during_eval is None:
            self.use_autocast_during_eval = parse_flag_from_env(env_prefix + "USE_AUTOCAST_DURING_EVAL")
        if self.margin is None:
            self.margin = int(os.environ.get(env_prefix + "MARGIN", 0))
        if self.interval is None:
            self.interval = int(os.environ.get(env_prefix + "INTERVAL", 1))
        if self.fp8_format is None:
            self.fp8_format = os.environ.get(env_prefix + "FORMAT", "HYBRID")
        self.fp8_format = self.fp8_format.upper()
        if self.fp8_format not in ("HYBRID", "E4M3", "E5M2"):
            raise ValueError("`fp8_format` must be one of HYBRID, E4M3, E5M2")
        if self.amax_compute_algo is None:
            self.amax_compute_algo = os.environ.get(env_prefix + "AMAX_COMPUTE_ALGO", "most_recent")
        self.amax_compute_algo = self.amax_compute_algo.lower()
        if self.amax_compute_algo not in ("max", "most_recent"):
            raise ValueError("`amax_compute_algo` must be 'max' or 'most_recent'")
        if self.amax_history_len is None:
            self.amax_history_len = int(os.environ.get(env_prefix + "AMAX_HISTORY_LEN", 1024))
        if self.override_linear_precision is None:
            self.override_linear_precision = (
                parse_flag_from_env(env_prefix + "OVERRIDE_FPROP"),
                parse_flag_from_env(env_prefix + "OVERRIDE_DGRAD"),
                parse_flag_from_env(env_prefix + "OVERRIDE_WGRAD"),
            )
        if self.use_mxfp8_block_scaling is None:
            self.use_mxfp8_block_scaling = parse_flag_from_env(env_prefix + "USE_MXFP8_BLOCK_SCALING")


@dataclass
class MSAMPRecipeKwargs(KwargsHandler):
    opt_level: Literal["O1", "O2"] = None

    def __post_init__(self):
        if self.opt_level is None:
            self.opt_level = os.environ.get("ACCELERATE_FP8_OPT_LEVEL", "O2")
        if self.opt_level not in ("O1", "O2"):
            raise ValueError("`opt_level` must be 'O1' or 'O2'")


@dataclass
class FP8RecipeKwargs(TERecipeKwargs, MSAMPRecipeKwargs):
    """Deprecated umbrella recipe; pick ``TERecipeKwargs`` / ``AORecipeKwargs`` / ``MSAMPRecipeKwargs``."""

    backend: Literal["MSAMP", "TE"] = None

    def __post_init__(self):
        warnings.warn("FP8RecipeKwargs is deprecated; use TERecipeKwargs, AORecipeKwargs or MSAMPRecipeKwargs.", FutureWarning)
        if self.backend is None:
            self.backend = os.environ.get("ACCELERATE_FP8_BACKEND", "TE")
        self.backend = self.backend.upper()
        if self.backend not in ("MSAMP", "TE"):
            raise ValueError("`backend` must be 'MSAMP' or 'TE'")
        TERecipeKwargs.__post_init__(self)
        MSAMPRecipeKwargs.__post_init__(self)


ProfilerActivity = Literal["cpu", "xpu", "mtia", "cuda", "hpu"]


@dataclass
class ProfileKwargs(KwargsHandler):
    """Factory for ``torch.profiler.profile`` used by ``Accelerator.profile()``."""

    activities: Optional[list[ProfilerActivity]] = None
    schedule_option: Optional[dict[str, int]] = None
    on_trace_ready: Optional[Callable] = None
    record_shapes: bool = False
    profile_memory: bool = False
    with_stack: bool = False
    with_flops: bool = False
    with_modules: bool = False
    output_trace_dir: Optional[str] = None

    def _get_profiler_activity(self, activity: str):
        table = {"cpu": torch.profiler.ProfilerActivity.CPU, "cuda": torch.profiler.ProfilerActivity.CUDA}
        for extra in ("xpu", "mtia", "hpu"):
            member = getattr(torch.profiler.ProfilerActivity, extra.upper(), None)
            if member is not None:
                table[extra] = member
        if activity not in table:
            raise ValueError(f"Invalid profiler activity: {activity}. Must be one of {list(table)}.")
        return table[activity]

    def build(self) -> torch.profiler.profile:
        activities = None
        if self.activities is not None:
            activities = [self._get_profiler_activity(a) for a in self.activities]
        schedule = None
        if self.schedule_option is not None:
            schedule = torch.profiler.schedule(**self.schedule_option)
        return torch.profiler.profile(
            activities=activities, schedule=schedule, on_trace_ready=self.on_trace_ready, record_shapes=self.record_shapes,
            profile_memory=self.profile_memory, with_stack=self.with_stack, with_flops=self.with_flops, with_modules=self.with_modules,
        )


# =============================================================================== small configs
@dataclass
class TensorInformation:
    shape: torch.Size
    dtype: torch.dtype


@dataclass
class DataLoaderConfiguration:
    split_batches: bool = False
    dispatch_batches: bool = None
    even_batches: bool = True
    use_seedable_sampler: bool = False
    data_seed: int = None
    non_blocking: bool = False
    use_stateful_dataloader: bool = False


@dataclass
class ProjectConfiguration:
    project_dir: str = None
    logging_dir: str = None
    automatic_checkpoint_naming: bool = False
    total_limit: int = None
    iteration: int = 0
    save_on_each_node: bool = False

    def set_directories(self, project_dir: str = None):
        self.project_dir = project_dir
        if self.logging_dir is None:
            self.logging_dir = project_dir

    def __post_init__(self):
        self.set_directories(self.project_dir)


@dataclass
class GradientAccumulationPlugin(KwargsHandler):
    num_steps: int = None
    adjust_scheduler: bool = True
    sync_with_dataloader: bool = True
    sync_each_batch: bool = False


@dataclass
class TorchDynamoPlugin(KwargsHandler):
    """Kept for config compatibility. The hot path of this framework is hand-written kernels, so a
    non-``NO`` backend only wraps the *user's* module with ``torch.compile`` when explicitly asked."""

    backend: DynamoBackend = None
    mode: str = None
    fullgraph: bool = None
    dynamic: bool = None
    options: Any = None
    disable: bool = False
    use_regional_compilation: bool = None

    def __post_init__(self):
        prefix = "ACCELERATE_DYNAMO_"
        if self.backend is None:
            self.backend = os.environ.get(prefix + "BACKEND", "no")
        self.backend = DynamoBackend(str(self.backend).upper())
        if self.mode is None:
            self.mode = os.environ.get(prefix + "MODE", "default")
        if self.fullgraph is None:
            self.fullgraph = str_to_bool(os.environ.get(prefix + "USE_FULLGRAPH", "False")) == 1
        if self.use_regional_compilation is None:
            self.use_regional_compilation = str_to_bool(os.environ.get(prefix + "USE_REGIONAL_COMPILATION", "False")) == 1
        if self.dynamic is None and os.environ.get(prefix + "USE_DYNAMIC", None) is not None:
            self.dynamic = str_to_bool(os.environ.get(prefix + "USE_DYNAMIC", "False")) == 1

    def to_dict(self):
        d = copy.deepcopy(self.__dict__)
        d["backend"] = d["backend"].value.lower()
        return d

    def to_kwargs(self):
        kw = super().to_kwargs()
        kw.pop("use_regional_compilation", None)
        return kw


# =============================================================================== DeepSpeed (config shell)
_DS_FIELD_PATHS = {  # plugin attribute -> dotted path of the same setting in a DeepSpeed config
    "gradient_accumulation_steps": "gradient_accumulation_steps",
    "gradient_clipping": "gradient_clipping",
    "zero_stage": "zero_optimization.stage",
    "offload_optimizer_device": "zero_optimization.offload_optimizer.device",
    "offload_param_device": "zero_optimization.offload_param.device",
    "offload_param_nvme_path": "zero_optimization.offload_param.nvme_path",
    "offload_optimizer_nvme_path": "zero_optimization.offload_optimizer.nvme_path",
    "zero3_save_16bit_model": "zero_optimization.stage3_gather_16bit_weights_on_model_save",
}


def _int_or_auto(text):
    return int(text) if str(text).isdigit() else text


def _float_or_auto(text):
    return text if text == "auto" else float(text)


_DS_ENV_DEFAULTS = (  # (attribute, environment variable, parser, default) for the fields the launcher exports
    ("gradient_accumulation_steps", "ACCELERATE_GRADIENT_ACCUMULATION_STEPS", _int_or_auto, "auto"),
    ("gradient_clipping", "ACCELERATE_GRADIENT_CLIPPING", _float_or_auto, "auto"),
    ("zero_stage", "ACCELERATE_DEEPSPEED_ZERO_STAGE", int, 2),
    ("offload_optimizer_device", "ACCELERATE_DEEPSPEED_OFFLOAD_OPTIMIZER_DEVICE", str, "none"),
    ("offload_param_device", "ACCELERATE_DEEPSPEED_OFFLOAD_PARAM_DEVICE", str, "none"),
    ("offload_optimizer_nvme_path", "ACCELERATE_DEEPSPEED_OFFLOAD_OPTIMIZER_NVME_PATH", str, "none"),
    ("offload_param_nvme_path", "ACCELERATE_DEEPSPEED_OFFLOAD_PARAM_NVME_PATH", str, "none"),
    ("zero3_save_16bit_model", "ACCELERATE_DEEPSPEED_ZERO3_SAVE_16BIT_MODEL", lambda text: str(text).lower() == "true", "false"),
    ("enable_msamp", "ACCELERATE_FP8_BACKEND", lambda text: text == "MSAMP", None),
    ("msamp_opt_level", "ACCELERATE_FP8_OPT_LEVEL", str, "O1"),
)
_DS_SHADOWED_BY_CONFIG_FILE = (
    "ACCELERATE_GRADIENT_ACCUMULATION_STEPS", "ACCELERATE_GRADIENT_CLIPPING", "ACCELERATE_DEEPSPEED_ZERO_STAGE",
    "ACCELERATE_DEEPSPEED_OFFLOAD_OPTIMIZER_DEVICE", "ACCELERATE_DEEPSPEED_OFFLOAD_PARAM_DEVICE",
    "ACCELERATE_DEEPSPEED_OFFLOAD_PARAM_NVME_PATH", "ACCELERATE_DEEPSPEED_OFFLOAD_OPTIMIZER_NVME_PATH",
    "ACCELERATE_DEEPSPEED_ZERO3_SAVE_16BIT_MODEL", "ACCELERATE_MIXED_PRECISION",
)


@dataclass
class DeepSpeedPlugin:
    """DeepSpeed configuration shell.

    DeepSpeed itself is not a dependency of this build: ZeRO-3 maps onto the native FSDP2-style
    engine (stage 3 == full shard, stage 2 == ``reshard_after_forward=False``), ZeRO-1/2 onto the
    DDP engine with a sharded optimizer, and Ulysses sequence parallelism onto
    :mod:`..parallel.ulysses`.  The fields and ``ACCELERATE_DEEPSPEED_*`` variables match the
    reference (src/accelerate/utils/dataclasses.py:1113-1256) so existing configs load.
    """

    hf_ds_config: Any = None
    gradient_accumulation_steps: int = None
    gradient_clipping: float = None
    zero_stage: int = None
    is_train_batch_min: bool = True
    offload_optimizer_device: str = None
    offload_param_device: str = None
    offload_optimizer_nvme_path: str = None
    offload_param_nvme_path: str = None
    zero3_init_flag: bool = None
    zero3_save_16bit_model: bool = None
    transformer_moe_cls_names: str = None
    enable_msamp: bool = None
    msamp_opt_level: Optional[Literal["O1", "O2"]] = None

    def __post_init__(self):
        for attribute, variable, parse, default in _DS_ENV_DEFAULTS:
            if getattr(self, attribute) is None:
                setattr(self, attribute, parse(os.environ.get(variable, default)))
        source = self.hf_ds_config
        if source is None:
            source = os.environ.get("ACCELERATE_DEEPSPEED_CONFIG_FILE", "none")
        if isinstance(source, str) and source == "none":
            self._compose_config()
        else:
            self._adopt_config(source)
        self.deepspeed_config = self.hf_ds_config.config
        self.deepspeed_config["steps_per_print"] = float("inf")
        if self.zero3_init_flag is None:
            self.zero3_init_flag = str_to_bool(os.environ.get("ACCELERATE_DEEPSPEED_ZERO3_INIT", str(self.hf_ds_config.is_zero3()))) == 1
        if self.zero3_init_flag and not self.hf_ds_config.is_zero3():
            warnings.warn("DeepSpeed Zero3 Init flag is only applicable for ZeRO Stage 3. Setting it to False.")
            self.zero3_init_flag = False
        self._selected = False
        self.dschf = None
        if self.enable_msamp and "msamp" not in self.deepspeed_config:
            if self.zero_stage == 3:
                raise NotImplementedError("MS-AMP is not supported for ZeRO Stage 3. Please use ZeRO Stage 0, 1, or 2 instead.")
            if self.msamp_opt_level not in ("O1", "O2"):
                raise ValueError("Invalid optimization level for MS-AMP. Please use one of ['O1' or'O2'].")
            self.deepspeed_config["msamp"] = {"enabled": True, "opt_level": self.msamp_opt_level}

    # ------------------------------------------------------------------ where the configuration comes from
    def _compose_config(self) -> None:
        """No file: the configuration is written from the plugin's own fields."""
        from .deepspeed import HfDeepSpeedConfig

        def nvme(device, path):
            return path if device == "nvme" else None

        config = {
            "train_batch_size": "auto",
            "train_micro_batch_size_per_gpu": "auto",
            "gradient_accumulation_steps": self.gradient_accumulation_steps,
            "zero_optimization": {
                "stage": self.zero_stage,
                "offload_optimizer": {"device": self.offload_optimizer_device,
                                      "nvme_path": nvme(self.offload_optimizer_device, self.offload_optimizer_nvme_path)},
                "offload_param": {"device": self.offload_param_device, "nvme_path": nvme(self.offload_param_device, self.offload_param_nvme_path)},
                "stage3_gather_16bit_weights_on_model_save": self.zero3_save_16bit_model,
            },
        }
        if self.gradient_clipping:
            config["gradient_clipping"] = self.gradient_clipping
        self.hf_ds_config = HfDeepSpeedConfig(config)

    def _adopt_config(self, source) -> None:
        """A file / dict / encoded config rules: its ``auto`` entries are filled from the plugin's fields, and the fields then take
        every value the config states, so ``plugin.zero_stage`` and ``config["zero_optimization"]["stage"]`` cannot disagree."""
        from .deepspeed import HfDeepSpeedConfig

        self.hf_ds_config = source if isinstance(source, HfDeepSpeedConfig) else HfDeepSpeedConfig(source)
        config = self.hf_ds_config.config
        config.setdefault("gradient_accumulation_steps", 1)
        if "zero_optimization" not in config:
            raise ValueError("Please specify the ZeRO optimization config in the DeepSpeed config.")
        self._deepspeed_config_checks()
        stated = {path: getattr(self, attribute) for attribute, path in _DS_FIELD_PATHS.items() if getattr(self, attribute) is not None}
        for path in stated:
            self.fill_match(path, must_match=False, **stated)
        self.hf_ds_config.set_stage_and_offload()
        for attribute, path in _DS_FIELD_PATHS.items():
            value = self.hf_ds_config.get_value(path)
            if value is not None and value != "auto":
                setattr(self, attribute, value)

    def _deepspeed_config_checks(self) -> None:
        """With a DeepSpeed config file, the same settings must not also come from the accelerate config (the launcher lists the
        ones it found in ``ACCELERATE_CONFIG_DS_FIELDS``)."""
        shadowed = [name.replace("ACCELERATE_", "").replace("DEEPSPEED_", "").lower() for name in _DS_SHADOWED_BY_CONFIG_FILE]
        given = os.environ.get("ACCELERATE_CONFIG_DS_FIELDS", "").split(",")
        if any(name in shadowed for name in given):
            raise ValueError(
                f"When using `deepspeed_config_file`, the following accelerate config variables will be ignored: {shadowed}.\n"
                "Please specify them appropriately in the DeepSpeed config file.\n"
                "If you are using an accelerate config file, remove others config variables mentioned in the above specified list.\n"
                "The easiest method is to create a new config following the questionnaire via `accelerate config`.\n"
                "It will only ask for the necessary config variables when using `deepspeed_config_file`."
            )

    # ------------------------------------------------------------------ reading and filling
    def fill_match(self, ds_key_long, mismatches=None, must_match=True, **kwargs):
        """An ``"auto"`` entry takes the value ``kwargs`` has under the same dotted name (an error if it has none); a stated entry is
        compared with it when ``must_match`` and a difference is appended to ``mismatches``."""
        mismatches = [] if mismatches is None else mismatches
        holder, leaf = self.hf_ds_config.find_config_node(ds_key_long)
        if holder is None:
            return
        if holder.get(leaf) == "auto":
            if ds_key_long not in kwargs:
                raise ValueError(f"`{ds_key_long}` not found in kwargs. Please specify `{ds_key_long}` without `auto` (set to correct value) "
                                 "in the DeepSpeed config file or pass it in kwargs.")
            holder[leaf] = kwargs[ds_key_long]
            return
        if must_match and holder.get(leaf) is not None and ds_key_long in kwargs and holder[leaf] != kwargs[ds_key_long]:
            mismatches.append(f"- ds {ds_key_long}={holder[leaf]} vs arg {ds_key_long}={kwargs[ds_key_long]}")

    def deepspeed_config_process(self, prefix="", mismatches=None, config=None, must_match=True, **kwargs):
        """``fill_match`` for every leaf of the configuration; raises once, at the top level, listing every mismatch."""
        top = mismatches is None
        mismatches = [] if mismatches is None else mismatches
        config = self.deepspeed_config if config is None else config
        for key, value in config.items():
            if isinstance(value, dict):
                self.deepspeed_config_process(prefix=f"{prefix}{key}.", mismatches=mismatches, config=value, must_match=must_match, **kwargs)
            else:
                self.fill_match(prefix + key, mismatches, must_match=must_match, **kwargs)
        if mismatches and (top or prefix == ""):
            listing = "\n".join(mismatches)
            raise ValueError("Please correct the following DeepSpeed config values that mismatch kwargs "
                             f" values:\n{listing}\nThe easiest method is to set these DeepSpeed config values to 'auto'.")

    def get_value(self, ds_key_long, default=None):
        return self.hf_ds_config.get_value(ds_key_long, default)

    def is_auto(self, ds_key_long) -> bool:
        return self.hf_ds_config.get_value(ds_key_long) == "auto"

    def stated_precision(self) -> Optional[str]:
        """``"fp16"`` / ``"bf16"`` when the configuration itself switches one of them on, else ``None``."""
        for name in ("fp16", "bf16"):
            if str(self.deepspeed_config.get(name, {}).get("enabled", False)).lower() == "true":
                return name
        return None

    def set_mixed_precision(self, mixed_precision) -> None:
        """Write the precision into the ``fp16`` / ``bf16`` sections (filling ``auto``); a section that enables the *other* 16-bit
        type contradicts ``mixed_precision`` and is an error."""
        config = self.deepspeed_config
        wanted = {"fp16.enabled": mixed_precision == "fp16", "bf16.enabled": mixed_precision in ("bf16", "fp8")}  # fp8 rides on bf16 autocast
        if mixed_precision != "no":  # checked first, so that a refused call leaves the configuration as it was
            other = "bf16" if mixed_precision == "fp16" else "fp16"
            if str(config.get(other, {}).get("enabled", "False")).lower() == "true":
                raise ValueError(f"`--mixed_precision` arg cannot be set to `{mixed_precision}` when `{other}` is set in the DeepSpeed config file.")
        if mixed_precision == "fp16":
            config.setdefault("fp16", {"enabled": True, "auto_cast": True})
        elif mixed_precision in ("bf16", "fp8"):
            config.setdefault("bf16", {"enabled": True})
        if mixed_precision == "fp8" and self.enable_msamp:
            config.setdefault("msamp", {"enabled": True, "opt_level": self.msamp_opt_level})
        for section in ("fp16", "bf16"):
            config.setdefault(section, {"enabled": False})
        self.fill_match("fp16.enabled", must_match=False, **wanted)
        self.fill_match("bf16.enabled", must_match=False, **wanted)

    def set_deepspeed_weakref(self) -> None:
        """The reference parks a resolved copy of the config where ``transformers`` finds it, so that ``from_pretrained`` builds ZeRO-3
        models partitioned.  There is no DeepSpeed here to partition for — big models are built under ``init_empty_weights`` and
        sharded by the native engine — so the resolved copy (``auto`` batch entries set to 1) is only kept on the plugin, ``dschf``."""
        from .deepspeed import HfDeepSpeedConfig

        resolved = copy.deepcopy(self.deepspeed_config)
        for key in ("gradient_accumulation_steps", "train_micro_batch_size_per_gpu"):
            if resolved.get(key, "auto") == "auto":
                resolved[key] = 1
        if resolved.get("train_batch_size") == "auto":
            del resolved["train_batch_size"]
        self.dschf = HfDeepSpeedConfig(resolved)

    def is_zero3_init_enabled(self):
        return self.zero3_init_flag

    @contextmanager
    def zero3_init_context_manager(self, enable=False):
        """Temporarily switch ``zero3_init_flag`` (e.g. off, to build an auxiliary model whole)."""
        before = self.zero3_init_flag
        if before == enable:
            yield
            return
        self.zero3_init_flag = enable
        self.set_deepspeed_weakref()
        try:
            yield
        finally:
            self.zero3_init_flag = before
            self.set_deepspeed_weakref()

    def set_moe_leaf_modules(self, model):
        """Mixture-of-experts blocks named in ``transformer_moe_cls_names`` must not be gathered expert by expert (which experts run
        depends on the data, so ranks would disagree on the order of gathers): every instance is marked ``_z3_leaf`` and the
        sharded engine keeps a marked module in one unit.  Returns the classes found."""
        if self.transformer_moe_cls_names is None:
            self.transformer_moe_cls_names = os.environ.get("ACCELERATE_DEEPSPEED_MOE_LAYER_CLS_NAMES", None)
        if self.transformer_moe_cls_names is None:
            return []
        found = []
        for name in self.transformer_moe_cls_names.split(","):
            cls = get_module_class_from_name(model, name.strip())
            if cls is None:
                raise Exception(f"Could not find a transformer layer class called '{name.strip()}' to wrap in the model.")
            found.append(cls)
        for module in model.modules():
            if isinstance(module, tuple(found)):
                module._z3_leaf = True
        return found

    # ------------------------------------------------------------------ selection among several plugins
    def select(self, _from_accelerator_state: bool = False):
        if not _from_accelerator_state:
            raise ValueError("A `DeepSpeedPlugin` object must be enabled manually by calling `AcceleratorState().enable_deepspeed_plugin(plugin_key)`.")
        self.set_deepspeed_weakref()
        self._selected = True

    def _unselect(self):
        self._selected = False

    def _set_selected(self, value: bool):
        self._selected = value

    @property
    def selected(self) -> bool:
        return self._selected

    @selected.setter
    def selected(self, value):
        raise NotImplementedError("'enabled' can only be set through calling 'AcceleratorState().enable_deepspeed_plugin(key)'.")

    def to_fsdp_plugin(self) -> "FullyShardedDataParallelPlugin":
        """The native-engine equivalent of this ZeRO configuration."""
        return FullyShardedDataParallelPlugin(
            fsdp_version=2,
            reshard_after_forward=self.zero_stage >= 3,
            cpu_offload=self.offload_param_device not in (None, "none"),
            auto_wrap_policy="no_wrap",
        )


# =============================================================================== FSDP
def _resolve_dtype(x):
    if isinstance(x, torch.dtype) or x is None:
        return x
    table = {"fp32": torch.float32, "float32": torch.float32, "fp16": torch.float16, "float16": torch.float16,
             "bf16": torch.bfloat16, "bfloat16": torch.bfloat16}
    if str(x).lower() in table:
        return table[str(x).lower()]
    raise ValueError(f"unknown dtype spec {x!r}")


@dataclass
class MixedPrecisionPolicy:
    """param / reduce / output dtypes of a sharded unit (mirrors ``torch.distributed.fsdp.MixedPrecisionPolicy``)."""

    param_dtype: Optional[torch.dtype] = None
    reduce_dtype: Optional[torch.dtype] = None
    output_dtype: Optional[torch.dtype] = None
    cast_forward_inputs: bool = True


@dataclass
class FullyShardedDataParallelPlugin:
    """Configuration of sharded data parallelism.

    ``fsdp_version=2`` (per-parameter dim-0 sharding) is what the native engine implements; FSDP1
    option names are accepted and mapped the way the reference's ``accelerate to-fsdp2`` does
    (reference src/accelerate/commands/to_fsdp2.py:51-66): FULL_SHARD / HYBRID_SHARD ->
    ``reshard_after_forward=True``, SHARD_GRAD_OP / HYBRID_SHARD_ZERO2 -> ``False``, NO_SHARD -> DDP.
    """

    fsdp_version: int = None
    sharding_strategy: Union[str, Any] = None
    reshard_after_forward: Union[str, bool, Any] = None
    backward_prefetch: Optional[Union[str, Any]] = None
    mixed_precision_policy: Optional[Union[dict, str, Any]] = None
    auto_wrap_policy: Optional[Union[Callable, str]] = None
    cpu_offload: Union[bool, Any] = None
    ignored_modules: Optional[Union[Iterable[torch.nn.Module], str]] = None
    state_dict_type: Union[str, Any] = None
    state_dict_config: Optional[Any] = None
    optim_state_dict_config: Optional[Any] = None
    limit_all_gathers: bool = True
    use_orig_params: Optional[bool] = None
    param_init_fn: Optional[Callable[[torch.nn.Module], None]] = None
    sync_module_states: Optional[bool] = None
    forward_prefetch: bool = None
    activation_checkpointing: bool = None
    cpu_ram_efficient_loading: bool = None
    transformer_cls_names_to_wrap: Optional[list[str]] = None
    min_num_params: Optional[int] = None

    def __post_init__(self):
        p = "FSDP_"
        notes = []
        if self.fsdp_version is None:
            self.fsdp_version = int(os.environ.get(p + "VERSION", "1"))
        if self.fsdp_version not in (1, 2):
            raise ValueError(f"fsdp_version must be 1 or 2, got {self.fsdp_version}")

        # ---- sharding strategy / reshard_after_forward
        if self.fsdp_version == 1:
            if self.sharding_strategy is None:
                self.sharding_strategy = os.environ.get(p + "SHARDING_STRATEGY", "FULL_SHARD")
            if isinstance(self.sharding_strategy, int) and not isinstance(self.sharding_strategy, bool):
                self.sharding_strategy = str(self.sharding_strategy)
            if isinstance(self.sharding_strategy, str):
                s = self.sharding_strategy.upper()
                if s.isdigit():
                    s = FSDP_SHARDING_STRATEGY[int(s) - 1]
                if s not in FSDP_SHARDING_STRATEGY:
                    raise ValueError(f"sharding_strategy must be one of {FSDP_SHARDING_STRATEGY}")
                self.sharding_strategy = s
            elif hasattr(self.sharding_strategy, "name"):
                self.sharding_strategy = self.sharding_strategy.name
        elif self.sharding_strategy is not None:
            notes.append("sharding_strategy is deprecated with FSDP2; use reshard_after_forward")
        if self.reshard_after_forward is None:
            if self.fsdp_version == 2:
                if self.sharding_strategy is not None:
                    s = str(getattr(self.sharding_strategy, "name", self.sharding_strategy)).upper()
                    self.reshard_after_forward = s in ("FULL_SHARD", "HYBRID_SHARD")
                else:
                    self.reshard_after_forward = str_to_bool(os.environ.get(p + "RESHARD_AFTER_FORWARD", "true")) == 1
            else:
                self.reshard_after_forward = self.sharding_strategy
        if isinstance(self.reshard_after_forward, str):
            if self.fsdp_version == 2:
                try:
                    self.reshard_after_forward = str_to_bool(self.reshard_after_forward.lower(), to_bool=True)
                except ValueError:
                    pass  # not a truth value (an FSDP1 strategy name, say): reported below with the field's name
            else:
                self.reshard_after_forward = self.reshard_after_forward.upper()
        if self.fsdp_version == 2 and not isinstance(self.reshard_after_forward, bool):
            raise ValueError(f"reshard_after_forward must be a bool with fsdp_version=2, got {self.reshard_after_forward!r}")
        if self.fsdp_version == 1 and isinstance(self.reshard_after_forward, bool):
            raise ValueError("reshard_after_forward must be a sharding-strategy name with fsdp_version=1")

        # ---- offload / prefetch
        if self.cpu_offload is None:
            self.cpu_offload = str_to_bool(os.environ.get(p + "OFFLOAD_PARAMS", "False")) == 1
        if self.backward_prefetch is None:
            self.backward_prefetch = os.environ.get(p + "BACKWARD_PREFETCH", None)
        if isinstance(self.backward_prefetch, str):
            b = self.backward_prefetch.upper()
            if b.isdigit():
                b = FSDP_BACKWARD_PREFETCH[int(b) - 1]
            self.backward_prefetch = None if b == "NO_PREFETCH" else b
        if self.fsdp_version == 2 and self.backward_prefetch is not None:
            notes.append("backward_prefetch is implicit in FSDP2 (the native engine always prefetches the next unit)")
            self.backward_prefetch = None

        # ---- wrap policy
        if self.auto_wrap_policy is None:
            self.auto_wrap_policy = os.environ.get(p + "AUTO_WRAP_POLICY", "NO_WRAP")
        if isinstance(self.auto_wrap_policy, str):
            w = self.auto_wrap_policy.upper()
            if w not in FSDP_AUTO_WRAP_POLICY:
                raise ValueError(f"auto_wrap_policy must be one of {FSDP_AUTO_WRAP_POLICY} or a callable")
            self.auto_wrap_policy = w
            if w == "TRANSFORMER_BASED_WRAP":
                if self.transformer_cls_names_to_wrap is None:
                    self.transformer_cls_names_to_wrap = os.environ.get(p + "TRANSFORMER_CLS_TO_WRAP", None)
                if isinstance(self.transformer_cls_names_to_wrap, str):
                    self.transformer_cls_names_to_wrap = [s.strip() for s in self.transformer_cls_names_to_wrap.split(",") if s.strip()]
            elif w == "SIZE_BASED_WRAP":
                if self.min_num_params is None:
                    self.min_num_params = int(os.environ.get(p + "MIN_NUM_PARAMS", 0))
                if self.min_num_params <= 0:
                    raise ValueError("min_num_params must be > 0 for SIZE_BASED_WRAP")
        self._wrap_fn: Optional[Callable[[torch.nn.Module], bool]] = None

        # ---- FSDP1-only booleans
        if self.use_orig_params is None and self.fsdp_version == 1:
            self.use_orig_params = str_to_bool(os.environ.get(p + "USE_ORIG_PARAMS", "False")) == 1
        if self.sync_module_states is None and self.fsdp_version == 1:
            self.sync_module_states = str_to_bool(os.environ.get(p + "SYNC_MODULE_STATES", "False")) == 1
        if self.forward_prefetch is None and self.fsdp_version == 1:
            self.forward_prefetch = str_to_bool(os.environ.get(p + "FORWARD_PREFETCH", "False")) == 1
        if self.fsdp_version == 2 and self.forward_prefetch:
            raise ValueError("forward_prefetch is an FSDP1 option; with fsdp_version=2 the engine prefetches by itself")
        if self.activation_checkpointing is None:
            self.activation_checkpointing = str_to_bool(os.environ.get(p + "ACTIVATION_CHECKPOINTING", "False")) == 1
        if self.ignored_modules is None:
            self.ignored_modules = os.environ.get(p + "IGNORED_MODULES", None)
        if self.cpu_ram_efficient_loading is None:
            self.cpu_ram_efficient_loading = str_to_bool(os.environ.get(p + "CPU_RAM_EFFICIENT_LOADING", "False")) == 1
        else:
            os.environ[p + "CPU_RAM_EFFICIENT_LOADING"] = str(self.cpu_ram_efficient_loading)

        # ---- state dict type
        if self.state_dict_type is None:
            self.state_dict_type = os.environ.get(p + "STATE_DICT_TYPE", "FULL_STATE_DICT" if self.fsdp_version == 1 else "SHARDED_STATE_DICT")
        self.set_state_dict_type(self.state_dict_type)

        if isinstance(self.mixed_precision_policy, dict):
            self.set_mixed_precision(self.mixed_precision_policy)
        elif isinstance(self.mixed_precision_policy, str):
            self.set_mixed_precision(self.mixed_precision_policy)
        if notes:
            logger.warning("FSDP2 conversion notes:\n" + "\n".join(notes))

    # ------------------------------------------------------------------ helpers
    def set_state_dict_type(self, state_dict_type=None):
        if state_dict_type is None:
            state_dict_type = "SHARDED_STATE_DICT"
        name = str(getattr(state_dict_type, "name", state_dict_type)).upper()
        if name.isdigit():
            from .constants import FSDP_STATE_DICT_TYPE

            name = FSDP_STATE_DICT_TYPE[int(name) - 1]
        if name not in ("FULL_STATE_DICT", "LOCAL_STATE_DICT", "SHARDED_STATE_DICT"):
            raise ValueError(f"unknown state_dict_type {state_dict_type}")
        if name == "LOCAL_STATE_DICT" and self.fsdp_version == 2:
            raise ValueError("LOCAL_STATE_DICT does not exist with fsdp_version=2")
        self.state_dict_type = name

    def set_auto_wrap_policy(self, model):
        """Resolve the configured policy to ``self._wrap_fn(module) -> bool`` for ``model``."""
        pol = self.auto_wrap_policy
        if callable(pol):
            self._wrap_fn = lambda m, _pol=pol: bool(_call_wrap_policy(_pol, m))
            return
        if pol == "TRANSFORMER_BASED_WRAP":
            names = self.transformer_cls_names_to_wrap
            if names is None:
                names = list(getattr(model, "_no_split_modules", None) or [])
            classes = set()
            for n in names:
                cls = get_module_class_from_name(model, n)
                if cls is None:
                    raise ValueError(f"Could not find the transformer layer class {n} in the model.")
                classes.add(cls)
            classes = tuple(classes)
            self._wrap_fn = lambda m, _c=classes: isinstance(m, _c)
        elif pol == "SIZE_BASED_WRAP":
            thr = self.min_num_params
            self._wrap_fn = lambda m, _t=thr: sum(p.numel() for p in m.parameters()) >= _t
        else:
            self._wrap_fn = None

    def should_wrap(self, module) -> bool:
        return self._wrap_fn is not None and self._wrap_fn(module)

    def set_mixed_precision(self, mixed_precision, buffer_autocast=False, override=False):
        if isinstance(mixed_precision, MixedPrecisionPolicy) and not override:
            self.mixed_precision_policy = mixed_precision
            return
        if isinstance(mixed_precision, dict):
            allowed = {"param_dtype", "reduce_dtype", "output_dtype", "buffer_dtype", "cast_forward_inputs"}
            unknown = set(mixed_precision) - allowed
            if unknown:
                raise ValueError(f"Invalid mixed precision policy keys {sorted(unknown)}; allowed {sorted(allowed)}")
            kw = {k: _resolve_dtype(v) if k.endswith("dtype") else v for k, v in mixed_precision.items() if k != "buffer_dtype"}
            self.mixed_precision_policy = MixedPrecisionPolicy(**kw)
            return
        if isinstance(mixed_precision, str):
            dt = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32, "no": None, "fp8": torch.bfloat16}.get(mixed_precision.lower(), "bad")
            if dt == "bad":
                raise ValueError(f"Unknown mixed precision {mixed_precision}")
            if dt is None:
                return
            if self.mixed_precision_policy is None or override:
                self.mixed_precision_policy = MixedPrecisionPolicy(param_dtype=dt, reduce_dtype=dt, output_dtype=dt)
            return
        if mixed_precision is not None and (self.mixed_precision_policy is None or override):
            # a torch MixedPrecision / MixedPrecisionPolicy object: copy its dtypes
            self.mixed_precision_policy = MixedPrecisionPolicy(
                param_dtype=getattr(mixed_precision, "param_dtype", None),
                reduce_dtype=getattr(mixed_precision, "reduce_dtype", None),
                output_dtype=getattr(mixed_precision, "output_dtype", None),
            )

    def set_cpu_offload(self):
        """``cpu_offload`` as torch's policy object (what ``fully_shard`` / ``FullyShardedDataParallel`` take); booleans are what the
        native engine reads (``True`` sends the job to the torch fall-back, which receives the policy object)."""
        if not isinstance(self.cpu_offload, bool):
            return
        if self.fsdp_version == 2:
            from torch.distributed.fsdp import CPUOffloadPolicy, OffloadPolicy

            self.cpu_offload = CPUOffloadPolicy() if self.cpu_offload else OffloadPolicy()
        else:
            from torch.distributed.fsdp import CPUOffload

            self.cpu_offload = CPUOffload(offload_params=self.cpu_offload)

    def validate_cpu_offload(self):
        if self.fsdp_version == 2:
            from torch.distributed.fsdp import OffloadPolicy

            if not isinstance(self.cpu_offload, OffloadPolicy):
                raise ValueError(f"`cpu_offload` must be an instance of `torch.distributed.fsdp.OffloadPolicy` in FSDP2, got {self.cpu_offload}")
        else:
            from torch.distributed.fsdp import CPUOffload

            if not isinstance(self.cpu_offload, CPUOffload):
                raise ValueError(f"`cpu_offload` must be an instance of `torch.distributed.fsdp.CPUOffload` in FSDP1, got {self.cpu_offload}")

    @property
    def offloads_parameters(self) -> bool:
        """Whether parameters live on the CPU between uses, whichever way ``cpu_offload`` is spelled (bool or torch policy object)."""
        value = self.cpu_offload
        if isinstance(value, bool) or value is None:
            return bool(value)
        if hasattr(value, "offload_params"):
            return bool(value.offload_params)
        return type(value).__name__ == "CPUOffloadPolicy"

    def validate_mixed_precision_policy(self):
        pol = self.mixed_precision_policy
        if pol is not None and not isinstance(pol, MixedPrecisionPolicy):
            raise ValueError("mixed_precision_policy must be a dict, a precision string or a MixedPrecisionPolicy")

    @property
    def is_no_shard(self) -> bool:
        return self.fsdp_version == 1 and str(self.sharding_strategy).upper() == "NO_SHARD"

    @property
    def effective_reshard_after_forward(self) -> bool:
        if self.fsdp_version == 2:
            return bool(self.reshard_after_forward)
        return str(self.reshard_after_forward).upper() in ("FULL_SHARD", "HYBRID_SHARD")


def _call_wrap_policy(policy, module):
    try:
        return policy(module)
    except TypeError:
        nonwrapped = sum(p.numel() for p in module.parameters())
        return policy(module=module, recurse=False, nonwrapped_numel=nonwrapped)


# =============================================================================== TP / CP / SP
@dataclass
class TorchTensorParallelPlugin:
    """Deprecated: use ``ParallelismConfig(tp_size=...)``."""

    tp_size: int = 1
    torch_device_mesh: Optional[Any] = None

    def __post_init__(self):
        warnings.warn("TorchTensorParallelPlugin is deprecated; pass ParallelismConfig(tp_size=...) instead.", FutureWarning)
        if not isinstance(self.tp_size, int) or self.tp_size < 1:
            raise ValueError("tp_size must be a positive integer")


@dataclass
class TorchContextParallelConfig:
    """``cp_comm_strategy``: ``allgather`` (gather all K/V once per attention call) or ``alltoall`` (ring)."""

    cp_comm_strategy: Optional[str] = None

    def __post_init__(self):
        if self.cp_comm_strategy is None:
            self.cp_comm_strategy = os.environ.get("PARALLELISM_CONFIG_CP_COMM_STRATEGY", "allgather")
        if self.cp_comm_strategy not in ("allgather", "alltoall"):
            raise ValueError(f"cp_comm_strategy must be 'allgather' or 'alltoall', got {self.cp_comm_strategy}")


@dataclass
class DeepSpeedSequenceParallelConfig:
    """Ulysses / ALST options. Executed by :mod:`..parallel.ulysses` (no DeepSpeed dependency)."""

    sp_seq_length: Optional[int] = None
    sp_seq_length_is_variable: Optional[bool] = None
    sp_attn_implementation: Optional[str] = None

    def __post_init__(self):
        p = "PARALLELISM_CONFIG_"
        if self.sp_seq_length_is_variable is None:
            self.sp_seq_length_is_variable = os.environ.get(p + "SP_SEQ_LENGTH_IS_VARIABLE", "true").lower() == "true"
        if not self.sp_seq_length_is_variable and self.sp_seq_length is None:
            if p + "SP_SEQ_LENGTH" not in os.environ:
                raise ValueError("`sp_seq_length` must be given when `sp_seq_length_is_variable` is False")
            self.sp_seq_length = os.environ.get(p + "SP_SEQ_LENGTH")
            self.sp_seq_length = None if self.sp_seq_length is None else int(self.sp_seq_length)
        if self.sp_attn_implementation is None:
            self.sp_attn_implementation = os.environ.get(p + "SP_ATTN_IMPLEMENTATION", "sdpa")
        if self.sp_attn_implementation not in ("flash_attention_2", "flash_attention_3", "sdpa", "b200_flash"):
            raise ValueError("sp_attn_implementation must be one of sdpa, flash_attention_2, flash_attention_3, b200_flash")


@dataclass
class TorchTensorParallelConfig:
    """``enable_async_tp`` selects the fused all-gather->GEMM / GEMM->reduce-scatter kernels."""

    enable_async_tp: bool = False

    def __post_init__(self):
        if not isinstance(self.enable_async_tp, bool):
            raise ValueError("enable_async_tp must be a bool")


# =============================================================================== Megatron-LM (config shell)
_MEGATRON_LOGGING_SWITCHES = (
    "log_params_norm", "log_num_zeros_in_grad", "log_throughput", "log_progress", "log_timers_to_tensorboard",
    "log_batch_size_to_tensorboard", "log_learning_rate_to_tensorboard", "log_loss_scale_to_tensorboard", "log_validation_ppl_to_tensorboard",
    "log_memory_to_tensorboard", "log_world_size_to_tensorboard",
)


@dataclass
class MegatronLMPlugin:
    """Megatron-LM option bag.  Megatron itself is not a dependency: tp/cp/sp degrees are mapped
    onto ``ParallelismConfig``; the remaining fields are carried for config compatibility
    (reference src/accelerate/utils/dataclasses.py:2286-2650)."""

    tp_degree: int = None
    pp_degree: int = None
    num_micro_batches: int = None
    gradient_clipping: float = None
    sequence_parallelism: bool = None
    recompute_activations: bool = None
    use_distributed_optimizer: bool = None
    pipeline_model_parallel_split_rank: int = None
    num_layers_per_virtual_pipeline_stage: int = None
    is_train_batch_min: str = True
    train_iters: int = None
    train_samples: int = None
    weight_decay_incr_style: str = "constant"
    start_weight_decay: float = None
    end_weight_decay: float = None
    lr_decay_style: str = "linear"
    lr_decay_iters: int = None
    lr_decay_samples: int = None
    lr_warmup_iters: int = None
    lr_warmup_samples: int = None
    lr_warmup_fraction: float = None
    min_lr: float = 0
    consumed_samples: list = None
    no_wd_decay_cond: Optional[Callable] = None
    scale_lr_cond: Optional[Callable] = None
    lr_mult: float = 1.0
    megatron_dataset_flag: bool = False
    seq_length: int = None
    encoder_seq_length: int = None
    decoder_seq_length: int = None
    tensorboard_dir: str = None
    set_all_logging_options: bool = False
    eval_iters: int = 100
    eval_interval: int = 1000
    return_logits: bool = False
    custom_train_step_class: Optional[Any] = None
    custom_train_step_kwargs: Optional[dict] = None
    custom_model_provider_function: Optional[Callable] = None
    custom_prepare_model_function: Optional[Callable] = None
    custom_megatron_datasets_provider_function: Optional[Callable] = None
    custom_get_batch_function: Optional[Callable] = None
    custom_loss_function: Optional[Callable] = None
    other_megatron_args: Optional[dict] = None
    use_custom_fsdp: bool = None
    overlap_cpu_optimizer_d2h_h2d: bool = None
    no_load_optim: bool = None
    eod_mask_loss: bool = None
    no_save_optim: bool = None
    optimizer_cpu_offload: bool = None
    use_precision_aware_optimizer: bool = None
    decoder_last_pipeline_num_layers: int = None
    recompute_granularity: str = None
    recompute_method: str = None
    recompute_num_layers: int = None
    attention_backend: bool = None
    expert_model_parallel_size: int = None
    context_parallel_size: int = None
    attention_dropout: float = None
    hidden_dropout: float = None
    attention_softmax_in_fp32: bool = None
    expert_tensor_parallel_size: int = None
    calculate_per_token_loss: bool = None
    use_rotary_position_embeddings: bool = None

    def __post_init__(self):
        p = "MEGATRON_LM_"
        if self.tp_degree is None:
            self.tp_degree = int(os.environ.get(p + "TP_DEGREE", 1))
        if self.pp_degree is None:
            self.pp_degree = int(os.environ.get(p + "PP_DEGREE", 1))
        if self.num_micro_batches is None:
            self.num_micro_batches = int(os.environ.get(p + "NUM_MICRO_BATCHES", 1))
        if self.gradient_clipping is None:
            self.gradient_clipping = float(os.environ.get(p + "GRADIENT_CLIPPING", 1.0))
        if self.recompute_activations is None:
            self.recompute_activations = str_to_bool(os.environ.get(p + "RECOMPUTE_ACTIVATIONS", "False")) == 1
        if self.use_distributed_optimizer is None:
            self.use_distributed_optimizer = str_to_bool(os.environ.get(p + "USE_DISTRIBUTED_OPTIMIZER", "False")) == 1
        if self.sequence_parallelism is None:
            self.sequence_parallelism = str_to_bool(os.environ.get(p + "SEQUENCE_PARALLELISM", "False")) == 1
        if self.context_parallel_size is None:
            self.context_parallel_size = int(os.environ.get(p + "CONTEXT_PARALLEL_SIZE", 1))
        if self.expert_model_parallel_size is None:
            self.expert_model_parallel_size = int(os.environ.get(p + "EXPERT_MODEL_PARALLEL_SIZE", 1))
        if self.expert_tensor_parallel_size is None:
            self.expert_tensor_parallel_size = int(os.environ.get(p + "EXPERT_TENSOR_PARALLEL_SIZE", 1))
        # the long tail: booleans, strings and numbers the launcher exports as MEGATRON_LM_<FIELD>
        tail = {"use_custom_fsdp": bool, "overlap_cpu_optimizer_d2h_h2d": bool, "no_load_optim": bool, "eod_mask_loss": bool, "no_save_optim": bool, "optimizer_cpu_offload": bool,
                "use_precision_aware_optimizer": bool, "attention_backend": bool, "attention_softmax_in_fp32": bool,
                "calculate_per_token_loss": bool, "use_rotary_position_embeddings": bool, "decoder_last_pipeline_num_layers": int,
                "recompute_num_layers": int, "recompute_granularity": str, "recompute_method": str, "attention_dropout": float,
                "hidden_dropout": float}
        for name, kind in tail.items():
            raw = os.environ.get(p + name.upper())
            if getattr(self, name, None) is None and raw is not None:
                setattr(self, name, (str_to_bool(raw) == 1) if kind is bool else kind(raw))
        if self.pp_degree > 1 or self.use_distributed_optimizer:
            self.DDP_impl = "local"
        else:
            self.DDP_impl = "torch"
        self.megatron_lm_default_args = {
            "tensor_model_parallel_size": self.tp_degree,
            "pipeline_model_parallel_size": self.pp_degree,
            "context_parallel_size": self.context_parallel_size,
            "expert_model_parallel_size": self.expert_model_parallel_size,
            "sequence_parallel": self.sequence_parallelism,
            "clip_grad": self.gradient_clipping,
            "num_micro_batches": self.num_micro_batches,
            "use_distributed_optimizer": self.use_distributed_optimizer,
            "recompute_activations": self.recompute_activations,
        }
        if self.other_megatron_args:
            self.megatron_lm_default_args.update(self.other_megatron_args)

    # ------------------------------------------------------------------ what a prepared job records on the plugin
    def set_network_size_args(self, model, batch_data=None):
        """Translate ``model.config`` into Megatron's argument names with the parser registered for its ``model_type``."""
        kind = model.config.model_type.lower()
        parser = next((fn for name, fn in MODEL_CONFIGS_TO_MEGATRON_PARSERS.items() if name in kind), None)
        if parser is None:
            raise ValueError(f"Accelerate Megatron-LM integration not supports {kind} model. You can add your own model config parser.")
        parser(self, model, batch_data)

    def set_mixed_precision(self, mixed_precision):
        if mixed_precision == "fp16":
            self.megatron_lm_default_args["fp16"] = True
        elif mixed_precision == "bf16":
            self.megatron_lm_default_args["bf16"] = True
            self.DDP_impl = "local"
            self.megatron_lm_default_args["DDP_impl"] = self.DDP_impl

    def set_training_args(self, micro_batch_size, dp_degree):
        """The batch geometry: global batch = data-parallel ranks x micro batch x micro batches per step."""
        self.data_parallel_size = dp_degree
        self.micro_batch_size = micro_batch_size
        self.global_batch_size = dp_degree * micro_batch_size * self.num_micro_batches
        self.megatron_lm_default_args.update(data_parallel_size=dp_degree, micro_batch_size=micro_batch_size,
                                             global_batch_size=self.global_batch_size)

    def set_optimizer_type(self, optimizer):
        """Record the update rule under Megatron's names (``adam`` / ``sgd`` and their hyper-parameters)."""
        name = type(optimizer).__name__.lower()
        args, defaults = self.megatron_lm_default_args, optimizer.defaults
        if "adam" in name:
            args.update(optimizer="adam", adam_beta1=defaults["betas"][0], adam_beta2=defaults["betas"][1], adam_eps=defaults["eps"])
        elif "sgd" in name:
            args.update(optimizer="sgd", sgd_momentum=defaults["momentum"])
        else:
            raise ValueError(f"Optimizer {name} is not supported by Megatron-LM")
        args.update(lr=defaults["lr"], weight_decay=defaults["weight_decay"])

    def set_scheduler_args(self, scheduler):
        """Iteration counts from a dummy scheduler's step counts (which are per process: divided by the data-parallel size)."""
        ranks = self.megatron_lm_default_args["data_parallel_size"]
        if self.train_iters is None:
            self.train_iters = scheduler.total_num_steps // ranks
            if self.train_samples is not None:
                self.train_samples = None
                warnings.warn("Ignoring `train_samples` as `train_iters` based on scheduler is being used for training.")
        if self.lr_warmup_iters is None:
            self.lr_warmup_iters = scheduler.warmup_num_steps // ranks
            if self.lr_warmup_samples is not None:
                warnings.warn("Ignoring `lr_warmup_samples` as `lr_warmup_iters` based on scheduler is being used for training.")
            self.lr_warmup_samples = 0
        for key in ("train_iters", "lr_warmup_iters", "train_samples", "lr_warmup_samples", "lr_decay_iters", "lr_decay_samples",
                    "lr_warmup_fraction", "lr_decay_style", "weight_decay_incr_style", "start_weight_decay", "end_weight_decay", "min_lr"):
            self.megatron_lm_default_args[key] = getattr(self, key)

    def set_tensorboard_logging_options(self):
        """Switch on every ``log_*`` option Megatron's tensorboard writer knows (the list is fixed here: Megatron is not imported)."""
        for key in _MEGATRON_LOGGING_SWITCHES:
            self.megatron_lm_default_args[key] = True

    def to_parallelism_kwargs(self) -> dict:
        """tp / cp sizes for ``ParallelismConfig`` (pipeline training has no native engine: pp must be 1)."""
        if self.pp_degree and self.pp_degree > 1:
            raise NotImplementedError("pipeline-parallel *training* is not provided by this framework (pp_degree must be 1)")
        return {"tp_size": self.tp_degree, "cp_size": self.context_parallel_size}


# =============================================================================== bitsandbytes (config shell)
@dataclass
class BnbQuantizationConfig:
    load_in_8bit: bool = False
    llm_int8_threshold: float = 6.0
    load_in_4bit: bool = False
    bnb_4bit_quant_type: str = "fp4"
    bnb_4bit_use_double_quant: bool = False
    bnb_4bit_compute_dtype: str = "fp16"
    torch_dtype: torch.dtype = None
    skip_modules: list = None
    keep_in_fp32_modules: list = None

    def __post_init__(self):
        if self.load_in_8bit and self.load_in_4bit:
            raise ValueError("load_in_8bit and load_in_4bit can't be both True")
        if not self.load_in_8bit and not self.load_in_4bit:
            raise ValueError("load_in_8bit and load_in_4bit can't be both False")
        if not isinstance(self.llm_int8_threshold, (int, float)):
            raise ValueError("llm_int8_threshold must be a float or an int")
        if self.bnb_4bit_quant_type not in ("fp4", "nf4"):
            raise ValueError(f"bnb_4bit_quant_type must be in ['fp4','nf4'] but found {self.bnb_4bit_quant_type}")
        if isinstance(self.bnb_4bit_compute_dtype, str):
            table = {"fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16}
            if self.bnb_4bit_compute_dtype not in table:
                raise ValueError("bnb_4bit_compute_dtype must be in ['fp32','fp16','bf16']")
            self.bnb_4bit_compute_dtype = table[self.bnb_4bit_compute_dtype]
        elif not isinstance(self.bnb_4bit_compute_dtype, torch.dtype):
            raise ValueError("bnb_4bit_compute_dtype must be a string or a torch.dtype")
        if self.skip_modules is not None and not isinstance(self.skip_modules, list):
            raise ValueError("skip_modules must be a list of strings")
        if self.keep_in_fp32_modules is not None and not isinstance(self.keep_in_fp32_modules, list):
            raise ValueError("keep_in_fp_32_modules must be a list of strings")
        if self.load_in_4bit:
            self.target_dtype = CustomDtype.INT4
        if self.load_in_8bit:
            self.target_dtype = torch.int8
        if self.load_in_4bit and self.llm_int8_threshold != 6.0:
            warnings.warn("llm_int8_threshold can only be used for model loaded in 8bit")
        if self.torch_dtype is None:
            self.torch_dtype = self.bnb_4bit_compute_dtype if self.load_in_4bit else torch.float16


def get_module_class_from_name(module, name):
    """Depth-first search of ``module`` for a sub-module whose class is called ``name``."""
    if module.__class__.__name__ == name:
        return module.__class__
    for child in module.children():
        found = get_module_class_from_name(child, name)
        if found is not None:
            return found
    return None


MODEL_CONFIGS_TO_MEGATRON_PARSERS = {}


def add_model_config_to_megatron_parser(model_type: str):
    """Register a function translating a transformers config of ``model_type`` into ``MegatronLMPlugin`` arguments
    (reference utils/dataclasses.py:2797-2806)."""

    def add_model_config_parser_helper(func):
        import functools

        @functools.wraps(func)
        def wrapper(*args, **kwargs):
            return func(*args, **kwargs)

        MODEL_CONFIGS_TO_MEGATRON_PARSERS[model_type] = func
        return wrapper

    return add_model_config_parser_helper


@dataclass(frozen=True)
class _MegatronModelSpec:
    """How one model family's config spells the quantities Megatron needs, and which sequence length governs it."""

    model_type_name: str                 # Megatron's name of the family: bert / gpt / t5
    layers: str                          # config attribute names ...
    hidden: str
    heads: str
    positions: str
    seq_field: Optional[str]             # plugin field that overrides `seq_length` (None: encoder/decoder lengths are kept apart)
    pretraining: Optional[bool] = True   # None: decided by the class name (masked-LM head = pre-training)
    extras: tuple = ()                   # (megatron key, source) pairs; source is a config attribute or a literal in a 1-tuple
    default_positions: int = 1024
    config_seq: Optional[str] = None     # config attribute that, when present, beats the plugin's family-specific length
    derive: Optional[Callable] = None    # config -> dict of entries that are computed rather than read


# GLM-4.5 mixture-of-experts (reference utils/dataclasses.py:2940-3021): a GPT-family decoder with grouped-query attention, half-rotary
# RoPE, sigmoid-scored routing with an expert bias, and shared experts of the routed experts' width.
_GLM4_MOE_EXTRAS = (
    ("tokenizer_type", ("HuggingFaceTokenizer",)), ("kv_channels", "head_dim"), ("original_model_type", "model_type"),
    ("position_embedding_type", ("rope",)), ("rope_type", ("rope",)), ("rotary_base", "rope_theta"),
    ("rotary-percent", (0.5,)), ("rotary_percent", "partial_rotary_factor"),
    ("qk_layernorm", "use_qk_norm"), ("add_bias_linear", (False,)), ("add_qkv_bias", (True,)),
    ("group_query_attention", (True,)), ("num_query_groups", "num_key_value_heads"), ("ffn_hidden_size", "intermediate_size"),
    ("normalization", ("RMSNorm",)), ("norm_epsilon", (1e-3,)), ("swiglu", (True,)), ("use_flash_attn", (True,)),
    ("moe_ffn_hidden_size", "moe_intermediate_size"), ("moe_shared_expert_intermediate_size", "moe_intermediate_size"),
    ("moe_router_pre_softmax", (True,)), ("moe_router_score_function", ("sigmoid",)), ("moe_router_enable_expert_bias", (True,)),
    ("moe_router_bias_update_rate", (0,)), ("moe_router_load_balancing_type", ("seq_aux_loss",)),
    ("moe_token_dispatcher_type", ("alltoall",)), ("moe_router_topk", "num_experts_per_tok"),
    ("moe_router_topk_scaling_factor", "routed_scaling_factor"), ("num_experts", "n_routed_experts"),
    ("moe_grouped_gemm", (True,)), ("moe_router_dtype", ("fp32",)), ("moe_permute_fusion", (True,)), ("moe_aux_loss_coeff", (0,)),
    ("eos_token_id", "eos_token_id"), ("bf16", "bf16"),
)


def _glm4_moe_derived(config) -> dict:
    dense = int(getattr(config, "first_k_dense_replace", 0))
    out = {"moe_layer_freq": [0] * dense + [1] * (config.num_hidden_layers - dense),
           "untie_embeddings_and_output_weights": not getattr(config, "tie_word_embeddings", False)}
    if getattr(config, "fp8_param", False):
        out.update({key: getattr(config, key) for key in ("fp8", "fp8_param", "fp8_param_gather", "fp8_recipe")})
    return out


_MEGATRON_MODEL_SPECS = {
    "megatron-bert": _MegatronModelSpec("bert", "num_hidden_layers", "hidden_size", "num_attention_heads", "max_position_embeddings",
                                        "encoder_seq_length", pretraining=None, extras=(("num_labels", "num_labels"),)),
    "bert": _MegatronModelSpec("bert", "num_hidden_layers", "hidden_size", "num_attention_heads", "max_position_embeddings",
                               "encoder_seq_length", pretraining=None, extras=(("num_labels", "num_labels"),)),
    "gpt2": _MegatronModelSpec("gpt", "n_layer", "n_embd", "n_head", "n_positions", "decoder_seq_length",
                               extras=(("tokenizer_type", ("GPT2BPETokenizer",)),)),
    "t5": _MegatronModelSpec("t5", "num_layers", "d_model", "num_heads", "n_positions", None),
    "llama": _MegatronModelSpec("gpt", "num_hidden_layers", "hidden_size", "num_attention_heads", "max_position_embeddings",
                                "decoder_seq_length",
                                extras=(("tokenizer_type", ("Llama2Tokenizer",)), ("num_query_groups", "num_key_value_heads"),
                                        ("ffn_hidden_size", "intermediate_size"), ("norm_epsilon", "rms_norm_eps"),
                                        ("rotary_base", "rope_theta"), ("swiglu", (True,)), ("normalization", ("RMSNorm",)),
                                        ("position_embedding_type", ("rope",)), ("untie_embeddings_and_output_weights", (True,)))),
    "glm4_moe": _MegatronModelSpec("gpt", "num_hidden_layers", "hidden_size", "num_attention_heads", "max_position_embeddings",
                                   "decoder_seq_length", pretraining=False, config_seq="max_sequence_length",
                                   extras=_GLM4_MOE_EXTRAS, derive=_glm4_moe_derived),
}


def _describe_for_megatron(spec: _MegatronModelSpec, plugin, model, batch_data) -> None:
    """Fill ``plugin.megatron_lm_default_args`` (and the plugin's sequence lengths) from ``model.config`` according to ``spec``."""
    config = model.config
    out = plugin.megatron_lm_default_args
    positions = getattr(config, spec.positions, None) or spec.default_positions

    def length_from(batch_key, fallback):
        if batch_data is not None and batch_key in batch_data:
            return batch_data[batch_key].shape[1]
        return fallback

    if spec.seq_field is None:
        # encoder-decoder: the two lengths stay separate
        if plugin.encoder_seq_length is None:
            plugin.encoder_seq_length = length_from("input_ids", positions)
        if plugin.decoder_seq_length is None:
            plugin.decoder_seq_length = length_from("labels", positions)
        out["encoder_seq_length"], out["decoder_seq_length"] = plugin.encoder_seq_length, plugin.decoder_seq_length
    else:
        specific = getattr(plugin, spec.seq_field)
        from_config = getattr(config, spec.config_seq, None) if spec.config_seq else None
        if spec.config_seq is not None:
            # this family settles the length once: plugin value, else the config's own, else the family field, the batch, the positions
            if plugin.seq_length is None:
                plugin.seq_length = from_config or specific or length_from("input_ids", positions)
        elif specific is not None:
            if plugin.seq_length is not None and plugin.seq_length != specific:
                warnings.warn(f"Both `seq_length` and `{spec.seq_field}` are set. Using `{spec.seq_field}`.")
            plugin.seq_length = specific
        elif plugin.seq_length is None:
            plugin.seq_length = length_from("input_ids", positions)
        out["seq_length"] = plugin.seq_length
    pretraining = spec.pretraining
    if pretraining is None:
        pretraining = "maskedlm" in type(model).__name__.lower()
    out.update(model_type_name=spec.model_type_name, num_layers=getattr(config, spec.layers), hidden_size=getattr(config, spec.hidden),
               num_attention_heads=getattr(config, spec.heads), max_position_embeddings=positions, pretraining_flag=pretraining,
               orig_vocab_size=config.vocab_size, model_return_dict=getattr(config, "return_dict", True))
    if spec.model_type_name == "gpt":
        out["return_logits"] = plugin.return_logits
    for key, source in spec.extras:
        if isinstance(source, tuple):
            out[key] = source[0]
        elif getattr(config, source, None) is not None:
            out[key] = getattr(config, source)
    if spec.derive is not None:
        out.update(spec.derive(config))


def _register_megatron_parsers() -> None:
    for model_type, spec in _MEGATRON_MODEL_SPECS.items():
        def parser(megatron_lm_plugin, model, batch_data, _spec=spec):
            _describe_for_megatron(_spec, megatron_lm_plugin, model, batch_data)

        parser.__name__ = f"parse_{model_type.replace('-', '_')}_config"
        add_model_config_to_megatron_parser(model_type)(parser)


_register_megatron_parsers()
parse_bert_config = MODEL_CONFIGS_TO_MEGATRON_PARSERS["megatron-bert"]
parse_gpt2_config = MODEL_CONFIGS_TO_MEGATRON_PARSERS["gpt2"]
parse_t5_config = MODEL_CONFIGS_TO_MEGATRON_PARSERS["t5"]
parse_llama_config = MODEL_CONFIGS_TO_MEGATRON_PARSERS["llama"]
parse_glm4_moe_config = MODEL_CONFIGS_TO_MEGATRON_PARSERS["glm4_moe"]
