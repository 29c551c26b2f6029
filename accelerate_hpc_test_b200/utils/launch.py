"""From launch arguments to a command line and an environment.

``accelerate launch`` never imports the training script: every setting crosses the process boundary as an environment
variable that ``PartialState`` / ``AcceleratorState`` / the plugin dataclasses read back.  This module
is that translation.  Most of it is tabular — *flag X becomes variable Y, formatted like Z* — so the mapping tables below
are the single place where a flag is wired (variable names are the interface shared with the reference,
``utils/launch.py:99-583``); the functions only decide which tables apply.

Not carried over: TPU, SageMaker and vendor-CCL launch paths (they raise).
"""

from __future__ import annotations

import argparse
import os
import subprocess
import sys
import warnings
from shutil import which
from typing import Callable

import torch

from .constants import DEEPSPEED_MULTINODE_LAUNCHERS
from .dataclasses import DistributedType, DynamoBackend, PrecisionType
from .other import is_port_in_use


# --------------------------------------------------------------------------- small formatters
def _text(value) -> str:
    return str(value)


def _lower(value) -> str:
    return str(value).lower()


def _truthy(value) -> bool:
    """Flags such as ``--fsdp_sync_module_states`` arrive as text ("true" / "false") from the CLI and as bool from YAML."""
    if isinstance(value, str):
        return value.strip().lower() in ("1", "true", "yes", "y", "on")
    return bool(value)


class EnvRule:
    """``args.<flag>`` → ``env[<variable>]`` through ``fmt``; skipped when the flag is absent or (``optional``) None."""

    __slots__ = ("flag", "variable", "fmt", "optional")

    def __init__(self, flag: str, variable: str, fmt: Callable = _text, optional: bool = False):
        self.flag, self.variable, self.fmt, self.optional = flag, variable, fmt, optional

    def apply(self, args, env: dict) -> None:
        if not hasattr(args, self.flag):
            return
        value = getattr(args, self.flag)
        if value is None and self.optional:
            return
        env[self.variable] = self.fmt(value)


def _apply(rules, args, env: dict) -> None:
    for rule in rules:
        rule.apply(args, env)


_FSDP_RULES = (
    EnvRule("fsdp_sharding_strategy", "FSDP_SHARDING_STRATEGY", optional=True),  # FSDP1 spelling
    EnvRule("fsdp_reshard_after_forward", "FSDP_RESHARD_AFTER_FORWARD", _lower, optional=True),  # FSDP2 spelling
    EnvRule("fsdp_offload_params", "FSDP_OFFLOAD_PARAMS", _lower),
    EnvRule("fsdp_min_num_params", "FSDP_MIN_NUM_PARAMS"),
    EnvRule("fsdp_auto_wrap_policy", "FSDP_AUTO_WRAP_POLICY", optional=True),
    EnvRule("fsdp_transformer_layer_cls_to_wrap", "FSDP_TRANSFORMER_CLS_TO_WRAP", optional=True),
    EnvRule("fsdp_backward_prefetch", "FSDP_BACKWARD_PREFETCH", optional=True),
    EnvRule("fsdp_state_dict_type", "FSDP_STATE_DICT_TYPE", optional=True),
    EnvRule("fsdp_forward_prefetch", "FSDP_FORWARD_PREFETCH", _lower),
    EnvRule("fsdp_use_orig_params", "FSDP_USE_ORIG_PARAMS", _lower),
    EnvRule("fsdp_cpu_ram_efficient_loading", "FSDP_CPU_RAM_EFFICIENT_LOADING", _lower),
    EnvRule("fsdp_sync_module_states", "FSDP_SYNC_MODULE_STATES", _lower),
    EnvRule("fsdp_activation_checkpointing", "FSDP_ACTIVATION_CHECKPOINTING", _lower),
    EnvRule("fsdp_ignored_modules", "FSDP_IGNORED_MODULES", optional=True),
)

_MEGATRON_RULES = (
    EnvRule("megatron_lm_tp_degree", "MEGATRON_LM_TP_DEGREE"),
    EnvRule("megatron_lm_pp_degree", "MEGATRON_LM_PP_DEGREE"),
    EnvRule("megatron_lm_gradient_clipping", "MEGATRON_LM_GRADIENT_CLIPPING"),
    EnvRule("megatron_lm_num_micro_batches", "MEGATRON_LM_NUM_MICRO_BATCHES", optional=True),
    EnvRule("megatron_lm_sequence_parallelism", "MEGATRON_LM_SEQUENCE_PARALLELISM", optional=True),
    EnvRule("megatron_lm_recompute_activations", "MEGATRON_LM_RECOMPUTE_ACTIVATIONS", optional=True),
    EnvRule("megatron_lm_use_distributed_optimizer", "MEGATRON_LM_USE_DISTRIBUTED_OPTIMIZER", optional=True),
) + tuple(
    EnvRule(f"megatron_lm_{name}", f"MEGATRON_LM_{name.upper()}", optional=True)
    for name in ("use_custom_fsdp", "no_load_optim", "no_save_optim", "eod_mask_loss", "optimizer_cpu_offload", "use_precision_aware_optimizer",
                 "overlap_cpu_optimizer_d2h_h2d", "decoder_last_pipeline_num_layers", "recompute_granularity", "recompute_method",
                 "recompute_num_layers", "attention_backend", "expert_model_parallel_size", "expert_tensor_parallel_size",
                 "context_parallel_size", "attention_dropout", "hidden_dropout", "attention_softmax_in_fp32", "calculate_per_token_loss",
                 "use_rotary_position_embeddings")
)

_PARALLELISM_RULES = (
    EnvRule("parallelism_config_dp_replicate_size", "PARALLELISM_CONFIG_DP_REPLICATE_SIZE"),
    EnvRule("parallelism_config_dp_shard_size", "PARALLELISM_CONFIG_DP_SHARD_SIZE"),
    EnvRule("parallelism_config_tp_size", "PARALLELISM_CONFIG_TP_SIZE"),
    EnvRule("parallelism_config_cp_size", "PARALLELISM_CONFIG_CP_SIZE"),
    EnvRule("parallelism_config_cp_backend", "PARALLELISM_CONFIG_CP_BACKEND"),
    EnvRule("parallelism_config_sp_size", "PARALLELISM_CONFIG_SP_SIZE"),
    EnvRule("parallelism_config_sp_backend", "PARALLELISM_CONFIG_SP_BACKEND"),
)
_CP_ONLY_RULES = (EnvRule("parallelism_config_cp_comm_strategy", "PARALLELISM_CONFIG_CP_COMM_STRATEGY"),)
_SP_ONLY_RULES = (
    EnvRule("parallelism_config_sp_seq_length", "PARALLELISM_CONFIG_SP_SEQ_LENGTH"),
    EnvRule("parallelism_config_sp_seq_length_is_variable", "PARALLELISM_CONFIG_SP_SEQ_LENGTH_IS_VARIABLE", _lower),
    EnvRule("parallelism_config_sp_attn_implementation", "PARALLELISM_CONFIG_SP_ATTN_IMPLEMENTATION"),
)

_DEEPSPEED_RULES = (
    EnvRule("zero_stage", "ACCELERATE_DEEPSPEED_ZERO_STAGE", optional=True),
    EnvRule("gradient_accumulation_steps", "ACCELERATE_GRADIENT_ACCUMULATION_STEPS", optional=True),
    EnvRule("gradient_clipping", "ACCELERATE_GRADIENT_CLIPPING", _lower, optional=True),
    EnvRule("offload_optimizer_device", "ACCELERATE_DEEPSPEED_OFFLOAD_OPTIMIZER_DEVICE", _lower, optional=True),
    EnvRule("offload_param_device", "ACCELERATE_DEEPSPEED_OFFLOAD_PARAM_DEVICE", _lower, optional=True),
    EnvRule("zero3_init_flag", "ACCELERATE_DEEPSPEED_ZERO3_INIT", _lower, optional=True),
    EnvRule("zero3_save_16bit_model", "ACCELERATE_DEEPSPEED_ZERO3_SAVE_16BIT_MODEL", _lower, optional=True),
    EnvRule("deepspeed_config_file", "ACCELERATE_DEEPSPEED_CONFIG_FILE", optional=True),
    EnvRule("deepspeed_moe_layer_cls_names", "ACCELERATE_DEEPSPEED_MOE_LAYER_CLS_NAMES", optional=True),
)

_DYNAMO_RULES = (
    EnvRule("dynamo_mode", "ACCELERATE_DYNAMO_MODE", lambda v: v),
    EnvRule("dynamo_use_fullgraph", "ACCELERATE_DYNAMO_USE_FULLGRAPH"),
    EnvRule("dynamo_use_dynamic", "ACCELERATE_DYNAMO_USE_DYNAMIC"),
    EnvRule("dynamo_use_regional_compilation", "ACCELERATE_DYNAMO_USE_REGIONAL_COMPILATION"),
)


# --------------------------------------------------------------------------- pieces shared by every launcher
def _filter_args(args: argparse.Namespace, parser, default_args=()):
    """A namespace for ``parser`` (e.g. ``torch.distributed.run``'s) holding its defaults overridden by the attributes of
    ``args`` that it knows; everything else in ``args`` is dropped."""
    accepted, _ = parser.parse_known_args(list(default_args))
    for name in vars(accepted):
        if hasattr(args, name):
            setattr(accepted, name, getattr(args, name))
    return accepted


def setup_fp8_env(args: argparse.Namespace, current_env: dict) -> dict:
    """Every ``--fp8_*`` flag that was given becomes ``ACCELERATE_FP8_*`` (the linear-precision triple becomes three variables)."""
    for flag, value in vars(args).items():
        if not flag.startswith("fp8_") or value is None:
            continue
        if flag == "fp8_override_linear_precision":
            for part, item in zip(("FPROP", "DGRAD", "WGRAD"), value):
                current_env[f"ACCELERATE_FP8_OVERRIDE_{part}"] = str(item)
        else:
            current_env[f"ACCELERATE_{flag.upper()}"] = str(value)
    return current_env


def _precision_and_compile_env(args, env: dict) -> None:
    wanted = str(args.mixed_precision).lower()
    try:
        env["ACCELERATE_MIXED_PRECISION"] = str(PrecisionType(wanted))
    except ValueError:
        raise ValueError(f"Unknown mixed_precision mode: {wanted}. Choose between {PrecisionType.list()}.") from None
    if wanted == "fp8":
        setup_fp8_env(args, env)
    backend = str(args.dynamo_backend).upper()
    try:
        env["ACCELERATE_DYNAMO_BACKEND"] = DynamoBackend(backend).value
    except ValueError:
        raise ValueError(f"Unknown dynamo backend: {backend}. Choose between {DynamoBackend.list()}.") from None
    _apply(_DYNAMO_RULES, args, env)


def _common_env(args) -> dict:
    env = os.environ.copy()
    if args.debug:
        env["ACCELERATE_DEBUG_MODE"] = "true"
    visible = getattr(args, "gpu_ids", "all")
    if visible not in ("all", None):
        env["CUDA_VISIBLE_DEVICES"] = visible
    _precision_and_compile_env(args, env)
    return env


def _check_module_flags(args) -> None:
    if args.module and args.no_python:
        raise ValueError("--module and --no_python cannot be used together")


# --------------------------------------------------------------------------- single process / mpirun
def _get_mpirun_args():
    """(launcher binary, hostfile flag, total-process flag, per-node flag, binding flag) of the installed MPI flavour."""
    found = next((exe for exe in ("mpirun", "mpiexec") if which(exe)), None)
    if found is None:
        raise OSError("mpirun or mpiexec were not found. Ensure that Intel MPI, Open MPI, or MVAPICH are installed.")
    banner = subprocess.check_output([found, "--version"])
    if b"Open MPI" in banner:
        return found, "--hostfile", "-n", "--npernode", "--bind-to"
    return found, "-f", "-n", "-ppn", ""


def _mpirun_prefix(args) -> list:
    exe, hostfile_flag, total_flag, per_node_flag, bind_flag = _get_mpirun_args()
    total, machines = getattr(args, "num_processes", None), args.num_machines
    per_node = str(total // machines) if total and machines else "1"
    prefix = [exe, hostfile_flag, args.mpirun_hostfile, per_node_flag, per_node]
    if total:
        prefix += [total_flag, str(total)]
    if bind_flag:
        prefix += [bind_flag, "socket", "-x", "MASTER_ADDR", "-x", "MASTER_PORT"]
    return prefix


def prepare_simple_launcher_cmd_env(args: argparse.Namespace) -> tuple[list[str], dict[str, str]]:
    """``python script.py ...`` for one process, or the same behind ``mpirun`` for a multi-CPU job with a hostfile."""
    _check_module_flags(args)
    via_mpi = args.mpirun_hostfile is not None
    cmd = _mpirun_prefix(args) if via_mpi else []
    if not args.no_python:
        cmd.append(sys.executable)
        if args.module:
            cmd.append("-m")
    cmd += [args.training_script, *args.training_script_args]

    env = _common_env(args)
    env["ACCELERATE_USE_CPU"] = str(args.cpu or getattr(args, "use_cpu", False))
    if args.num_machines > 1:
        env["MASTER_ADDR"], env["MASTER_PORT"] = args.main_process_ip, str(args.main_process_port)
    elif via_mpi:
        env["MASTER_ADDR"], env["MASTER_PORT"] = args.main_process_ip or "127.0.0.1", str(args.main_process_port or 29500)
    env["OMP_NUM_THREADS"] = str(getattr(args, "num_cpu_threads_per_process", 1) or 1)
    if getattr(args, "enable_cpu_affinity", False):
        env["ACCELERATE_CPU_AFFINITY"] = "1"
    return cmd, env


# --------------------------------------------------------------------------- torchrun
_PORT_BUSY_ONE_NODE = (
    "Port `{port}` is already in use. Accelerate will attempt to launch in a standalone-like mode by finding an open port "
    "automatically for this session. If this current attempt fails, or for more control in future runs, please specify a different "
    "port (e.g., `--main_process_port <your_chosen_port>`) in your launch command or Accelerate config file."
)
_PORT_BUSY_MULTI_NODE = (
    "Tried to launch distributed communication on port `{port}`, but another process is utilizing it. Please specify a different "
    "port (such as using the `--main_process_port` flag or specifying a different `main_process_port` in your config file) and "
    "rerun your script. To automatically use the next open port (on a single node), you can set this to `0`."
)


def _set_rendezvous(args) -> None:
    """Fill the attributes ``torch.distributed.run`` reads (``nproc_per_node``, ``nnodes``, ``master_*`` / ``rdzv_endpoint``)."""
    machines, total = args.num_machines, args.num_processes
    host, port = args.main_process_ip, args.main_process_port
    if machines > 1:
        args.nproc_per_node, args.nnodes, args.node_rank = str(total // machines), str(machines), int(args.machine_rank)
        if getattr(args, "same_network", False):
            args.master_addr, args.master_port = str(host), str(port)
        else:
            args.rdzv_endpoint = f"{host}:{port}"
    else:
        args.nproc_per_node = str(total)
        args.master_addr = str(host) if host else "127.0.0.1"
        if port is not None:
            args.master_port = str(port)
    probe = 29500 if port is None else port
    first_node = machines <= 1 or int(args.machine_rank) == 0
    if first_node and is_port_in_use(probe):
        if machines > 1:
            raise ConnectionError(_PORT_BUSY_MULTI_NODE.format(port=probe))
        args.standalone = True  # torchrun then picks a free port itself
        warnings.warn(_PORT_BUSY_ONE_NODE.format(port=probe), UserWarning)


def prepare_multi_gpu_env(args: argparse.Namespace) -> dict[str, str]:
    """Completes ``args`` for ``torch.distributed.run`` and returns the environment of a multi-GPU (DDP / FSDP / Megatron) job."""
    _set_rendezvous(args)
    _check_module_flags(args)
    args.module, args.no_python = bool(args.module), bool(args.no_python)

    env = _common_env(args)
    if args.use_fsdp:
        if _truthy(args.fsdp_cpu_ram_efficient_loading) and not _truthy(args.fsdp_sync_module_states):
            raise ValueError("When using `--fsdp_cpu_ram_efficient_loading` set `--fsdp_sync_module_states` to `True`")
        env["ACCELERATE_USE_FSDP"] = "true"
        env["FSDP_VERSION"] = str(getattr(args, "fsdp_version", 1))
        _apply(_FSDP_RULES, args, env)
    if args.use_megatron_lm:
        env["ACCELERATE_USE_MEGATRON_LM"] = "true"
        _apply(_MEGATRON_RULES, args, env)
    env["OMP_NUM_THREADS"] = str(args.num_cpu_threads_per_process)
    if args.enable_cpu_affinity:
        env["ACCELERATE_CPU_AFFINITY"] = "1"
    if args.use_parallelism_config:
        prepare_extend_env_parallelism_config(args, env)
    return env


def prepare_extend_env_parallelism_config(args: argparse.Namespace, current_env: dict) -> None:
    """``--parallelism_config_*`` → ``PARALLELISM_CONFIG_*`` (in place); the cp / sp detail variables only when that axis is used."""
    current_env["ACCELERATE_USE_PARALLELISM_CONFIG"] = "true"
    _apply(_PARALLELISM_RULES, args, current_env)
    if args.parallelism_config_cp_size > 1:
        _apply(_CP_ONLY_RULES, args, current_env)
    if args.parallelism_config_sp_size > 1:
        _apply(_SP_ONLY_RULES, args, current_env)


def prepare_deepspeed_cmd_env(args: argparse.Namespace) -> tuple[list[str], dict[str, str]]:
    """A DeepSpeed-configured job.  DeepSpeed is a configuration dialect here (ZeRO stages map onto the native sharded engine),
    so the job is a torchrun job whose environment additionally carries the ``ACCELERATE_DEEPSPEED_*`` variables the plugin
    reads.  Returns ``([], env)``: no external launcher command is needed."""
    launcher = getattr(args, "deepspeed_multinode_launcher", None)
    if launcher is not None and launcher not in DEEPSPEED_MULTINODE_LAUNCHERS:
        raise ValueError(f"Unknown multinode launcher {launcher}; choose from {DEEPSPEED_MULTINODE_LAUNCHERS}")
    env = prepare_multi_gpu_env(args)
    env["ACCELERATE_USE_DEEPSPEED"] = "true"
    # which DeepSpeed settings came from the accelerate config file: next to a `deepspeed_config_file` the plugin refuses the ones
    # that file is supposed to state itself
    env["ACCELERATE_CONFIG_DS_FIELDS"] = str(getattr(args, "deepspeed_fields_from_accelerate_config", "") or "").lower()
    _apply(_DEEPSPEED_RULES, args, env)
    return [], env


def prepare_tpu(args, current_env, pod: bool = False):
    raise NotImplementedError("TPU launch is out of scope for the B200 build (SURVEY.md §7.1)")


def prepare_sagemager_args_inputs(sagemaker_config, args):
    raise NotImplementedError("SageMaker launch is out of scope for the B200 build (SURVEY.md §7.1)")


def env_var_path_add(env_var_name: str, path_to_add: str) -> str:
    """Value of the ``:``-separated variable ``env_var_name`` with ``path_to_add`` appended (empty entries dropped)."""
    entries = [entry for entry in os.environ.get(env_var_name, "").split(":") if entry]
    return ":".join(entries + [str(path_to_add)])


# --------------------------------------------------------------------------- in-process (forked) workers
class PrepareForLaunch:
    """The callable ``notebook_launcher`` / ``debug_launcher`` hand to ``torch.multiprocessing.start_processes``: it gives the
    forked worker its rank variables (and, in debug mode, a gloo group rendezvousing over a file), then runs ``launcher``."""

    def __init__(self, launcher, distributed_type="NO", debug: bool = False):
        self.launcher = launcher
        self.distributed_type = DistributedType(distributed_type)
        self.debug = debug

    def _debug_group(self, index: int) -> None:
        world = int(os.environ["WORLD_SIZE"])
        store = torch.distributed.FileStore(os.environ["ACCELERATE_DEBUG_RDV_FILE"], world)
        os.environ["RANK"] = os.environ["LOCAL_RANK"] = str(index)
        torch.distributed.init_process_group("gloo", rank=index, world_size=world, store=store)

    def __call__(self, index, *args):
        if self.debug:
            self._debug_group(index)
        elif self.distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
            per_node, node = int(os.environ.get("NPROC", 1)), int(os.environ.get("NODE_RANK", 0))
            os.environ["LOCAL_RANK"] = str(index)
            os.environ["RANK"] = str(per_node * node + index)
        os.environ["FORK_LAUNCHED"] = "1"
        self.launcher(*args)
