"""Turning launch arguments into commands and environment variables.

The launcher never imports the training script: everything crosses the process boundary as
environment variables read by ``PartialState`` / the plugins.  Parity with reference
src/accelerate/utils/launch.py: ``_filter_args`` :46-54, ``setup_fp8_env`` :81-96,
``prepare_simple_launcher_cmd_env`` :99-195, ``prepare_multi_gpu_env`` :198-394,
``prepare_extend_env_parallelism_config`` :397-421, ``prepare_deepspeed_cmd_env`` :424-583,
``PrepareForLaunch`` :776-819.  TPU / SageMaker / vendor-CCL branches are not carried over.
"""

from __future__ import annotations

import argparse
import os
import subprocess
import sys
import warnings
from ast import literal_eval
from shutil import which
from typing import Any

import torch

from .constants import DEEPSPEED_MULTINODE_LAUNCHERS
from .dataclasses import DistributedType
from .environment import patch_environment
from .other import get_free_port, is_port_in_use


def _filter_args(args: argparse.Namespace, parser, default_args=()):
    """Keep only the attributes of ``args`` that ``parser`` (e.g. torch.distributed.run's) knows."""
    new_args, _ = parser.parse_known_args(list(default_args))
    for key, value in vars(args).items():
        if key in vars(new_args).keys():
            setattr(new_args, key, value)
    return new_args


def _get_mpirun_args():
    """(mpirun binary, hostfile flag, -n flag, -ppn flag, bind-to flag) for the MPI flavour installed."""
    mpi_apps = [x for x in ["mpirun", "mpiexec"] if which(x)]
    if len(mpi_apps) == 0:
        raise OSError("mpirun or mpiexec were not found. Ensure that Intel MPI, Open MPI, or MVAPICH are installed.")
    mpi_app = mpi_apps[0]
    mpirun_version = subprocess.check_output([mpi_app, "--version"])
    if b"Open MPI" in mpirun_version:
        return mpi_app, "--hostfile", "-n", "--npernode", "--bind-to"
    return mpi_app, "-f", "-n", "-ppn", ""


def setup_fp8_env(args: argparse.Namespace, current_env: dict):
    """``--fp8_*`` flags -> ``ACCELERATE_FP8_*`` variables."""
    prefix = "ACCELERATE_"
    for arg in vars(args):
        if arg.startswith("fp8_"):
            value = getattr(args, arg)
            if value is not None:
                if arg == "fp8_override_linear_precision":
                    current_env[prefix + "FP8_OVERRIDE_FPROP"] = str(value[0])
                    current_env[prefix + "FP8_OVERRIDE_DGRAD"] = str(value[1])
                    current_env[prefix + "FP8_OVERRIDE_WGRAD"] = str(value[2])
                else:
                    current_env[f"{prefix}{arg.upper()}"] = str(getattr(args, arg))
    return current_env


def _mixed_precision(args) -> str:
    try:
        from .dataclasses import PrecisionType

        mixed_precision = PrecisionType(str(args.mixed_precision).lower())
    except ValueError:
        from .dataclasses import PrecisionType

        raise ValueError(f"Unknown mixed_precision mode: {str(args.mixed_precision).lower()}. Choose between {PrecisionType.list()}.")
    return str(mixed_precision)


def _dynamo_env(args, current_env):
    from .dataclasses import DynamoBackend

    try:
        dynamo_backend = DynamoBackend(str(args.dynamo_backend).upper())
    except ValueError:
        raise ValueError(f"Unknown dynamo backend: {str(args.dynamo_backend).upper()}. Choose between {DynamoBackend.list()}.")
    current_env["ACCELERATE_DYNAMO_BACKEND"] = dynamo_backend.value
    current_env["ACCELERATE_DYNAMO_MODE"] = args.dynamo_mode
    current_env["ACCELERATE_DYNAMO_USE_FULLGRAPH"] = str(args.dynamo_use_fullgraph)
    current_env["ACCELERATE_DYNAMO_USE_DYNAMIC"] = str(args.dynamo_use_dynamic)
    current_env["ACCELERATE_DYNAMO_USE_REGIONAL_COMPILATION"] = str(args.dynamo_use_regional_compilation)


def prepare_simple_launcher_cmd_env(args: argparse.Namespace) -> tuple[list[str], dict[str, str]]:
    """Command + env of a single-process (or mpirun multi-CPU) launch."""
    cmd = []
    if args.no_python and args.module:
        raise ValueError("--module and --no_python cannot be used together")
    num_processes = getattr(args, "num_processes", None)
    num_machines = args.num_machines
    if args.mpirun_hostfile is not None:
        mpi_app_name, hostfile_arg, num_proc_arg, proc_per_node_arg, bind_to_arg = _get_mpirun_args()
        bind_to = getattr(args, "mpirun_ccl", None)
        nproc_per_node = str(num_processes // num_machines) if num_processes and num_machines else "1"
        cmd += [mpi_app_name, hostfile_arg, args.mpirun_hostfile, proc_per_node_arg, nproc_per_node]
        if num_processes:
            cmd += [num_proc_arg, str(num_processes)]
        if bind_to_arg:
            cmd += [bind_to_arg, "socket", "-x", "MASTER_ADDR", "-x", "MASTER_PORT"]
    if not args.no_python:
        cmd.append(sys.executable)
        if args.module:
            cmd.append("-m")
    cmd.append(args.training_script)
    cmd.extend(args.training_script_args)

    current_env = os.environ.copy()
    current_env["ACCELERATE_USE_CPU"] = str(args.cpu or getattr(args, "use_cpu", False))
    if args.debug:
        current_env["ACCELERATE_DEBUG_MODE"] = "true"
    if args.gpu_ids != "all" and args.gpu_ids is not None:
        current_env["CUDA_VISIBLE_DEVICES"] = args.gpu_ids
    if num_machines > 1:
        current_env["MASTER_ADDR"] = args.main_process_ip
        current_env["MASTER_PORT"] = str(args.main_process_port)
    elif args.mpirun_hostfile is not None:
        current_env["MASTER_ADDR"] = args.main_process_ip or "127.0.0.1"
        current_env["MASTER_PORT"] = str(args.main_process_port or 29500)
    current_env["ACCELERATE_MIXED_PRECISION"] = _mixed_precision(args)
    if str(args.mixed_precision).lower() == "fp8":
        setup_fp8_env(args, current_env)
    _dynamo_env(args, current_env)
    current_env["OMP_NUM_THREADS"] = str(getattr(args, "num_cpu_threads_per_process", 1) or 1)
    if getattr(args, "enable_cpu_affinity", False):
        current_env["ACCELERATE_CPU_AFFINITY"] = "1"
    return cmd, current_env


def prepare_multi_gpu_env(args: argparse.Namespace) -> dict[str, str]:
    """Sets the torchrun-facing attributes on ``args`` and returns the env of a multi-GPU launch."""
    num_processes = args.num_processes
    num_machines = args.num_machines
    main_process_ip = args.main_process_ip
    main_process_port = args.main_process_port
    if num_machines > 1:
        args.nproc_per_node = str(num_processes // num_machines)
        args.nnodes = str(num_machines)
        args.node_rank = int(args.machine_rank)
        if getattr(args, "same_network", False):
            args.master_addr = str(main_process_ip)
            args.master_port = str(main_process_port)
        else:
            args.rdzv_endpoint = f"{main_process_ip}:{main_process_port}"
    else:
        args.nproc_per_node = str(num_processes)
        if main_process_port is not None:
            args.master_port = str(main_process_port)
        args.master_addr = str(main_process_ip) if main_process_ip else "127.0.0.1"
    if main_process_port is None:
        main_process_port = 29500
    # a busy port on a single node: let torchrun pick one (--standalone) instead of failing later
    need_port_check = num_machines <= 1 or int(args.machine_rank) == 0
    if need_port_check and is_port_in_use(main_process_port):
        if num_machines <= 1:
            args.standalone = True
            warnings.warn(f"Port `{main_process_port}` is already in use. Accelerate will attempt to launch in a standalone-like mode by finding an open port automatically for this session. If this current attempt fails, or for more control in future runs, please specify a different port (e.g., `--main_process_port <your_chosen_port>`) in your launch command or Accelerate config file.", UserWarning)
        else:
            raise ConnectionError(f"Tried to launch distributed communication on port `{main_process_port}`, but another process is utilizing it. Please specify a different port (such as using the `--main_process_port` flag or specifying a different `main_process_port` in your config file) and rerun your script. To automatically use the next open port (on a single node), you can set this to `0`.")
    if args.module and args.no_python:
        raise ValueError("--module and --no_python cannot be used together")
    elif args.module:
        args.module = True
    elif args.no_python:
        args.no_python = True

    current_env = os.environ.copy()
    if args.debug:
        current_env["ACCELERATE_DEBUG_MODE"] = "true"
    gpu_ids = getattr(args, "gpu_ids", "all")
    if gpu_ids != "all" and args.gpu_ids is not None:
        current_env["CUDA_VISIBLE_DEVICES"] = gpu_ids
    current_env["ACCELERATE_MIXED_PRECISION"] = _mixed_precision(args)
    if str(args.mixed_precision).lower() == "fp8":
        setup_fp8_env(args, current_env)
    _dynamo_env(args, current_env)

    if args.use_fsdp:
        current_env["ACCELERATE_USE_FSDP"] = "true"
        if args.fsdp_cpu_ram_efficient_loading and not args.fsdp_sync_module_states:
            raise ValueError("When using `--fsdp_cpu_ram_efficient_loading` set `--fsdp_sync_module_states` to `True`")
        current_env["FSDP_VERSION"] = str(args.fsdp_version) if hasattr(args, "fsdp_version") else "1"
        # FSDP1 uses sharding_strategy, FSDP2 reshard_after_forward; both travel when given
        if getattr(args, "fsdp_sharding_strategy", None) is not None:
            current_env["FSDP_SHARDING_STRATEGY"] = str(args.fsdp_sharding_strategy)
        if getattr(args, "fsdp_reshard_after_forward", None) is not None:
            current_env["FSDP_RESHARD_AFTER_FORWARD"] = str(args.fsdp_reshard_after_forward).lower()
        current_env["FSDP_OFFLOAD_PARAMS"] = str(args.fsdp_offload_params).lower()
        current_env["FSDP_MIN_NUM_PARAMS"] = str(args.fsdp_min_num_params)
        if args.fsdp_auto_wrap_policy is not None:
            current_env["FSDP_AUTO_WRAP_POLICY"] = str(args.fsdp_auto_wrap_policy)
        if args.fsdp_transformer_layer_cls_to_wrap is not None:
            current_env["FSDP_TRANSFORMER_CLS_TO_WRAP"] = str(args.fsdp_transformer_layer_cls_to_wrap)
        if args.fsdp_backward_prefetch is not None:
            current_env["FSDP_BACKWARD_PREFETCH"] = str(args.fsdp_backward_prefetch)
        if args.fsdp_state_dict_type is not None:
            current_env["FSDP_STATE_DICT_TYPE"] = str(args.fsdp_state_dict_type)
        current_env["FSDP_FORWARD_PREFETCH"] = str(args.fsdp_forward_prefetch).lower()
        current_env["FSDP_USE_ORIG_PARAMS"] = str(args.fsdp_use_orig_params).lower()
        current_env["FSDP_CPU_RAM_EFFICIENT_LOADING"] = str(args.fsdp_cpu_ram_efficient_loading).lower()
        current_env["FSDP_SYNC_MODULE_STATES"] = str(args.fsdp_sync_module_states).lower()
        current_env["FSDP_ACTIVATION_CHECKPOINTING"] = str(args.fsdp_activation_checkpointing).lower()
        if getattr(args, "fsdp_ignored_modules", None) is not None:
            current_env["FSDP_IGNORED_MODULES"] = str(args.fsdp_ignored_modules)

    if args.use_megatron_lm:
        prefix = "MEGATRON_LM_"
        current_env["ACCELERATE_USE_MEGATRON_LM"] = "true"
        current_env[prefix + "TP_DEGREE"] = str(args.megatron_lm_tp_degree)
        current_env[prefix + "PP_DEGREE"] = str(args.megatron_lm_pp_degree)
        current_env[prefix + "GRADIENT_CLIPPING"] = str(args.megatron_lm_gradient_clipping)
        if args.megatron_lm_num_micro_batches is not None:
            current_env[prefix + "NUM_MICRO_BATCHES"] = str(args.megatron_lm_num_micro_batches)
        if args.megatron_lm_sequence_parallelism is not None:
            current_env[prefix + "SEQUENCE_PARALLELISM"] = str(args.megatron_lm_sequence_parallelism)
        if args.megatron_lm_recompute_activations is not None:
            current_env[prefix + "RECOMPUTE_ACTIVATIONS"] = str(args.megatron_lm_recompute_activations)
        if args.megatron_lm_use_distributed_optimizer is not None:
            current_env[prefix + "USE_DISTRIBUTED_OPTIMIZER"] = str(args.megatron_lm_use_distributed_optimizer)

    current_env["OMP_NUM_THREADS"] = str(args.num_cpu_threads_per_process)
    if args.enable_cpu_affinity:
        current_env["ACCELERATE_CPU_AFFINITY"] = "1"
    if not args.use_parallelism_config:
        return current_env
    prepare_extend_env_parallelism_config(args, current_env)
    return current_env


def prepare_extend_env_parallelism_config(args: argparse.Namespace, current_env: dict) -> None:
    """``--parallelism_config_*`` flags -> ``PARALLELISM_CONFIG_*`` variables (in place)."""
    prefix = "PARALLELISM_CONFIG_"
    current_env["ACCELERATE_USE_PARALLELISM_CONFIG"] = "true"
    current_env[prefix + "DP_REPLICATE_SIZE"] = str(args.parallelism_config_dp_replicate_size)
    current_env[prefix + "DP_SHARD_SIZE"] = str(args.parallelism_config_dp_shard_size)
    current_env[prefix + "TP_SIZE"] = str(args.parallelism_config_tp_size)
    current_env[prefix + "CP_SIZE"] = str(args.parallelism_config_cp_size)
    current_env[prefix + "CP_BACKEND"] = str(args.parallelism_config_cp_backend)
    current_env[prefix + "SP_SIZE"] = str(args.parallelism_config_sp_size)
    current_env[prefix + "SP_BACKEND"] = str(args.parallelism_config_sp_backend)
    if args.parallelism_config_cp_size > 1:
        current_env[prefix + "CP_COMM_STRATEGY"] = str(args.parallelism_config_cp_comm_strategy)
    if args.parallelism_config_sp_size > 1:
        current_env[prefix + "SP_SEQ_LENGTH"] = str(args.parallelism_config_sp_seq_length)
        current_env[prefix + "SP_SEQ_LENGTH_IS_VARIABLE"] = str(args.parallelism_config_sp_seq_length_is_variable).lower()
        current_env[prefix + "SP_ATTN_IMPLEMENTATION"] = str(args.parallelism_config_sp_attn_implementation)


def prepare_deepspeed_cmd_env(args: argparse.Namespace) -> tuple[list[str], dict[str, str]]:
    """DeepSpeed-configured launch.  DeepSpeed is a config shell here (ZeRO maps onto the native sharded
    engine), so the job is started with torchrun and the ``ACCELERATE_DEEPSPEED_*`` variables are set
    exactly as the reference sets them for the plugin to read."""
    if getattr(args, "deepspeed_multinode_launcher", None) not in (None, *DEEPSPEED_MULTINODE_LAUNCHERS):
        raise ValueError(f"Unknown multinode launcher {args.deepspeed_multinode_launcher}; choose from {DEEPSPEED_MULTINODE_LAUNCHERS}")
    current_env = prepare_multi_gpu_env(args)
    current_env["ACCELERATE_USE_DEEPSPEED"] = "true"
    prefix = "ACCELERATE_DEEPSPEED_"
    if args.zero_stage is not None:
        current_env[prefix + "ZERO_STAGE"] = str(args.zero_stage)
    if args.gradient_accumulation_steps is not None:
        current_env["ACCELERATE_GRADIENT_ACCUMULATION_STEPS"] = str(args.gradient_accumulation_steps)
    if args.gradient_clipping is not None:
        current_env["ACCELERATE_GRADIENT_CLIPPING"] = str(args.gradient_clipping).lower()
    if args.offload_optimizer_device is not None:
        current_env[prefix + "OFFLOAD_OPTIMIZER_DEVICE"] = str(args.offload_optimizer_device).lower()
    if args.offload_param_device is not None:
        current_env[prefix + "OFFLOAD_PARAM_DEVICE"] = str(args.offload_param_device).lower()
    if args.zero3_init_flag is not None:
        current_env[prefix + "ZERO3_INIT"] = str(args.zero3_init_flag).lower()
    if args.zero3_save_16bit_model is not None:
        current_env[prefix + "ZERO3_SAVE_16BIT_MODEL"] = str(args.zero3_save_16bit_model).lower()
    if args.deepspeed_config_file is not None:
        current_env[prefix + "CONFIG_FILE"] = str(args.deepspeed_config_file)
    if getattr(args, "deepspeed_moe_layer_cls_names", None) is not None:
        current_env[prefix + "MOE_LAYER_CLS_NAMES"] = str(args.deepspeed_moe_layer_cls_names)
    return [], current_env


def prepare_tpu(args, current_env, pod: bool = False):
    raise NotImplementedError("TPU launch is out of scope for the B200 build (SURVEY.md §7.1)")


def prepare_sagemager_args_inputs(sagemaker_config, args):
    raise NotImplementedError("SageMaker launch is out of scope for the B200 build (SURVEY.md §7.1)")


def env_var_path_add(env_var_name: str, path_to_add: str) -> str:
    """``path_to_add`` appended to the ``:``-separated variable ``env_var_name``."""
    paths = [p for p in os.environ.get(env_var_name, "").split(":") if len(p) > 0]
    paths.append(str(path_to_add))
    return ":".join(paths)


class PrepareForLaunch:
    """Callable handed to ``torch.multiprocessing.start_processes``: sets the rank variables, then runs ``launcher``."""

    def __init__(self, launcher, distributed_type="NO", debug: bool = False):
        self.launcher = launcher
        self.distributed_type = DistributedType(distributed_type)
        self.debug = debug

    def __call__(self, index, *args):
        if self.debug:
            world_size = int(os.environ.get("WORLD_SIZE"))
            rdv_file = os.environ.get("ACCELERATE_DEBUG_RDV_FILE")
            os.environ["RANK"] = str(index)
            os.environ["LOCAL_RANK"] = str(index)
            torch.distributed.init_process_group("gloo", rank=index, store=torch.distributed.FileStore(rdv_file, world_size), world_size=world_size)
        elif self.distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
            os.environ["LOCAL_RANK"] = str(index)
            nproc = int(os.environ.get("NPROC", 1))
            node_rank = int(os.environ.get("NODE_RANK", 0))
            os.environ["RANK"] = str(nproc * node_rank + index)
        os.environ["FORK_LAUNCHED"] = str(1)
        self.launcher(*args)
