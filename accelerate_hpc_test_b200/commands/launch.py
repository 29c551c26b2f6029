"""``accelerate launch``: merge the YAML config with the command line, export the result as environment
variables and hand off to ``torch.distributed.run`` (one process per GPU) or a plain subprocess.

Flag surface and merge rules follow reference src/accelerate/commands/launch.py: argument groups
:141-983, ``_validate_launch_command`` :1196-1373 (YAML value unless the flag was typed explicitly —
``args.nondefault``), launcher selection :1376-1399, ``simple_launcher`` :986-995,
``multi_gpu_launcher`` :998-1030, ``deepspeed_launcher`` :1033-1083.  TPU and SageMaker launchers are
not part of the B200 build.
"""

from __future__ import annotations

import argparse
import logging
import os
import subprocess
import sys
from pathlib import Path

import torch

from ..utils.constants import TORCH_DYNAMO_MODES
from ..utils.dataclasses import ComputeEnvironment, DistributedType, DynamoBackend
from ..utils.environment import check_cuda_p2p_ib_support, patch_environment, str_to_bool
from ..utils.launch import _filter_args, prepare_deepspeed_cmd_env, prepare_multi_gpu_env, prepare_simple_launcher_cmd_env
from .config.config_args import default_config_file, load_config_from_file
from .utils import CustomArgumentParser, CustomHelpFormatter

logger = logging.getLogger(__name__)

options_to_group = {
    "multi_gpu": "Distributed GPUs",
    "use_deepspeed": "DeepSpeed Arguments",
    "use_fsdp": "FSDP Arguments",
    "use_megatron_lm": "Megatron-LM Arguments",
    "fp8_backend": "FP8 Arguments",
}


def clean_option(option):
    """``--foo-bar=1`` -> ``__foo_bar`` style normalisation used by the help formatter."""
    if "fp8_backend" in option:
        option = "--fp8_backend"
    if option.startswith("--"):
        return option[2:].replace("-", "_")


def launch_command_parser(subparsers=None):
    description = "Launch a python script in a distributed scenario. Arguments can be passed in with either hyphens (`--num-processes=2`) or underscores (`--num_processes=2`)"
    if subparsers is not None:
        parser = subparsers.add_parser("launch", description=description, add_help=False, allow_abbrev=False, formatter_class=CustomHelpFormatter)
    else:
        parser = CustomArgumentParser("Accelerate launch command", description=description, add_help=False, allow_abbrev=False,
                                      formatter_class=CustomHelpFormatter)
    parser.add_argument("-h", "--help", action="help", help="Show this help message and exit.")
    parser.add_argument("--config_file", default=None, help="The config file to use for the default values in the launching script.")
    parser.add_argument("--quiet", "-q", action="store_true", help="Silence subprocess errors from the launch stack trace and only show the relevant tracebacks. (Only applicable to DeepSpeed and single-process configurations)")

    hardware_args = parser.add_argument_group("Hardware Selection Arguments", "Arguments for selecting the hardware to be used.")
    hardware_args.add_argument("--cpu", default=False, action="store_true", help="Whether or not to force the training on the CPU.")
    hardware_args.add_argument("--multi_gpu", default=False, action="store_true", help="Whether or not this should launch a distributed GPU training.")
    hardware_args.add_argument("--tpu", default=False, action="store_true", help="(unsupported on this build) Whether or not this should launch a TPU training.")

    resource_args = parser.add_argument_group("Resource Selection Arguments", "Arguments for fine-tuning how available hardware should be used.")
    resource_args.add_argument("--mixed_precision", type=str, choices=["no", "fp16", "bf16", "fp8"],
                               help="Whether or not to use mixed precision training. Choose between FP16, BF16 and FP8 (MXFP8 block-scaled on B200) training.")
    resource_args.add_argument("--num_processes", type=int, default=None, help="The total number of processes to be launched in parallel.")
    resource_args.add_argument("--num_machines", type=int, default=None, help="The total number of machines used in this training.")
    resource_args.add_argument("--num_cpu_threads_per_process", type=int, default=None,
                               help="The number of CPU threads per process. Can be tuned for optimal performance.")
    resource_args.add_argument("--enable_cpu_affinity", default=False, action="store_true",
                               help="Whether or not CPU affinity and balancing should be enabled. Currently only supported on NVIDIA hardware.")
    resource_args.add_argument("--dynamo_backend", type=str, choices=["no"] + [b.lower() for b in DynamoBackend.list()],
                               help="Choose a backend to optimize your training with dynamo (not used on the native hot path).")
    resource_args.add_argument("--dynamo_mode", type=str, default="default", choices=TORCH_DYNAMO_MODES, help="Choose a mode to optimize your training with dynamo.")
    resource_args.add_argument("--dynamo_use_fullgraph", default=False, action="store_true", help="Whether to use full graph mode for dynamo or it is ok to break model into several subgraphs")
    resource_args.add_argument("--dynamo_use_dynamic", default=False, action="store_true", help="Whether to enable dynamic shape tracing.")
    resource_args.add_argument("--dynamo_use_regional_compilation", default=False, action="store_true", help="Whether to enable regional compilation.")

    paradigm_args = parser.add_argument_group("Training Paradigm Arguments", "Arguments for selecting which training paradigm to be used.")
    paradigm_args.add_argument("--use_deepspeed", default=False, action="store_true", help="Whether to use a DeepSpeed (ZeRO) configuration; it runs on the native sharded engine.")
    paradigm_args.add_argument("--use_fsdp", default=False, action="store_true", help="Whether to use fully sharded data parallelism.")
    paradigm_args.add_argument("--use_parallelism_config", default=False, action="store_true", help="Whether to use the parallelism config to configure the N-d distributed training.")
    paradigm_args.add_argument("--use_megatron_lm", default=False, action="store_true", help="Whether to use a Megatron-LM configuration (tp/cp degrees map onto ParallelismConfig).")
    paradigm_args.add_argument("--use_xpu", default=None, action="store_true", help="(unsupported on this build)")

    distributed_args = parser.add_argument_group("Distributed GPUs", "Arguments related to distributed GPU training.")
    distributed_args.add_argument("--gpu_ids", default=None, help="What GPUs (by id) should be used for training on this machine as a comma-separated list")
    distributed_args.add_argument("--same_network", default=False, action="store_true", help="Whether all machines used for multinode training exist on the same local network.")
    distributed_args.add_argument("--machine_rank", type=int, default=None, help="The rank of the machine on which this script is launched.")
    distributed_args.add_argument("--main_process_ip", type=str, default=None, help="The IP address of the machine of rank 0.")
    distributed_args.add_argument("--main_process_port", type=int, default=None, help="The port to use to communicate with the machine of rank 0.")
    distributed_args.add_argument("-t", "--tee", default="0", type=str, help="Tee std streams into a log file and also to console.")
    distributed_args.add_argument("--log_dir", type=str, default=None, help="Base directory to use for log files when using torchrun/torch.distributed.run as launcher. Use with --tee to redirect std streams info log files.")
    distributed_args.add_argument("--role", type=str, default="default", help="User-defined role for the workers.")
    distributed_args.add_argument("--rdzv_backend", type=str, default="static", help="The rendezvous method to use, such as 'static' (the default) or 'c10d'")
    distributed_args.add_argument("--rdzv_conf", type=str, default="", help="Additional rendezvous configuration (<key1>=<value1>,<key2>=<value2>,...).")
    distributed_args.add_argument("--max_restarts", type=int, default=0, help="Maximum number of worker group restarts before failing.")
    distributed_args.add_argument("--monitor_interval", type=float, default=0.1, help="Interval, in seconds, to monitor the state of workers.")
    parser.add_argument("-m", "--module", action="store_true", help="Change each process to interpret the launch script as a Python module, executing with the same behavior as 'python -m'.")
    parser.add_argument("--no_python", action="store_true", help="Skip prepending the training script with 'python' - just execute it directly. Useful when the script is not a Python script.")

    tpu_args = parser.add_argument_group("TPU", "Arguments related to TPU (unsupported on this build; accepted so old command lines parse).")
    tpu_args.add_argument("--tpu_cluster", action="store_true", dest="tpu_use_cluster", help=argparse.SUPPRESS)
    tpu_args.add_argument("--no_tpu_cluster", action="store_false", dest="tpu_use_cluster", help=argparse.SUPPRESS)
    tpu_args.add_argument("--tpu_use_sudo", action="store_true", help=argparse.SUPPRESS)
    tpu_args.add_argument("--vm", type=str, action="append", help=argparse.SUPPRESS)
    tpu_args.add_argument("--env", type=str, action="append", help=argparse.SUPPRESS)
    tpu_args.add_argument("--main_training_function", type=str, default=None, help=argparse.SUPPRESS)
    tpu_args.add_argument("--downcast_bf16", action="store_true", help=argparse.SUPPRESS)

    deepspeed_args = parser.add_argument_group("DeepSpeed Arguments", "Arguments related to DeepSpeed.")
    deepspeed_args.add_argument("--deepspeed_config_file", default=None, type=str, help="DeepSpeed config file.")
    deepspeed_args.add_argument("--zero_stage", default=None, type=int, help="DeepSpeed's ZeRO optimization stage (useful only when `use_deepspeed` flag is passed). If unspecified, will default to `2`.")
    deepspeed_args.add_argument("--offload_optimizer_device", default=None, type=str, help="Decides where (none|cpu|nvme) to offload optimizer states.")
    deepspeed_args.add_argument("--offload_param_device", default=None, type=str, help="Decides where (none|cpu|nvme) to offload parameters.")
    deepspeed_args.add_argument("--offload_optimizer_nvme_path", default=None, type=str, help="Decides Nvme Path to offload optimizer states.")
    deepspeed_args.add_argument("--offload_param_nvme_path", default=None, type=str, help="Decides Nvme Path to offload parameters.")
    deepspeed_args.add_argument("--gradient_accumulation_steps", default=None, type=int, help="No of gradient_accumulation_steps used in your training script.")
    deepspeed_args.add_argument("--gradient_clipping", default=None, type=float, help="gradient clipping value used in your training script.")
    deepspeed_args.add_argument("--zero3_init_flag", default=None, type=str, help="Decides Whether (true|false) to enable `deepspeed.zero.Init` for constructing massive models.")
    deepspeed_args.add_argument("--zero3_save_16bit_model", default=None, type=str, help="Decides Whether (true|false) to save 16-bit model weights when using ZeRO Stage-3.")
    deepspeed_args.add_argument("--deepspeed_hostfile", default=None, type=str, help="DeepSpeed hostfile for configuring multi-node compute resources.")
    deepspeed_args.add_argument("--deepspeed_exclusion_filter", default=None, type=str, help="DeepSpeed exclusion filter string when using mutli-node setup.")
    deepspeed_args.add_argument("--deepspeed_inclusion_filter", default=None, type=str, help="DeepSpeed inclusion filter string when using mutli-node setup.")
    deepspeed_args.add_argument("--deepspeed_multinode_launcher", default=None, type=str, help="DeepSpeed multi-node launcher to use, e.g. `pdsh`, `standard`, `openmpi`, `mvapich`, `mpich`, `slurm`, `nossh` (requires DeepSpeed >= 0.14.5). If unspecified, will default to `pdsh`.")
    deepspeed_args.add_argument("--deepspeed_moe_layer_cls_names", default=None, type=str, help="comma-separated list of transformer MoE layer class names (case-sensitive) to wrap")

    fsdp_args = parser.add_argument_group("FSDP Arguments", "Arguments related to Fully Shared Data Parallelism.")
    fsdp_args.add_argument("--fsdp_version", type=str, default="1", choices=["1", "2"], help="FSDP version to use. (useful only when `use_fsdp` flag is passed).")
    fsdp_args.add_argument("--fsdp_offload_params", default="false", type=str, help="Decides Whether (true|false) to offload parameters and gradients to CPU.")
    fsdp_args.add_argument("--fsdp_min_num_params", type=int, default=1e8, help="FSDP's minimum number of parameters for Default Auto Wrapping.")
    fsdp_args.add_argument("--fsdp_sharding_strategy", type=str, default="FULL_SHARD", help="FSDP's sharding strategy (FSDP1 only).")
    fsdp_args.add_argument("--fsdp_reshard_after_forward", type=str, default="true", help="FSDP's Reshard After Forward Strategy. Supports either boolean (FSDP2) or `FULL_SHARD | SHARD_GRAD_OP | NO_RESHARD` (FSDP1).")
    fsdp_args.add_argument("--fsdp_auto_wrap_policy", type=str, default=None, help="FSDP's auto wrap policy.")
    fsdp_args.add_argument("--fsdp_transformer_layer_cls_to_wrap", default=None, type=str, help="Transformer layer class name (case-sensitive) to wrap ,e.g, `LlamaDecoderLayer`, `BertLayer` ....")
    fsdp_args.add_argument("--fsdp_backward_prefetch", default=None, type=str, help="FSDP's backward prefetch policy (FSDP1 only).")
    fsdp_args.add_argument("--fsdp_state_dict_type", default=None, type=str, help="FSDP's state dict type.")
    fsdp_args.add_argument("--fsdp_forward_prefetch", default="false", type=str, help="If True, then FSDP explicitly prefetches the next upcoming all-gather while executing in the forward pass (FSDP1 only).")
    fsdp_args.add_argument("--fsdp_use_orig_params", default="true", type=str, help="If True, allows non-uniform `requires_grad` during init (FSDP1 only).")
    fsdp_args.add_argument("--fsdp_cpu_ram_efficient_loading", default="true", type=str, help="If True, only the first process loads the pretrained model checkpoint while all other processes have empty weights.")
    fsdp_args.add_argument("--fsdp_sync_module_states", default="true", type=str, help="If True, each individually wrapped FSDP unit will broadcast module parameters from rank 0 (FSDP1 only).")
    fsdp_args.add_argument("--fsdp_activation_checkpointing", default="false", type=str, help="Decides Whether (true|false) intermediate activations are freed during the forward pass, and a checkpoint is left as a placeholder.")
    fsdp_args.add_argument("--fsdp_ignored_modules", default=None, type=str, help="Regex of module names to exclude from sharding.")

    megatron_lm_args = parser.add_argument_group("Megatron-LM Arguments", "Arguments related to Megatron-LM.")
    megatron_lm_args.add_argument("--megatron_lm_tp_degree", type=int, default=1, help="Megatron-LM's Tensor Parallelism (TP) degree.")
    megatron_lm_args.add_argument("--megatron_lm_pp_degree", type=int, default=1, help="Megatron-LM's Pipeline Parallelism (PP) degree.")
    megatron_lm_args.add_argument("--megatron_lm_num_micro_batches", type=int, default=None, help="Megatron-LM's number of micro batches when PP degree > 1.")
    megatron_lm_args.add_argument("--megatron_lm_sequence_parallelism", default=None, type=str, help="Decides Whether (true|false) to enable Sequence Parallelism when TP degree > 1.")
    megatron_lm_args.add_argument("--megatron_lm_recompute_activations", default=None, type=str, help="Decides Whether (true|false) to enable Selective Activation Recomputation.")
    megatron_lm_args.add_argument("--megatron_lm_use_distributed_optimizer", default=None, type=str, help="Decides Whether (true|false) to use distributed optimizer which shards optimizer state and gradients across Data Pralellel (DP) ranks.")
    megatron_lm_args.add_argument("--megatron_lm_gradient_clipping", default=1.0, type=float, help="Megatron-LM's gradient clipping value based on global L2 Norm (0 to disable).")
    # the long tail of Megatron-LM switches: carried into the plugin (and its `megatron_lm_default_args`) for configuration
    # compatibility; the degrees among them (context / expert parallel sizes) feed `ParallelismConfig`
    tail = "(useful only when `use_megatron_lm` flag is passed)."
    for flag, kind, default, what in _MEGATRON_TAIL:
        megatron_lm_args.add_argument(f"--megatron_lm_{flag}", default=default, type=kind, help=f"{what} {tail}")

    fp8_args = parser.add_argument_group("FP8 Arguments", "Arguments related to FP8 training (native MXFP8 block scaling on B200, or TransformerEngine / torchao / MS-AMP)")
    fp8_args.add_argument("--fp8_backend", type=str, choices=["native", "ao", "te", "msamp"], help="Choose a backend to train with FP8")
    fp8_args.add_argument("--fp8_use_autocast_during_eval", default=False, action="store_true", help="Whether to use FP8 autocast during eval mode. Generally better metrics are found when this is `False`.")
    fp8_args.add_argument("--fp8_margin", type=int, default=0, help="The margin to use for the gradient scaling.")
    fp8_args.add_argument("--fp8_interval", type=int, default=1, help="The interval to use for how often the scaling factor is recomputed.")
    fp8_args.add_argument("--fp8_format", type=str, default="HYBRID", choices=["HYBRID", "E4M3", "E5M2"], help="The format to use for the FP8 recipe.")
    fp8_args.add_argument("--fp8_amax_history_len", type=int, default=1024, help="The length of the history to use for the scaling factor computation.")
    fp8_args.add_argument("--fp8_amax_compute_algo", type=str, default="most_recent", choices=["max", "most_recent"], help="The algorithm to use for the scaling factor computation.")
    fp8_args.add_argument("--fp8_override_linear_precision", type=lambda x: tuple(map(str_to_bool, x.split(","))), default=(False, False, False),
                          help="Whether or not to execute `fprop`, `dgrad`, and `wgrad` GEMMS in higher precision. Should be passed in a comma-separated string of booleans.")
    fp8_args.add_argument("--fp8_use_mxfp8_block_scaling", default=False, action="store_true", help="Use MXFP8 1x32 block scaling (UE8M0 scales), the Blackwell-native fp8 recipe.")
    fp8_args.add_argument("--fp8_opt_level", type=str, default="O2", choices=["O1", "O2"], help="What level of 8-bit collective communication should be used with MS-AMP.")
    fp8_args.add_argument("--fp8_pad_inner_dim", default=True, action="store_true", help="torchao: pad the inner dims of the GEMMs to multiples of 16.")
    fp8_args.add_argument("--fp8_enable_fsdp_float8_all_gather", default=True, action="store_true", help="Gather sharded parameters in fp8.")

    aws_args = parser.add_argument_group("AWS Arguments", "Arguments related to AWS (unsupported on this build).")
    aws_args.add_argument("--aws_access_key_id", type=str, default=None, help=argparse.SUPPRESS)
    aws_args.add_argument("--aws_secret_access_key", type=str, default=None, help=argparse.SUPPRESS)
    parser.add_argument("--debug", action="store_true", help="Whether to print out the torch.distributed stack trace when something fails.")
    mpirun_args = parser.add_argument_group("MPI Arguments", "Arguments related to mpirun for Multi-CPU")
    mpirun_args.add_argument("--mpirun_hostfile", type=str, default=None, help="Location for a hostfile for using Accelerate to launch a multi-CPU training job with mpirun.")
    mpirun_args.add_argument("--mpirun_ccl", type=int, default=1, help="The number of oneCCL worker threads (ignored).")

    parallelism_config_args = parser.add_argument_group("ParallelismConfig Arguments", "Arguments related to the ParallelismConfig used for distributed training.")
    parallelism_config_args.add_argument("--parallelism_config_dp_replicate_size", type=int, default=1, help="The number of processes for data parallel training. Defaults to 1 (no data parallelism).")
    parallelism_config_args.add_argument("--parallelism_config_dp_shard_size", type=int, default=1, help="The number of processes for FSDP sharding. Defaults to 1 (No FSDP sharding).")
    parallelism_config_args.add_argument("--parallelism_config_tp_size", type=int, default=1, help="The number of processes for tensor parallel training. Defaults to 1 (no tensor parallelism).")
    parallelism_config_args.add_argument("--parallelism_config_cp_size", type=int, default=1, help="The number of processese for context parallel training. Defaults to 1 (no context parallelism).")
    parallelism_config_args.add_argument("--parallelism_config_cp_backend", type=str, choices=["torch", "b200"], default="torch", help="Context Parallelism backend.")
    parallelism_config_args.add_argument("--parallelism_config_cp_comm_strategy", type=str, default="allgather", help="The communication strategy for context parallel training. Defaults to 'allgather'. Other option is alltoall")
    parallelism_config_args.add_argument("--parallelism_config_sp_size", type=int, default=1, help="The number of processese for sequence parallel training. Defaults to 1 (no sequence parallelism).")
    parallelism_config_args.add_argument("--parallelism_config_sp_backend", type=str, choices=["deepspeed", "b200"], default="deepspeed", help="Sequence Parallelism backend (both run on the native Ulysses engine).")
    parallelism_config_args.add_argument("--parallelism_config_sp_seq_length", type=str, default=None, help="Sequence length for when `--parallelism_config_sp_seq_length_is_variable=false`.")
    parallelism_config_args.add_argument("--parallelism_config_sp_seq_length_is_variable", type=bool, default=True, help="If `True` will work with a sequence length that may change between batches.")
    parallelism_config_args.add_argument("--parallelism_config_sp_attn_implementation", type=str, default="sdpa", help="Attention implementation to use. Can be one of 'flash_attention_2', 'flash_attention_3', 'sdpa' or 'b200_flash'.")

    parser.add_argument("training_script", type=str, help="The full path to the script to be launched in parallel, followed by all the arguments for the training script.")
    parser.add_argument("training_script_args", nargs=argparse.REMAINDER, help="Arguments of the training script.")
    if subparsers is not None:
        parser.set_defaults(func=launch_command)
    return parser


def simple_launcher(args):
    cmd, current_env = prepare_simple_launcher_cmd_env(args)
    process = subprocess.Popen(cmd, env=current_env)
    process.wait()
    if process.returncode != 0:
        if not args.quiet:
            raise subprocess.CalledProcessError(returncode=process.returncode, cmd=cmd)
        else:
            sys.exit(1)


def multi_gpu_launcher(args):
    import torch.distributed.run as distrib_run

    current_env = prepare_multi_gpu_env(args)
    if not check_cuda_p2p_ib_support():
        message = "Using RTX 4000 series which doesn't support faster communication speedups. Ensuring P2P and IB communications are disabled."
        warn = False
        if "NCCL_P2P_DISABLE" not in current_env:
            current_env["NCCL_P2P_DISABLE"] = "1"
            warn = True
        if "NCCL_IB_DISABLE" not in current_env:
            current_env["NCCL_IB_DISABLE"] = "1"
            warn = True
        if warn:
            logger.warning(message)
    debug = getattr(args, "debug", False)
    args = _filter_args(args, distrib_run.get_args_parser(), ["--training_script", args.training_script, "--training_script_args", args.training_script_args])
    with patch_environment(**current_env):
        try:
            distrib_run.run(args)
        except Exception:
            if debug:
                import traceback

                traceback.print_exc()
            raise


def deepspeed_launcher(args):
    """A DeepSpeed-configured job: same torchrun hand-off, plus the ``ACCELERATE_DEEPSPEED_*`` variables the
    plugin reads; ZeRO stages execute on the native sharded engine."""
    import torch.distributed.run as distrib_run

    _, current_env = prepare_deepspeed_cmd_env(args)
    debug = getattr(args, "debug", False)
    args = _filter_args(args, distrib_run.get_args_parser(), ["--training_script", args.training_script, "--training_script_args", args.training_script_args])
    with patch_environment(**current_env):
        try:
            distrib_run.run(args)
        except Exception:
            if debug:
                import traceback

                traceback.print_exc()
            raise


def tpu_launcher(args):
    raise NotImplementedError("TPU launch is out of scope for the B200 build")


def sagemaker_launcher(sagemaker_config, args):
    raise NotImplementedError("SageMaker launch is out of scope for the B200 build")


def _truthy_flag(text) -> bool:
    return str(text).lower() in ("1", "true", "yes", "y", "on")


# (flag, type, default, help) of the less common Megatron-LM options; "true|false" strings are parsed by the plugin
_MEGATRON_TAIL = (
    ("use_custom_fsdp", _truthy_flag, False, "Whether to use Megatron's own FSDP."),
    ("no_load_optim", _truthy_flag, False, "Do not load the optimizer state from a checkpoint."),
    ("no_save_optim", _truthy_flag, False, "Do not save the optimizer state in checkpoints."),
    ("eod_mask_loss", _truthy_flag, False, "Mask the loss on end-of-document tokens."),
    ("optimizer_cpu_offload", _truthy_flag, False, "Keep optimizer state in host memory."),
    ("use_precision_aware_optimizer", _truthy_flag, False, "Use the precision-aware optimizer."),
    ("overlap_cpu_optimizer_d2h_h2d", _truthy_flag, False, "Overlap the host<->device copies of an offloaded optimizer with compute."),
    ("decoder_last_pipeline_num_layers", int, None, "Number of decoder layers on the last pipeline stage."),
    ("recompute_granularity", str, None, "Activation recompute granularity (full, selective)."),
    ("recompute_method", str, None, "Activation recompute method (uniform, block)."),
    ("recompute_num_layers", int, None, "Number of layers whose activations are recomputed."),
    ("attention_backend", str, None, "Decides Whether (true|false) to enable attention backend."),
    ("expert_model_parallel_size", int, None, "Expert model parallel size."),
    ("expert_tensor_parallel_size", int, None, "Expert tensor parallel size."),
    ("context_parallel_size", int, None, "Context parallel size."),
    ("attention_dropout", float, None, "Attention dropout."),
    ("hidden_dropout", float, None, "Hidden dropout."),
    ("attention_softmax_in_fp32", str, None, "Decides Whether (true|false) to run the attention softmax in fp32."),
    ("calculate_per_token_loss", str, None, "Decides Whether (true|false) to average the loss per token."),
    ("use_rotary_position_embeddings", str, None, "Decides Whether (true|false) to use rotary position embeddings."),
)

_SECTION_PREFIX = {"fsdp_config": "fsdp", "fp8_config": "fp8", "parallelism_config": "parallelism_config", "megatron_lm_config": "megatron_lm"}
_STRATEGY_FLAGS = ("multi_gpu", "tpu", "use_deepspeed", "use_fsdp", "use_megatron_lm")
_NOT_COPIED = ("compute_environment", "mixed_precision", "distributed_type")


class _LaunchResolver:
    """Completes the launch namespace from three sources, strongest first: flags typed on the command line
    (``args.nondefault``), the YAML config (``--config_file`` or the default file), built-in defaults that are reported back
    to the user.  One instance per launch; ``resolve()`` returns ``(args, config or None, mixed_precision_came_from_config)``."""

    def __init__(self, args):
        self.args = args
        self.typed = getattr(args, "nondefault", None) or set()
        self.notes: list = []
        self.config = None
        self.precision_from_config = False

    # ---- rules that need no config ------------------------------------------------------------------------
    def _reject_contradictions(self) -> None:
        a = self.args
        if sum(bool(x) for x in (a.multi_gpu, a.cpu, a.tpu, a.use_deepspeed, a.use_fsdp)) > 1:
            raise ValueError("You can only use one of `--cpu`, `--multi_gpu`, `--tpu`, `--use_deepspeed`, `--use_fsdp` at a time.")
        if a.multi_gpu and a.num_processes is not None and a.num_processes < 2:
            raise ValueError("You need to use at least 2 processes to use `--multi_gpu`.")
        fsdp2 = a.use_fsdp and str(a.fsdp_version) != "1"
        if a.use_parallelism_config and not fsdp2:
            raise ValueError("You cannot use `--use_parallelism_config` without `--use_fsdp` and `--fsdp_version=2`. ")

    def _config_applies(self) -> bool:
        a = self.args
        return a.config_file is not None or (os.path.isfile(default_config_file) and not a.cpu)

    # ---- with a config file --------------------------------------------------------------------------------
    def _adopt_strategy(self, cfg) -> None:
        a = self.args
        if any(getattr(a, flag) for flag in _STRATEGY_FLAGS):
            return  # the command line already picked one
        kind = cfg.distributed_type
        a.multi_gpu = kind == DistributedType.MULTI_GPU
        a.use_deepspeed = kind == DistributedType.DEEPSPEED
        a.use_fsdp = kind == DistributedType.FSDP
        a.use_megatron_lm = kind == DistributedType.MEGATRON_LM
        a.tpu = kind == DistributedType.XLA
        a.tpu_use_cluster = cfg.tpu_use_cluster if a.tpu else False
        a.use_parallelism_config = bool(cfg.parallelism_config)

    def _adopt_gpu_ids(self, cfg) -> None:
        a = self.args
        if a.gpu_ids is None:
            a.gpu_ids = cfg.gpu_ids if cfg.gpu_ids is not None else "all"
        one_machine = a.num_machines is not None and a.num_machines <= 1
        if a.multi_gpu and one_machine and a.gpu_ids != "all" and len(a.gpu_ids.split(",")) < 2:
            raise ValueError("Less than two GPU ids were configured and tried to run on on multiple GPUs. Please ensure at least two "
                             "are specified for `--gpu_ids`, or use `--gpu_ids='all'`.")

    def _adopt_fields(self, cfg) -> None:
        """Top-level keys fill flags that are still unset; the nested sections (``fsdp_config.x`` → ``--fsdp_x`` …) fill every
        flag the user did not type."""
        a = self.args
        for name, value in vars(cfg).items():
            if isinstance(value, dict):
                prefix = _SECTION_PREFIX.get(name)
                for key, item in value.items():
                    flag = key if (prefix is None or key.startswith(prefix)) else f"{prefix}_{key}"
                    if flag not in self.typed:
                        setattr(a, flag, item)
            elif name not in _NOT_COPIED and getattr(a, name, None) is None:
                setattr(a, name, value)

    def _from_config(self) -> None:
        a = self.args
        cfg = self.config = load_config_from_file(a.config_file)
        self._adopt_strategy(cfg)
        self._adopt_gpu_ids(cfg)
        if cfg.compute_environment == ComputeEnvironment.LOCAL_MACHINE:
            self._adopt_fields(cfg)
        a.debug = a.debug or cfg.debug
        if not a.mixed_precision:
            a.mixed_precision = cfg.mixed_precision if cfg.mixed_precision is not None else "no"
            self.precision_from_config = cfg.mixed_precision is not None
        if a.dynamo_backend is None:
            a.dynamo_backend = "no"
        if a.num_processes == -1:
            raise ValueError("You need to manually pass in `--num_processes` using this config yaml.")

    # ---- without a config file -----------------------------------------------------------------------------
    def _default(self, flag: str, value, shown: str = None) -> None:
        if getattr(self.args, flag, None) is None:
            setattr(self.args, flag, value)
            if shown is not None:
                self.notes.append(f"\t`--{flag}` was set to a value of `{shown}`")

    def _from_defaults(self) -> None:
        a = self.args
        gpus = torch.cuda.device_count() if (torch.cuda.is_available() and not a.cpu) else 0
        self._default("num_processes", gpus or 1, str(gpus or 1))
        self._default("debug", False)
        picked = a.multi_gpu or a.use_deepspeed or a.use_fsdp or a.use_megatron_lm
        if not picked and gpus and a.num_processes > 1:
            a.multi_gpu = True
            self.notes.append("\t\tMore than one GPU was found, enabling multi-GPU training.\n"
                              "\t\tIf this was unintended please pass in `--num_processes=1`.")
        self._default("num_machines", 1, "1")
        self._default("mixed_precision", "no", "'no'")
        if not hasattr(a, "use_cpu"):
            a.use_cpu = a.cpu
        self._default("dynamo_backend", "no", "'no'")
        self._default("gpu_ids", "all")

    # ---- common tail ---------------------------------------------------------------------------------------------
    def _cpu_threads(self) -> None:
        a = self.args
        on_sagemaker = self.config is not None and self.config.compute_environment == ComputeEnvironment.AMAZON_SAGEMAKER
        if on_sagemaker or a.num_cpu_threads_per_process is not None:
            return
        a.num_cpu_threads_per_process = 1
        if not (a.cpu and a.num_processes >= 1):
            return
        per_node = max(1, a.num_processes // (a.num_machines or 1))
        try:
            import psutil

            share = int(psutil.cpu_count(logical=False) / per_node)
        except Exception:  # noqa: BLE001 - psutil missing or no physical-core count: keep one thread
            share = 1
        if share > 1:
            a.num_cpu_threads_per_process = share
            self.notes.append(f"\t`--num_cpu_threads_per_process` was set to `{share}` to improve out-of-box performance when training on CPUs")

    def resolve(self):
        a = self.args
        self._reject_contradictions()
        if self._config_applies():
            self._from_config()
        else:
            self._from_defaults()
        if a.debug:
            logger.debug("Running script in debug mode, expect distributed operations to be slightly slower.")
        self._cpu_threads()
        if a.num_machines is None:
            a.num_machines = 1
        if a.machine_rank is None:
            a.machine_rank = 0
        if self.notes:
            logger.warning("The following values were not passed to `accelerate launch` and had defaults used instead:\n"
                           + "\n".join(self.notes)
                           + "\nTo avoid this warning pass in values for each of the problematic parameters or run `accelerate config`.")
        return a, self.config, self.precision_from_config


def _validate_launch_command(args):
    """Fill ``args`` from the config file (explicitly typed flags win), infer process counts, reject contradictory flags."""
    return _LaunchResolver(args).resolve()


def launch_command(args):
    args, defaults, mp_from_config_flag = _validate_launch_command(args)
    if args.tpu:
        tpu_launcher(args)
    elif args.use_deepspeed and not args.cpu:
        args.deepspeed_fields_from_accelerate_config = list(defaults.deepspeed_config.keys()) if defaults else []
        if mp_from_config_flag:
            args.deepspeed_fields_from_accelerate_config.append("mixed_precision")
        args.deepspeed_fields_from_accelerate_config = ",".join(args.deepspeed_fields_from_accelerate_config)
        deepspeed_launcher(args)
    elif args.use_fsdp and not args.cpu:
        multi_gpu_launcher(args)
    elif args.use_megatron_lm and not args.cpu:
        multi_gpu_launcher(args)
    elif args.multi_gpu and not args.cpu:
        multi_gpu_launcher(args)
    elif args.cpu and args.num_processes is not None and args.num_processes > 1 and getattr(args, "mpirun_hostfile", None) is None:
        # multi-process CPU (gloo) on one machine: same torchrun path
        multi_gpu_launcher(args)
    elif defaults is not None and defaults.compute_environment == ComputeEnvironment.AMAZON_SAGEMAKER:
        sagemaker_launcher(defaults, args)
    else:
        simple_launcher(args)


def main():
    parser = launch_command_parser()
    args = parser.parse_args()
    launch_command(args)


if __name__ == "__main__":
    main()
