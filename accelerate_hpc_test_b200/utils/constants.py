"""File names, option lists and version floors shared by the checkpoint, launch and configuration layers.

The *values* are a wire format: checkpoint directories and YAML configs must interchange with the reference, so the strings are the
reference's (src/accelerate/utils/constants.py:20-87; checkpoint layout table in SURVEY.md §5.4).  Grouped here by who reads them.
"""

import operator

# ---- comparison operators accepted by utils.versions.compare_versions
STR_OPERATION_TO_FUNC = {"<": operator.lt, "<=": operator.le, "==": operator.eq, "!=": operator.ne, ">=": operator.ge, ">": operator.gt}

# ---- version floors
MIN_TORCH_VERSION = "2.6.0"  # DeviceMesh, DCP, symmetric memory: this framework targets torch >= 2.6
FSDP2_PYTORCH_VERSION = "2.6.0"
FSDP_PYTORCH_VERSION = "2.1.0"
ELASTIC_LOG_LINE_PREFIX_TEMPLATE_PYTORCH_VERSION = "2.2.0"
XPU_PROFILING_AVAILABLE_PYTORCH_VERSION = "2.4.0"
MITA_PROFILING_AVAILABLE_PYTORCH_VERSION = "2.1.0"
DTENSOR_PYTORCH_VERSION = "2.5.0"
BETA_TP_AVAILABLE_PYTORCH_VERSION = "2.3.0"
BETA_TP_AVAILABLE_TRANSFORMERS_VERSION = "4.52.0"
BETA_CP_AVAILABLE_PYTORCH_VERSION = "2.6.0"
BETA_SP_AVAILABLE_DEEPSPEED_VERSION = "0.18.2"  # the reference's Ulysses needs DeepSpeed from this version on; here it needs none

# ---- what save_state / load_state / save_model call their files
MODEL_NAME = "pytorch_model"
SAFE_MODEL_NAME = "model"
WEIGHTS_NAME = f"{MODEL_NAME}.bin"
SAFE_WEIGHTS_NAME = f"{SAFE_MODEL_NAME}.safetensors"
WEIGHTS_INDEX_NAME = f"{WEIGHTS_NAME}.index.json"
SAFE_WEIGHTS_INDEX_NAME = f"{SAFE_WEIGHTS_NAME}.index.json"
WEIGHTS_PATTERN_NAME = MODEL_NAME + "{suffix}.bin"
SAFE_WEIGHTS_PATTERN_NAME = SAFE_MODEL_NAME + "{suffix}.safetensors"
FSDP_MODEL_NAME = f"{MODEL_NAME}_fsdp"
OPTIMIZER_NAME = "optimizer"
SCHEDULER_NAME = "scheduler"
SAMPLER_NAME = "sampler"
DATALOADER_STATE_NAME = "dl_state_dict"
SCALER_NAME = "scaler.pt"
RNG_STATE_NAME = "random_states"
CUSTOM_CHECKPOINT_PATTERN = "custom_checkpoint_{index}.pkl"
PROFILE_PATTERN_NAME = "profile_{suffix}.json"

# ---- choices offered by `accelerate config` / accepted in YAML files (position = the integer older configs store)
FSDP_SHARDING_STRATEGY = ["FULL_SHARD", "SHARD_GRAD_OP", "NO_SHARD", "HYBRID_SHARD", "HYBRID_SHARD_ZERO2"]
FSDP_AUTO_WRAP_POLICY = ["TRANSFORMER_BASED_WRAP", "SIZE_BASED_WRAP", "NO_WRAP"]
FSDP_BACKWARD_PREFETCH = ["BACKWARD_PRE", "BACKWARD_POST", "NO_PREFETCH"]
FSDP_STATE_DICT_TYPE = ["FULL_STATE_DICT", "LOCAL_STATE_DICT", "SHARDED_STATE_DICT"]
FSDP2_STATE_DICT_TYPE = ["SHARDED_STATE_DICT", "FULL_STATE_DICT"]
DEEPSPEED_MULTINODE_LAUNCHERS = ["pdsh", "standard", "openmpi", "mvapich", "mpich", "nossh", "slurm"]
TORCH_DYNAMO_MODES = ["default", "reduce-overhead", "max-autotune"]

# ---- distributed types that run on CUDA devices / that launch through torch.distributed.run
CUDA_DISTRIBUTED_TYPES = ["DEEPSPEED", "MULTI_GPU", "FSDP", "MEGATRON_LM", "TP"]
TORCH_DISTRIBUTED_OPERATION_TYPES = [*CUDA_DISTRIBUTED_TYPES, "MULTI_CPU"]

# ---- `accelerate launch` flags handed to torch.distributed.run unchanged
TORCH_LAUNCH_PARAMS = (
    "nproc_per_node nnodes node_rank master_addr master_port rdzv_backend rdzv_endpoint rdzv_id rdzv_conf standalone max_restarts "
    "monitor_interval start_method role module m no_python run_path log_dir r redirects t tee"
).split()

# ---- SageMaker images the reference's launcher would ask for (kept so that configuration files naming them still load; this build
# launches on local B200 nodes only)
SAGEMAKER_PYTORCH_VERSION = "1.10.2"
SAGEMAKER_PYTHON_VERSION = "py38"
SAGEMAKER_TRANSFORMERS_VERSION = "4.17.0"
SAGEMAKER_PARALLEL_EC2_INSTANCES = ["ml.p3.16xlarge", "ml.p3dn.24xlarge", "ml.p4dn.24xlarge"]


def __getattr__(name):
    """``SUPPORTED_PYTORCH_LAYERS_FOR_UPCASTING``: the layer classes whose weights layerwise casting stores small and computes big.
    Built on first use so that this module stays importable without torch."""
    if name == "SUPPORTED_PYTORCH_LAYERS_FOR_UPCASTING":
        from torch import nn

        return (nn.Conv1d, nn.Conv2d, nn.Conv3d, nn.ConvTranspose1d, nn.ConvTranspose2d, nn.ConvTranspose3d, nn.Linear)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
