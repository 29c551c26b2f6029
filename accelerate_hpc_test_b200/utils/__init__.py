"""Public ``utils`` namespace (mirrors reference src/accelerate/utils/__init__.py)."""

from .constants import (
    MODEL_NAME,
    OPTIMIZER_NAME,
    PROFILE_PATTERN_NAME,
    RNG_STATE_NAME,
    SAFE_MODEL_NAME,
    SAFE_WEIGHTS_INDEX_NAME,
    SAFE_WEIGHTS_NAME,
    SAFE_WEIGHTS_PATTERN_NAME,
    SAMPLER_NAME,
    SCALER_NAME,
    SCHEDULER_NAME,
    TORCH_DISTRIBUTED_OPERATION_TYPES,
    TORCH_LAUNCH_PARAMS,
    WEIGHTS_INDEX_NAME,
    WEIGHTS_NAME,
    WEIGHTS_PATTERN_NAME,
)
from .dataclasses import (
    AORecipeKwargs,
    AutocastKwargs,
    BnbQuantizationConfig,
    ComputeEnvironment,
    CustomDtype,
    DataLoaderConfiguration,
    DDPCommunicationHookType,
    DeepSpeedPlugin,
    DeepSpeedSequenceParallelConfig,
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
    LoggerType,
    MegatronLMPlugin,
    MixedPrecisionPolicy,
    MSAMPRecipeKwargs,
    PrecisionType,
    ProfileKwargs,
    ProjectConfiguration,
    RNGType,
    SageMakerDistributedType,
    TensorInformation,
    TERecipeKwargs,
    TorchContextParallelConfig,
    TorchDynamoPlugin,
    TorchTensorParallelConfig,
    TorchTensorParallelPlugin,
    get_module_class_from_name,
)
from .environment import (
    are_libraries_initialized,
    check_cuda_fp8_capability,
    check_cuda_p2p_ib_support,
    clear_environment,
    convert_dict_to_env_variables,
    get_cpu_distributed_information,
    get_gpu_info,
    get_int_from_env,
    parse_choice_from_env,
    parse_flag_from_env,
    patch_environment,
    purge_accelerate_environment,
    set_numa_affinity,
    str_to_bool,
)
from .imports import *  # noqa: F401,F403 - the is_*_available() probes
from .memory import clear_device_cache, find_executable_batch_size, release_memory, should_reduce_batch_size
from .operations import (
    CannotPadNestedTensorWarning,
    DistributedOperationException,
    GatheredParameters,
    broadcast,
    broadcast_object_list,
    concatenate,
    convert_outputs_to_fp32,
    convert_to_fp32,
    copy_tensor_to_devices,
    find_batch_size,
    find_device,
    gather,
    gather_object,
    get_data_structure,
    honor_type,
    ignorant_find_batch_size,
    initialize_tensors,
    is_namedtuple,
    is_tensor_information,
    is_torch_tensor,
    listify,
    pad_across_processes,
    pad_input_tensors,
    recursively_apply,
    reduce,
    send_to_device,
    slice_tensors,
)
from .other import (
    check_os_kernel,
    clean_state_dict_for_safetensors,
    compile_regions,
    convert_bytes,
    extract_model_from_parallel,
    get_free_port,
    get_module_children_bottom_up,
    get_pretty_name,
    has_compiled_regions,
    is_compiled_module,
    is_port_in_use,
    load,
    merge_dicts,
    recursive_getattr,
    save,
    wait_for_everyone,
    write_basic_config,
)
from .random import set_seed, synchronize_rng_state, synchronize_rng_states
from .tqdm import tqdm
from .versions import compare_versions, is_torch_version


def __getattr__(name):
    """Heavier helpers (device maps, offload stores, FSDP utilities, launch helpers) load on first use."""
    import importlib

    lazy = {
        "modeling": ("check_device_map check_tied_parameters_in_config check_tied_parameters_on_same_device compute_module_sizes "
                     "convert_file_size_to_int dtype_byte_size find_tied_parameters get_balanced_memory get_max_layer_size get_max_memory "
                     "get_mixed_precision_context_manager has_offloaded_params id_tensor_storage infer_auto_device_map is_peft_model "
                     "load_checkpoint_in_model load_offloaded_weights load_state_dict named_module_tensors retie_parameters "
                     "set_module_tensor_to_device shard_checkpoint get_state_dict_offloaded_model get_grad_scaler clean_device_map "
                     "align_module_device get_state_dict_from_offload calculate_maximum_sizes fallback_allocate get_module_size_with_ties").split(),
        "offload": "OffloadedWeightsLoader PrefixedDataset extract_submodules_state_dict load_offloaded_weight offload_state_dict offload_weight save_offload_index".split(),
        "fsdp_utils": ("disable_fsdp_ram_efficient_loading enable_fsdp_ram_efficient_loading ensure_weights_retied fsdp2_apply_ac "
                       "fsdp2_load_full_state_dict fsdp2_prepare_model fsdp2_switch_optimizer_parameters get_fsdp2_grad_scaler "
                       "load_fsdp_model load_fsdp_optimizer merge_fsdp_weights save_fsdp_model save_fsdp_optimizer fsdp2_canonicalize_names").split(),
        "launch": ("PrepareForLaunch _filter_args prepare_deepspeed_cmd_env prepare_multi_gpu_env prepare_sagemager_args_inputs "
                   "prepare_simple_launcher_cmd_env prepare_tpu").split(),
        "deepspeed": "DeepSpeedEngineWrapper DeepSpeedOptimizerWrapper DeepSpeedSchedulerWrapper DummyOptim DummyScheduler HfDeepSpeedConfig get_active_deepspeed_plugin map_pytorch_optim_to_deepspeed".split(),
        "bnb": "has_4bit_bnb_layers load_and_quantize_model".split(),
        "transformer_engine": "apply_fp8_autowrap contextual_fp8_autocast convert_model has_transformer_engine_layers".split(),
        "ao": "convert_model_to_fp8_ao filter_first_and_last_linear_layers has_ao_layers".split(),
        "megatron_lm": ("MegatronLMDummyDataLoader MegatronLMDummyScheduler AbstractTrainStep BertTrainStep GPTTrainStep T5TrainStep "
                        "MegatronEngine MegatronLMOptimizerWrapper MegatronLMSchedulerWrapper avg_losses_across_data_parallel_group "
                        "gather_across_data_parallel_groups").split(),
        "imports": "torchao_required deepspeed_required is_torchvision_available is_triton_available is_bitsandbytes_multi_backend_available".split(),
        "environment": "get_current_device_type".split(),
        "other": "model_has_dtensor compile_regions_deepspeed".split(),
        "dataclasses": "add_model_config_to_megatron_parser".split(),
        "constants": "XPU_PROFILING_AVAILABLE_PYTORCH_VERSION MITA_PROFILING_AVAILABLE_PYTORCH_VERSION".split(),
    }
    for mod, names in lazy.items():
        if name in names:
            m = importlib.import_module(f"{__name__}.{mod}")
            return getattr(m, name)
    # the Megatron helpers carry a prefix here that they do not have in their module (reference utils/__init__.py:269-273)
    if name.startswith("megatron_lm_") and name[len("megatron_lm_"):] in ("initialize", "prepare_data_loader", "prepare_model_optimizer_scheduler",
                                                                         "prepare_optimizer", "prepare_scheduler"):
        return getattr(importlib.import_module(f"{__name__}.megatron_lm"), name[len("megatron_lm_"):])
    if name == "ParallelismConfig":
        return importlib.import_module(__name__.rsplit(".", 1)[0] + ".parallelism_config").ParallelismConfig
    if name == "install_xla":
        return importlib.import_module(f"{__name__}.torch_xla").install_xla
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
