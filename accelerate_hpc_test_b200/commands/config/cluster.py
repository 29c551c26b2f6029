"""The interactive questionnaire of ``accelerate config`` for a local machine / cluster
(reference commands/config/cluster.py:58-924, reduced to the CUDA + CPU paths of this build)."""

import os

from ...utils.constants import FSDP_AUTO_WRAP_POLICY, FSDP_BACKWARD_PREFETCH, FSDP_SHARDING_STRATEGY, FSDP_STATE_DICT_TYPE, TORCH_DYNAMO_MODES
from ...utils.dataclasses import ComputeEnvironment, DistributedType
from .config_args import ClusterConfig
from .config_utils import (
    DYNAMO_BACKENDS,
    _ask_field,
    _ask_options,
    _convert_distributed_mode,
    _convert_dynamo_backend,
    _convert_fp8_backend,
    _convert_mixed_precision,
    _convert_yes_no_to_bool,
)


def get_cluster_input():
    distributed_type = _ask_options("Which type of machine are you using?", ["No distributed training", "multi-CPU", "multi-GPU"], _convert_distributed_mode)
    machine_rank = 0
    num_machines = 1
    num_processes = 1
    gpu_ids = None
    main_process_ip = None
    main_process_port = None
    rdzv_backend = "static"
    same_network = True
    debug = False
    if distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
        num_machines = _ask_field("How many different machines will you use (use more than 1 for multi-node training)? [1]: ", int, default=1)
        if num_machines > 1:
            machine_rank = _ask_options("What is the rank of this machine?", list(range(num_machines)), int)
            main_process_ip = _ask_field("What is the IP address of the machine that will host the main process? ")
            main_process_port = _ask_field("What is the port you will use to communicate with the main process? ", int)
            same_network = _ask_field("Are all the machines on the same local network? Answer `no` if nodes are on the cloud and/or on different network hosts [YES/no]: ",
                                      _convert_yes_no_to_bool, default=True, error_message="Please enter yes or no.")
            if not same_network:
                rdzv_backend = _ask_field("What rendezvous backend will you use? ('static', 'c10d', ...): ", default="static")
        debug = _ask_field("Should distributed operations be checked while running for errors? This can avoid timeout issues but will be slower. [yes/NO]: ",
                           _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
    use_cpu = distributed_type == DistributedType.MULTI_CPU
    if distributed_type == DistributedType.NO:
        use_cpu = _ask_field("Do you want to run your training on CPU only (even if a GPU is available)? [yes/NO]:",
                             _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
    dynamo_config = {}
    use_dynamo = _ask_field("Do you wish to optimize your script with torch dynamo?[yes/NO]:", _convert_yes_no_to_bool, default=False,
                            error_message="Please enter yes or no.")
    if use_dynamo:
        prefix = "dynamo_"
        dynamo_config[prefix + "backend"] = _ask_options("Which dynamo backend would you like to use?", [x.lower() for x in DYNAMO_BACKENDS],
                                                         _convert_dynamo_backend, default=2)
        use_custom_options = _ask_field("Do you want to customize the defaults sent to torch.compile? [yes/NO]: ", _convert_yes_no_to_bool,
                                        default=False, error_message="Please enter yes or no.")
        if use_custom_options:
            dynamo_config[prefix + "mode"] = _ask_options("Which mode do you want to use?", TORCH_DYNAMO_MODES, lambda x: TORCH_DYNAMO_MODES[int(x)], default=0)
            dynamo_config[prefix + "use_fullgraph"] = _ask_field("Do you want the fullgraph mode or it is ok to break model into several subgraphs? [yes/NO]: ",
                                                                 _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
            dynamo_config[prefix + "use_dynamic"] = _ask_field("Do you want to enable dynamic shape tracing? [yes/NO]: ", _convert_yes_no_to_bool,
                                                               default=False, error_message="Please enter yes or no.")
            dynamo_config[prefix + "use_regional_compilation"] = _ask_field("Do you want to enable regional compilation? [yes/NO]: ",
                                                                            _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")

    deepspeed_config = {}
    fsdp_config = {}
    parallelism_config = {}
    megatron_lm_config = {}
    if distributed_type == DistributedType.MULTI_GPU:
        use_deepspeed = _ask_field("Do you want to use a DeepSpeed (ZeRO) style configuration? It runs on the native sharded engine. [yes/NO]: ",
                                   _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
        if use_deepspeed:
            distributed_type = DistributedType.DEEPSPEED
            deepspeed_config["zero_stage"] = _ask_options("What should be your DeepSpeed's ZeRO optimization stage?", [0, 1, 2, 3], int, default=2)
            deepspeed_config["gradient_accumulation_steps"] = _ask_field("How many gradient accumulation steps you're passing in your script? [1]: ", int, default=1)
            use_gradient_clipping = _ask_field("Do you want to use gradient clipping? [yes/NO]: ", _convert_yes_no_to_bool, default=False,
                                               error_message="Please enter yes or no.")
            if use_gradient_clipping:
                deepspeed_config["gradient_clipping"] = _ask_field("What is the gradient clipping value? [1.0]: ", float, default=1.0)
        if distributed_type == DistributedType.MULTI_GPU:
            use_fsdp = _ask_field("Do you want to use FullyShardedDataParallel? [yes/NO]: ", _convert_yes_no_to_bool, default=False,
                                  error_message="Please enter yes or no.")
            if use_fsdp:
                distributed_type = DistributedType.FSDP
                fsdp_config["fsdp_version"] = _ask_options("What should be your FSDP version? [2]: ", [1, 2], lambda x: int(x) + 1, default=1)
                fsdp_version = fsdp_config["fsdp_version"]
                if fsdp_version == 1:
                    fsdp_config["fsdp_reshard_after_forward"] = _ask_options("What should be your sharding strategy?", FSDP_SHARDING_STRATEGY,
                                                                              lambda x: FSDP_SHARDING_STRATEGY[int(x)])
                else:
                    fsdp_config["fsdp_reshard_after_forward"] = _ask_field("Do you want to enable resharding after forward? [YES/no]: ",
                                                                            _convert_yes_no_to_bool, default=True, error_message="Please enter yes or no.")
                fsdp_config["fsdp_offload_params"] = _ask_field("Do you want to offload parameters and gradients to CPU? [yes/NO]: ",
                                                                _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
                fsdp_config["fsdp_auto_wrap_policy"] = _ask_options("What should be your auto wrap policy?", FSDP_AUTO_WRAP_POLICY,
                                                                    lambda x: FSDP_AUTO_WRAP_POLICY[int(x)])
                if fsdp_config["fsdp_auto_wrap_policy"] == FSDP_AUTO_WRAP_POLICY[0]:
                    use_no_split_modules = _ask_field("Do you want to use the model's `_no_split_modules` to wrap. Only applicable for 🤗 Transformers [yes/NO]: ",
                                                      _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
                    if not use_no_split_modules:
                        fsdp_config["fsdp_transformer_layer_cls_to_wrap"] = _ask_field(
                            "Specify the comma-separated list of transformer layer class names (case-sensitive) to wrap ,e.g, :`LlamaDecoderLayer`, `BertLayer`...? : ", str)
                elif fsdp_config["fsdp_auto_wrap_policy"] == FSDP_AUTO_WRAP_POLICY[1]:
                    fsdp_config["fsdp_min_num_params"] = _ask_field("What should be your FSDP's minimum number of parameters for Default Auto Wrapping Policy? [1e8]: ",
                                                                    int, default=100000000)
                if fsdp_version == 1:
                    fsdp_config["fsdp_backward_prefetch"] = _ask_options("What should be your FSDP's backward prefetch policy?", FSDP_BACKWARD_PREFETCH,
                                                                         lambda x: FSDP_BACKWARD_PREFETCH[int(x)])
                fsdp_config["fsdp_state_dict_type"] = _ask_options("What should be your FSDP's state dict type?", FSDP_STATE_DICT_TYPE,
                                                                   lambda x: FSDP_STATE_DICT_TYPE[int(x)], default=2 if fsdp_version == 1 else 2)
                fsdp_config["fsdp_cpu_ram_efficient_loading"] = _ask_field("Do you want to enable CPU RAM efficient model loading? Only applicable for 🤗 Transformers models. [YES/no]: ",
                                                                           _convert_yes_no_to_bool, default=True, error_message="Please enter yes or no.")
                fsdp_config["fsdp_activation_checkpointing"] = _ask_field("Do you want to enable FSDP activation checkpointing? [yes/NO]: ",
                                                                          _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
                if fsdp_version == 2:
                    use_parallelism_config = _ask_field("Do you want to use the parallelism config? [yes/NO]: ", _convert_yes_no_to_bool, default=False,
                                                        error_message="Please enter yes or no.")
                    if use_parallelism_config:
                        prefix = "parallelism_config_"
                        parallelism_config[prefix + "dp_replicate_size"] = _ask_field("What is the data parallelism replicate size? [1]: ", int, default=1)
                        parallelism_config[prefix + "dp_shard_size"] = _ask_field("What is the FSDP shard size? [1]: ", int, default=1)
                        parallelism_config[prefix + "tp_size"] = _ask_field("What is the tensor parallelism size? [1]: ", int, default=1)
                        parallelism_config[prefix + "cp_size"] = _ask_field("What is the context parallelism size? [1]: ", int, default=1)
                        if parallelism_config[prefix + "cp_size"] > 1:
                            parallelism_config[prefix + "cp_comm_strategy"] = _ask_options("What is the compute parallelism communication strategy?",
                                                                                          ["allgather", "alltoall"], lambda x: ["allgather", "alltoall"][int(x)], default=0)
                        parallelism_config[prefix + "sp_size"] = _ask_field("What is the sequence parallelism (Ulysses) size? [1]: ", int, default=1)

    if distributed_type in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU, DistributedType.DEEPSPEED, DistributedType.FSDP):
        machine_type = "GPU" if distributed_type != DistributedType.MULTI_CPU else "CPU"
        num_processes = _ask_field(f"How many {machine_type}(s) should be used for distributed training? [1]:", int, default=1,
                                   error_message="Please enter an integer.")
    if distributed_type in (DistributedType.MULTI_GPU, DistributedType.NO, DistributedType.FSDP, DistributedType.DEEPSPEED) and not use_cpu:
        gpu_ids = _ask_field("What GPU(s) (by id) should be used for training on this machine as a comma-separated list? [all]:", default="all")
    enable_cpu_affinity = False
    if distributed_type in (DistributedType.NO, DistributedType.MULTI_GPU) and not use_cpu:
        enable_cpu_affinity = _ask_field("Would you like to enable numa efficiency? (Currently only supported on NVIDIA hardware). [yes/NO]: ",
                                         _convert_yes_no_to_bool, default=False, error_message="Please enter yes or no.")
    fp8_config = None
    if use_cpu:
        mixed_precision = "no"
    else:
        mixed_precision = _ask_options("Do you wish to use mixed precision?", ["no", "fp16", "bf16", "fp8"], _convert_mixed_precision)
        if mixed_precision == "fp8":
            fp8_config = {}
            fp8_config["backend"] = _ask_options("Which FP8 backend do you want to use?", ["native (MXFP8 on B200)", "te", "ao", "msamp"], _convert_fp8_backend)
            if str(fp8_config["backend"]).upper() in ("NATIVE", "FP8BACKENDTYPE.NATIVE", "TE", "FP8BACKENDTYPE.TE"):
                fp8_config["use_mxfp8_block_scaling"] = _ask_field("Do you want to use MXFP8 1x32 block scaling? [YES/no]: ", _convert_yes_no_to_bool,
                                                                   default=True, error_message="Please enter yes or no.")
            fp8_config["backend"] = getattr(fp8_config["backend"], "value", fp8_config["backend"])
    return ClusterConfig(
        compute_environment=ComputeEnvironment.LOCAL_MACHINE, distributed_type=distributed_type, num_processes=num_processes, gpu_ids=gpu_ids,
        mixed_precision=str(mixed_precision), downcast_bf16=False, machine_rank=machine_rank, num_machines=num_machines,
        main_process_ip=main_process_ip, main_process_port=main_process_port, main_training_function="main", fp8_config=fp8_config,
        deepspeed_config=deepspeed_config, fsdp_config=fsdp_config, parallelism_config=parallelism_config, megatron_lm_config=megatron_lm_config,
        use_cpu=use_cpu, rdzv_backend=rdzv_backend, same_network=same_network, tpu_use_cluster=False, tpu_use_sudo=False, dynamo_config=dynamo_config,
        debug=debug, enable_cpu_affinity=enable_cpu_affinity,
    )
