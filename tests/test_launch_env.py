"""Environment builders of the launcher (behaviour ports of the reference's tests/test_launch.py:21-73 plus the flag tables)."""

import argparse
import os
import socket
import sys

import pytest

from accelerate_hpc_test_b200.commands.launch import _validate_launch_command, launch_command_parser
from accelerate_hpc_test_b200.utils import launch as L
from accelerate_hpc_test_b200.utils.launch import (
    PrepareForLaunch,
    _filter_args,
    env_var_path_add,
    prepare_deepspeed_cmd_env,
    prepare_extend_env_parallelism_config,
    prepare_multi_gpu_env,
    prepare_simple_launcher_cmd_env,
    setup_fp8_env,
)


def parse(*argv):
    args = launch_command_parser().parse_args([*argv, "train.py", "--lr", "1"])
    args, _, _ = _validate_launch_command(args)
    return args


@pytest.fixture(autouse=True)
def no_default_config(monkeypatch, tmp_path):
    import accelerate_hpc_test_b200.commands.launch as launch_mod

    monkeypatch.setattr(launch_mod, "default_config_file", str(tmp_path / "absent.yaml"))


def test_simple_launcher_command_and_env():
    args = parse("--cpu", "--mixed_precision", "bf16", "--debug")
    cmd, env = prepare_simple_launcher_cmd_env(args)
    assert cmd == [sys.executable, "train.py", "--lr", "1"]
    assert env["ACCELERATE_USE_CPU"] == "True" and env["ACCELERATE_MIXED_PRECISION"] == "bf16" and env["ACCELERATE_DEBUG_MODE"] == "true"
    assert env["ACCELERATE_DYNAMO_BACKEND"] == "NO" and "OMP_NUM_THREADS" in env


def test_module_and_no_python_flags():
    args = parse("--cpu", "--module")
    cmd, _ = prepare_simple_launcher_cmd_env(args)
    assert cmd[:3] == [sys.executable, "-m", "train.py"]
    args = parse("--cpu", "--no_python")
    cmd, _ = prepare_simple_launcher_cmd_env(args)
    assert cmd[0] == "train.py"
    args = parse("--cpu", "--no_python", "--module")
    with pytest.raises(ValueError, match="cannot be used together"):
        prepare_simple_launcher_cmd_env(args)


def test_gpu_ids_become_cuda_visible_devices():
    args = parse("--cpu", "--gpu_ids", "1,3")
    _, env = prepare_simple_launcher_cmd_env(args)
    assert env["CUDA_VISIBLE_DEVICES"] == "1,3"
    args = parse("--cpu")
    _, env = prepare_simple_launcher_cmd_env(args)
    assert args.gpu_ids == "all" and env.get("CUDA_VISIBLE_DEVICES") == os.environ.get("CUDA_VISIBLE_DEVICES")


def test_unknown_precision_and_dynamo_backend_are_rejected():
    args = parse("--cpu")
    args.mixed_precision = "int3"
    with pytest.raises(ValueError, match="Unknown mixed_precision"):
        prepare_simple_launcher_cmd_env(args)
    args = parse("--cpu")
    args.dynamo_backend = "warp_drive"
    with pytest.raises(ValueError, match="Unknown dynamo backend"):
        prepare_simple_launcher_cmd_env(args)


def test_mpirun_prefix(monkeypatch):
    monkeypatch.setattr(L, "_get_mpirun_args", lambda: ("mpirun", "--hostfile", "-n", "--npernode", "--bind-to"))
    args = parse("--cpu", "--num_processes", "8", "--num_machines", "2", "--mpirun_hostfile", "/tmp/hosts", "--main_process_ip", "10.0.0.1",
                 "--main_process_port", "29400")
    cmd, env = prepare_simple_launcher_cmd_env(args)
    assert cmd[:7] == ["mpirun", "--hostfile", "/tmp/hosts", "--npernode", "4", "-n", "8"]
    assert cmd[7:13] == ["--bind-to", "socket", "-x", "MASTER_ADDR", "-x", "MASTER_PORT"]
    assert env["MASTER_ADDR"] == "10.0.0.1" and env["MASTER_PORT"] == "29400"


def test_multi_gpu_single_node_rendezvous():
    args = parse("--multi_gpu", "--num_processes", "4", "--main_process_port", "29511")
    env = prepare_multi_gpu_env(args)
    assert args.nproc_per_node == "4" and args.master_addr == "127.0.0.1" and args.master_port == "29511"
    assert "ACCELERATE_USE_FSDP" not in env and env["ACCELERATE_MIXED_PRECISION"] == "no"


def test_multi_node_uses_rdzv_endpoint_or_master_addr():
    args = parse("--multi_gpu", "--num_processes", "16", "--num_machines", "2", "--machine_rank", "1", "--main_process_ip", "10.1.1.1",
                 "--main_process_port", "29500")
    prepare_multi_gpu_env(args)
    assert (args.nproc_per_node, args.nnodes, args.node_rank, args.rdzv_endpoint) == ("8", "2", 1, "10.1.1.1:29500")
    args = parse("--multi_gpu", "--num_processes", "16", "--num_machines", "2", "--machine_rank", "1", "--main_process_ip", "10.1.1.1",
                 "--main_process_port", "29500", "--same_network")
    prepare_multi_gpu_env(args)
    assert args.master_addr == "10.1.1.1" and args.master_port == "29500"


def test_busy_port_switches_to_standalone_on_one_node_and_fails_on_many():
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("localhost", 0))
        s.listen(1)
        port = s.getsockname()[1]
        args = parse("--multi_gpu", "--num_processes", "2", "--main_process_port", str(port))
        with pytest.warns(UserWarning, match="already in use"):
            prepare_multi_gpu_env(args)
        assert args.standalone is True
        args = parse("--multi_gpu", "--num_processes", "4", "--num_machines", "2", "--machine_rank", "0", "--main_process_ip", "127.0.0.1",
                     "--main_process_port", str(port))
        with pytest.raises(ConnectionError):
            prepare_multi_gpu_env(args)


def test_fsdp_flags_travel_as_environment():
    args = parse("--use_fsdp", "--num_processes", "8", "--fsdp_version", "2", "--fsdp_auto_wrap_policy", "TRANSFORMER_BASED_WRAP",
                 "--fsdp_transformer_layer_cls_to_wrap", "LlamaDecoderLayer", "--fsdp_state_dict_type", "SHARDED_STATE_DICT",
                 "--fsdp_reshard_after_forward", "true", "--fsdp_activation_checkpointing", "true", "--fsdp_cpu_ram_efficient_loading", "true",
                 "--fsdp_sync_module_states", "true", "--main_process_port", "29512")
    env = prepare_multi_gpu_env(args)
    assert env["ACCELERATE_USE_FSDP"] == "true" and env["FSDP_VERSION"] == "2" and env["FSDP_AUTO_WRAP_POLICY"] == "TRANSFORMER_BASED_WRAP"
    assert env["FSDP_TRANSFORMER_CLS_TO_WRAP"] == "LlamaDecoderLayer" and env["FSDP_STATE_DICT_TYPE"] == "SHARDED_STATE_DICT"
    assert env["FSDP_RESHARD_AFTER_FORWARD"] == "true" and env["FSDP_ACTIVATION_CHECKPOINTING"] == "true"
    assert env["FSDP_OFFLOAD_PARAMS"] == "false" and env["FSDP_CPU_RAM_EFFICIENT_LOADING"] == "true"


def test_fsdp_ram_efficient_loading_needs_sync_module_states():
    args = parse("--use_fsdp", "--num_processes", "2", "--fsdp_cpu_ram_efficient_loading", "true", "--fsdp_sync_module_states", "false",
                 "--main_process_port", "29513")
    with pytest.raises(ValueError, match="fsdp_sync_module_states"):
        prepare_multi_gpu_env(args)


def test_parallelism_config_flags():
    args = parse("--use_fsdp", "--fsdp_version", "2", "--use_parallelism_config", "--num_processes", "8", "--parallelism_config_dp_shard_size", "4",
                 "--parallelism_config_tp_size", "2", "--main_process_port", "29514")
    env = prepare_multi_gpu_env(args)
    assert env["ACCELERATE_USE_PARALLELISM_CONFIG"] == "true" and env["PARALLELISM_CONFIG_DP_SHARD_SIZE"] == "4"
    assert env["PARALLELISM_CONFIG_TP_SIZE"] == "2" and env["PARALLELISM_CONFIG_CP_SIZE"] == "1"
    assert "PARALLELISM_CONFIG_CP_COMM_STRATEGY" not in env and "PARALLELISM_CONFIG_SP_SEQ_LENGTH" not in env
    ns = argparse.Namespace(parallelism_config_dp_replicate_size=1, parallelism_config_dp_shard_size=2, parallelism_config_tp_size=1,
                            parallelism_config_cp_size=2, parallelism_config_cp_backend="torch", parallelism_config_cp_comm_strategy="alltoall",
                            parallelism_config_sp_size=2, parallelism_config_sp_backend="b200", parallelism_config_sp_seq_length="4096",
                            parallelism_config_sp_seq_length_is_variable=False, parallelism_config_sp_attn_implementation="sdpa")
    env = {}
    prepare_extend_env_parallelism_config(ns, env)
    assert env["PARALLELISM_CONFIG_CP_COMM_STRATEGY"] == "alltoall" and env["PARALLELISM_CONFIG_SP_SEQ_LENGTH_IS_VARIABLE"] == "false"
    assert env["PARALLELISM_CONFIG_SP_ATTN_IMPLEMENTATION"] == "sdpa"


def test_use_parallelism_config_requires_fsdp2():
    with pytest.raises(ValueError, match="use_parallelism_config"):
        parse("--multi_gpu", "--num_processes", "2", "--use_parallelism_config")


def test_megatron_flags():
    args = parse("--use_megatron_lm", "--num_processes", "4", "--megatron_lm_tp_degree", "2", "--megatron_lm_pp_degree", "2",
                 "--megatron_lm_num_micro_batches", "8", "--main_process_port", "29515")
    env = prepare_multi_gpu_env(args)
    assert env["ACCELERATE_USE_MEGATRON_LM"] == "true" and env["MEGATRON_LM_TP_DEGREE"] == "2" and env["MEGATRON_LM_NUM_MICRO_BATCHES"] == "8"
    assert "MEGATRON_LM_SEQUENCE_PARALLELISM" not in env


def test_deepspeed_flags_are_a_config_dialect():
    args = parse("--use_deepspeed", "--num_processes", "4", "--zero_stage", "3", "--gradient_accumulation_steps", "2", "--offload_param_device", "CPU",
                 "--zero3_init_flag", "true", "--main_process_port", "29516")
    cmd, env = prepare_deepspeed_cmd_env(args)
    assert cmd == [] and env["ACCELERATE_USE_DEEPSPEED"] == "true" and env["ACCELERATE_DEEPSPEED_ZERO_STAGE"] == "3"
    assert env["ACCELERATE_GRADIENT_ACCUMULATION_STEPS"] == "2" and env["ACCELERATE_DEEPSPEED_OFFLOAD_PARAM_DEVICE"] == "cpu"
    assert env["ACCELERATE_DEEPSPEED_ZERO3_INIT"] == "true" and "ACCELERATE_DEEPSPEED_CONFIG_FILE" not in env
    assert env["ACCELERATE_CONFIG_DS_FIELDS"] == ""                          # nothing came from an accelerate config file
    args.deepspeed_fields_from_accelerate_config = "deepspeed_config_file,Zero3_Init_Flag"
    assert prepare_deepspeed_cmd_env(args)[1]["ACCELERATE_CONFIG_DS_FIELDS"] == "deepspeed_config_file,zero3_init_flag"
    args.deepspeed_multinode_launcher = "carrier_pigeon"
    with pytest.raises(ValueError, match="Unknown multinode launcher"):
        prepare_deepspeed_cmd_env(args)


def test_fp8_environment():
    ns = argparse.Namespace(fp8_backend="TE", fp8_format="HYBRID", fp8_margin=2, fp8_override_linear_precision=(True, False, True),
                            fp8_opt_level=None, unrelated=1)
    env = setup_fp8_env(ns, {})
    assert env == {"ACCELERATE_FP8_BACKEND": "TE", "ACCELERATE_FP8_FORMAT": "HYBRID", "ACCELERATE_FP8_MARGIN": "2",
                   "ACCELERATE_FP8_OVERRIDE_FPROP": "True", "ACCELERATE_FP8_OVERRIDE_DGRAD": "False", "ACCELERATE_FP8_OVERRIDE_WGRAD": "True"}


def test_filter_args_keeps_only_what_the_target_parser_knows():
    target = argparse.ArgumentParser()
    target.add_argument("--nnodes", default="1")
    target.add_argument("--tee", default="0")
    ours = argparse.Namespace(nnodes="4", mixed_precision="bf16")
    out = _filter_args(ours, target, ["--tee", "3"])
    assert vars(out) == {"nnodes": "4", "tee": "3"}


def test_env_var_path_add(monkeypatch):
    monkeypatch.setenv("MY_PATHS", "/a::/b")
    assert env_var_path_add("MY_PATHS", "/c") == "/a:/b:/c"
    monkeypatch.delenv("MY_PATHS")
    assert env_var_path_add("MY_PATHS", "/c") == "/c"


def test_prepare_for_launch_sets_rank_variables(monkeypatch):
    seen = {}
    monkeypatch.setenv("NPROC", "4")
    monkeypatch.setenv("NODE_RANK", "1")
    for k in ("RANK", "LOCAL_RANK", "FORK_LAUNCHED"):
        monkeypatch.delenv(k, raising=False)
    launcher = PrepareForLaunch(lambda *a: seen.update(rank=os.environ["RANK"], local=os.environ["LOCAL_RANK"], args=a), "MULTI_CPU")
    launcher(2, "x")
    assert seen == {"rank": "6", "local": "2", "args": ("x",)} and os.environ["FORK_LAUNCHED"] == "1"
    for k in ("RANK", "LOCAL_RANK", "FORK_LAUNCHED"):
        os.environ.pop(k, None)


def test_validate_reports_defaults_and_contradictions(caplog):
    args = launch_command_parser().parse_args(["--cpu", "--multi_gpu", "x.py"])
    with pytest.raises(ValueError, match="only use one of"):
        _validate_launch_command(args)
    args = launch_command_parser().parse_args(["--multi_gpu", "--num_processes", "1", "x.py"])
    with pytest.raises(ValueError, match="at least 2 processes"):
        _validate_launch_command(args)
    args = launch_command_parser().parse_args(["--cpu", "x.py"])
    args, cfg, from_cfg = _validate_launch_command(args)
    assert cfg is None and from_cfg is False and args.num_processes == 1 and args.num_machines == 1 and args.mixed_precision == "no"
    assert args.machine_rank == 0 and args.num_cpu_threads_per_process >= 1


def test_megatron_tail_flags_reach_the_plugin_through_the_environment():
    from accelerate_hpc_test_b200.commands.launch import launch_command_parser
    from accelerate_hpc_test_b200.utils import MegatronLMPlugin
    from accelerate_hpc_test_b200.utils.environment import patch_environment
    from accelerate_hpc_test_b200.utils.launch import prepare_multi_gpu_env

    args = launch_command_parser().parse_args([
        "--use_megatron_lm", "--num_processes", "8", "--num_machines", "1", "--main_process_port", "29999", "--megatron_lm_tp_degree", "2", "--megatron_lm_context_parallel_size", "2",
        "--megatron_lm_recompute_granularity", "selective", "--megatron_lm_attention_dropout", "0.1", "--megatron_lm_eod_mask_loss", "true",
        "--megatron_lm_use_rotary_position_embeddings", "true", "--megatron_lm_recompute_num_layers", "4", "train.py"])
    args, _, _ = _validate_launch_command(args)
    assert args.megatron_lm_eod_mask_loss is True and args.megatron_lm_no_save_optim is False and args.megatron_lm_hidden_dropout is None
    env = prepare_multi_gpu_env(args)
    assert env["ACCELERATE_USE_MEGATRON_LM"] == "true" and env["MEGATRON_LM_TP_DEGREE"] == "2" and env["MEGATRON_LM_CONTEXT_PARALLEL_SIZE"] == "2"
    assert env["MEGATRON_LM_RECOMPUTE_GRANULARITY"] == "selective" and env["MEGATRON_LM_ATTENTION_DROPOUT"] == "0.1"
    assert env["MEGATRON_LM_EOD_MASK_LOSS"] == "True" and "MEGATRON_LM_HIDDEN_DROPOUT" not in env
    with patch_environment(**{k: v for k, v in env.items() if k.startswith("MEGATRON_LM_")}):
        plugin = MegatronLMPlugin()
    assert plugin.tp_degree == 2 and plugin.context_parallel_size == 2 and plugin.recompute_granularity == "selective"
    assert plugin.attention_dropout == 0.1 and plugin.eod_mask_loss is True and plugin.use_rotary_position_embeddings is True
    assert plugin.recompute_num_layers == 4 and plugin.hidden_dropout is None
