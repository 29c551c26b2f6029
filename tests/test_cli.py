"""CLI: launch argument parsing / validation, env builders, config files, to-fsdp2, estimate-memory, real launches
(modelled on reference tests/test_cli.py and tests/test_launch.py)."""

import os
import subprocess
import sys

import pytest
import yaml

from accelerate_hpc_test_b200.commands.config.config_args import ClusterConfig, load_config_from_file
from accelerate_hpc_test_b200.commands.config.default import write_basic_config
from accelerate_hpc_test_b200.commands.estimate import estimate_command_parser, gather_data
from accelerate_hpc_test_b200.commands.launch import _validate_launch_command, launch_command_parser
from accelerate_hpc_test_b200.commands.to_fsdp2 import convert_config_to_fsdp2
from accelerate_hpc_test_b200.utils.launch import prepare_multi_gpu_env, prepare_simple_launcher_cmd_env

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))


def test_hyphen_and_underscore_flags_are_equivalent_and_recorded():
    parser = launch_command_parser()
    a = parser.parse_args(["--num-processes", "4", "--mixed-precision=bf16", "train.py", "--lr", "1"])
    b = parser.parse_args(["--num_processes", "4", "--mixed_precision=bf16", "train.py", "--lr", "1"])
    assert a.num_processes == b.num_processes == 4 and a.mixed_precision == b.mixed_precision == "bf16"
    assert a.training_script == "train.py" and a.training_script_args == ["--lr", "1"]
    assert {"num_processes", "mixed_precision"} <= a.nondefault


def test_validate_launch_fills_defaults_without_config(tmp_path, monkeypatch):
    monkeypatch.setattr("accelerate_hpc_test_b200.commands.launch.default_config_file", str(tmp_path / "nope.yaml"))
    parser = launch_command_parser()
    args = parser.parse_args(["--cpu", "train.py"])
    args, defaults, _ = _validate_launch_command(args)
    assert defaults is None and args.num_processes == 1 and args.mixed_precision == "no" and args.num_machines == 1
    with pytest.raises(ValueError):
        _validate_launch_command(parser.parse_args(["--cpu", "--multi_gpu", "train.py"]))


def test_config_file_values_and_cli_override(tmp_path):
    cfg = {
        "compute_environment": "LOCAL_MACHINE", "distributed_type": "FSDP", "mixed_precision": "bf16", "num_processes": 8, "num_machines": 1,
        "machine_rank": 0, "use_cpu": False, "debug": False,
        "fsdp_config": {"fsdp_version": 2, "fsdp_auto_wrap_policy": "TRANSFORMER_BASED_WRAP", "fsdp_transformer_layer_cls_to_wrap": "LlamaDecoderLayer",
                        "fsdp_state_dict_type": "SHARDED_STATE_DICT", "fsdp_reshard_after_forward": True},
        "parallelism_config": {"parallelism_config_dp_shard_size": 4, "parallelism_config_tp_size": 2},
    }
    path = tmp_path / "cfg.yaml"
    path.write_text(yaml.safe_dump(cfg))
    parser = launch_command_parser()
    args = parser.parse_args(["--config_file", str(path), "--num_processes", "4", "train.py"])
    args, defaults, mp_from_config = _validate_launch_command(args)
    assert args.use_fsdp and args.num_processes == 4 and args.mixed_precision == "bf16" and mp_from_config
    assert args.fsdp_version == 2 and args.fsdp_transformer_layer_cls_to_wrap == "LlamaDecoderLayer"
    assert args.use_parallelism_config and args.parallelism_config_tp_size == 2 and args.parallelism_config_dp_shard_size == 4
    env = prepare_multi_gpu_env(args)
    assert env["ACCELERATE_USE_FSDP"] == "true" and env["FSDP_VERSION"] == "2" and env["FSDP_AUTO_WRAP_POLICY"] == "TRANSFORMER_BASED_WRAP"
    assert env["ACCELERATE_MIXED_PRECISION"] == "bf16" and env["PARALLELISM_CONFIG_TP_SIZE"] == "2"
    assert env["ACCELERATE_USE_PARALLELISM_CONFIG"] == "true" and args.nproc_per_node == "4"


def test_unknown_config_keys_are_rejected(tmp_path):
    path = tmp_path / "bad.yaml"
    path.write_text(yaml.safe_dump({"compute_environment": "LOCAL_MACHINE", "distributed_type": "NO", "mixed_precision": "no", "use_cpu": True,
                                    "debug": False, "another_invalid_key": 1, "invalid_key": 2}))
    with pytest.raises(ValueError, match=r"unknown keys \(\['another_invalid_key', 'invalid_key'\]\)"):
        load_config_from_file(str(path))


def test_old_config_versions_load(tmp_path):
    old = {"compute_environment": "LOCAL_MACHINE", "distributed_type": "MULTI_GPU", "fp16": True, "machine_rank": 0, "main_process_ip": None,
           "main_process_port": None, "main_training_function": "main", "num_machines": 1, "num_processes": 2}
    path = tmp_path / "old.yaml"
    path.write_text(yaml.safe_dump(old))
    cfg = load_config_from_file(str(path))
    assert cfg.mixed_precision == "fp16" and cfg.use_cpu is False and cfg.debug is False and cfg.num_processes == 2


def test_write_basic_config_roundtrip(tmp_path):
    path = tmp_path / "default.yaml"
    assert write_basic_config("bf16", str(path)) == path
    cfg = load_config_from_file(str(path))
    assert isinstance(cfg, ClusterConfig) and cfg.mixed_precision == "bf16"
    assert write_basic_config("bf16", str(path)) is False  # never overwrites


def test_simple_launcher_env():
    parser = launch_command_parser()
    args = parser.parse_args(["--cpu", "--debug", "--mixed_precision", "no", "script.py", "--x"])
    args.num_machines = 1
    args.dynamo_backend = "no"
    args.num_cpu_threads_per_process = 3
    cmd, env = prepare_simple_launcher_cmd_env(args)
    assert cmd == [sys.executable, "script.py", "--x"]
    assert env["ACCELERATE_USE_CPU"] == "True" and env["ACCELERATE_DEBUG_MODE"] == "true" and env["OMP_NUM_THREADS"] == "3"
    assert env["ACCELERATE_DYNAMO_BACKEND"] == "NO"


def test_multi_node_env_uses_rdzv_endpoint():
    parser = launch_command_parser()
    args = parser.parse_args(["--multi_gpu", "--num_processes", "16", "--num_machines", "2", "--machine_rank", "1", "--main_process_ip", "10.0.0.1",
                              "--main_process_port", "29777", "--mixed_precision", "bf16", "t.py"])
    args.dynamo_backend = "no"
    args.num_cpu_threads_per_process = 1
    args.gpu_ids = "all"
    env = prepare_multi_gpu_env(args)
    assert args.nproc_per_node == "8" and args.nnodes == "2" and args.node_rank == 1 and args.rdzv_endpoint == "10.0.0.1:29777"
    assert env["ACCELERATE_MIXED_PRECISION"] == "bf16"


def test_to_fsdp2_conversion():
    cfg = {"fsdp_config": {"fsdp_sharding_strategy": "SHARD_GRAD_OP", "fsdp_backward_prefetch": "BACKWARD_PRE", "fsdp_use_orig_params": True,
                           "fsdp_sync_module_states": True, "fsdp_auto_wrap_policy": "TRANSFORMER_BASED_WRAP", "fsdp_offload_params": False}}
    out = convert_config_to_fsdp2(cfg)["fsdp_config"]
    assert out["fsdp_version"] == 2 and out["fsdp_reshard_after_forward"] is False
    assert "fsdp_backward_prefetch" not in out and "fsdp_use_orig_params" not in out and "fsdp_sync_module_states" not in out
    assert out["fsdp_auto_wrap_policy"] == "TRANSFORMER_BASED_WRAP"


def test_estimate_memory_builtin():
    args = estimate_command_parser().parse_args(["llama3-8b", "--dtypes", "float32", "bfloat16"])
    rows = gather_data(args)
    total_fp32 = rows[0][2]
    assert abs(total_fp32 - 8.03e9 * 4) / (8.03e9 * 4) < 0.01
    assert rows[1][2] == total_fp32 / 2
    assert rows[0][3]["optimizer"] == total_fp32 * 2


def test_launch_single_process_subprocess(tmp_path):
    script = tmp_path / "hello.py"
    script.write_text("import os\nprint('MP=' + os.environ['ACCELERATE_MIXED_PRECISION'], 'CPU=' + os.environ['ACCELERATE_USE_CPU'])\n")
    out = subprocess.run([sys.executable, "-m", "accelerate_hpc_test_b200.commands.launch", "--cpu", "--num_processes", "1", "--mixed_precision", "no",
                          "--config_file", _empty_cfg(tmp_path), str(script)], capture_output=True, text=True, env=ENV, cwd=ROOT, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "MP=no CPU=True" in out.stdout


def _empty_cfg(tmp_path):
    p = tmp_path / "base.yaml"
    p.write_text(yaml.safe_dump({"compute_environment": "LOCAL_MACHINE", "distributed_type": "NO", "mixed_precision": "no", "use_cpu": True,
                                 "debug": False, "num_processes": 1, "num_machines": 1, "machine_rank": 0}))
    return str(p)


def test_accelerate_test_script_two_cpu_processes(tmp_path):
    """`accelerate launch --cpu --num_processes 2 -m ...test_script` : the full in-package end-to-end script over gloo."""
    from accelerate_hpc_test_b200.utils.other import get_free_port

    out = subprocess.run([sys.executable, "-m", "accelerate_hpc_test_b200.commands.launch", "--config_file", _empty_cfg(tmp_path), "--cpu",
                          "--num_processes", "2", "--main_process_port", str(get_free_port()), "-m",
                          "accelerate_hpc_test_b200.test_utils.scripts.test_script"], capture_output=True, text=True, env=ENV, cwd=ROOT, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "test_script: all passed" in out.stdout


# ------------------------------------------------------------------------------------------------- config files on disk
CONFIGS = os.path.join(ROOT, "tests", "test_configs")
_VINTAGES = sorted(name for name in os.listdir(CONFIGS) if name.endswith(".yaml") and name != "invalid_keys.yaml")


@pytest.mark.parametrize("name", _VINTAGES)
def test_config_files_of_every_vintage_load_and_launch(name):
    """Every file `accelerate config` ever wrote still loads, and merges into a launch command line without complaint."""
    path = os.path.join(CONFIGS, name)
    cfg = load_config_from_file(path)
    assert isinstance(cfg, ClusterConfig)
    raw = yaml.safe_load(open(path))
    assert cfg.num_processes == raw["num_processes"] and str(cfg.distributed_type).endswith(str(raw["distributed_type"]))
    # keys a vintage did not know come out as today's defaults
    assert cfg.debug is bool(raw.get("debug", False)) and cfg.use_cpu is bool(raw.get("use_cpu", False))
    args, defaults, _ = _validate_launch_command(launch_command_parser().parse_args(["--config_file", path, "train.py"]))
    assert defaults is not None and args.num_processes == raw["num_processes"] and args.mixed_precision == raw.get("mixed_precision", "no")
    assert args.use_fsdp == (raw["distributed_type"] == "FSDP") and args.use_deepspeed == (raw["distributed_type"] == "DEEPSPEED")
    assert args.use_megatron_lm == (raw["distributed_type"] == "MEGATRON_LM")


def test_config_sections_reach_the_launch_arguments():
    parse = lambda name, *extra: _validate_launch_command(
        launch_command_parser().parse_args(["--config_file", os.path.join(CONFIGS, name), *extra, "train.py"]))[0]
    fsdp1 = parse("latest_fsdp1.yaml")
    assert fsdp1.fsdp_sharding_strategy == "SHARD_GRAD_OP" and fsdp1.fsdp_transformer_layer_cls_to_wrap == "LlamaDecoderLayer"
    assert fsdp1.fsdp_activation_checkpointing is True
    nd = parse("latest_fsdp2_nd.yaml")
    assert nd.fsdp_version == 2 and nd.use_parallelism_config and (nd.parallelism_config_dp_shard_size, nd.parallelism_config_tp_size) == (4, 2)
    zero3 = parse("latest_deepspeed.yaml")
    assert zero3.zero_stage == 3 and zero3.gradient_accumulation_steps == 4 and zero3.zero3_init_flag is True
    megatron = parse("latest_megatron.yaml")
    assert (megatron.megatron_lm_tp_degree, megatron.megatron_lm_pp_degree, megatron.megatron_lm_num_micro_batches) == (2, 2, 2)
    assert megatron.megatron_lm_sequence_parallelism is True and megatron.megatron_lm_use_distributed_optimizer is True
    mpi = load_config_from_file(os.path.join(CONFIGS, "v0_28_mpi.yaml"))
    assert mpi.mpirun_config == {"mpirun_hostfile": "/etc/cluster/hosts"} and mpi.num_machines == 2 and mpi.main_process_port == 29511
    fp8 = load_config_from_file(os.path.join(CONFIGS, "v0_34_fp8.yaml"))
    assert fp8.mixed_precision == "fp8" and fp8.fp8_config["backend"] == "TE" and fp8.fp8_config["fp8_format"] == "HYBRID"
    # sections that do not belong to the distributed type stay inert; a flag typed on the command line beats the file
    inert = parse("validate_launch_cmd.yaml", "--num_processes", "2", "--cpu")
    assert inert.debug is True and inert.num_processes == 2 and not inert.use_fsdp and not inert.use_deepspeed


def test_config_file_with_unknown_keys_names_them():
    with pytest.raises(ValueError, match=r"unknown keys \(\['nproc', 'shard_everything'\]\)"):
        load_config_from_file(os.path.join(CONFIGS, "invalid_keys.yaml"))


# ------------------------------------------------------------------------------------------------- every sub-command, end to end
class _Ran:
    def __init__(self, returncode, stdout):
        self.returncode, self.stdout, self.stderr = returncode, stdout, ""


def _cli(*argv):
    """Run ``accelerate <argv>`` in this process through the real parser and handler (a fresh interpreter per call costs seconds)."""
    import contextlib
    import io

    from accelerate_hpc_test_b200.commands.accelerate_cli import build_parser

    captured = io.StringIO()
    code = 0
    with contextlib.redirect_stdout(captured), contextlib.redirect_stderr(captured):
        try:
            parser = build_parser()
            args = parser.parse_args(list(argv))
            handler = getattr(args, "func", None)
            if handler is None:
                parser.print_help()
                code = 1
            else:
                handler(args)
        except SystemExit as stop:
            code = stop.code if isinstance(stop.code, int) else (0 if stop.code is None else 1)
        except Exception as error:  # noqa: BLE001 - a command that raises is a command that failed
            print(f"{type(error).__name__}: {error}")
            code = 1
    return _Ran(code, captured.getvalue())


def test_cli_env_estimate_and_help(tmp_path):
    top = subprocess.run([sys.executable, "-m", "accelerate_hpc_test_b200.commands.accelerate_cli"], cwd=ROOT, env=ENV, capture_output=True,
                         text=True, timeout=300)                                             # once for real: the module entry point
    assert top.returncode == 1 and "accelerate <command> [<args>]" in top.stdout          # no sub-command: help, non-zero
    for command in ("config", "launch", "env", "test", "estimate-memory", "merge-weights", "to-fsdp2", "tpu-config"):
        assert command in top.stdout, command
    config_file = tmp_path / "cfg.yaml"
    assert _cli("config", "default", "--config_file", str(config_file), "--mixed_precision", "bf16").returncode == 0
    env = _cli("env", "--config_file", str(config_file))
    assert env.returncode == 0 and "Copy-and-paste the text below" in env.stdout and "PyTorch version" in env.stdout
    assert "`Accelerate` config passed:" in env.stdout and "mixed_precision: bf16" in env.stdout
    estimate = _cli("estimate-memory", "llama3-1b", "--dtypes", "float32", "int4")
    assert estimate.returncode == 0 and "Largest Layer" in estimate.stdout and "int4" in estimate.stdout and "float32" in estimate.stdout
    unknown = _cli("estimate-memory", "no-such-model")
    assert unknown.returncode != 0
    assert _cli("tpu-config").returncode != 0


def test_cli_config_update_and_to_fsdp2(tmp_path):
    old = tmp_path / "old.yaml"
    old.write_text(yaml.safe_dump({"compute_environment": "LOCAL_MACHINE", "distributed_type": "FSDP", "mixed_precision": "bf16", "num_processes": 2,
                                   "fsdp_config": {"fsdp_sharding_strategy": "FULL_SHARD", "fsdp_backward_prefetch": "BACKWARD_PRE",
                                                   "fsdp_auto_wrap_policy": "TRANSFORMER_BASED_WRAP", "fsdp_use_orig_params": True,
                                                   "fsdp_transformer_layer_cls_to_wrap": "LlamaDecoderLayer"}}))
    updated = _cli("config", "update", "--config_file", str(old))
    assert updated.returncode == 0 and "Sucessfully updated the configuration file" in updated.stdout
    refreshed = yaml.safe_load(old.read_text())
    assert refreshed["debug"] is False and refreshed["use_cpu"] is False and refreshed["num_processes"] == 2      # today's keys were added
    new = tmp_path / "new.yaml"
    converted = _cli("to-fsdp2", "--config_file", str(old), "--output_file", str(new))
    assert converted.returncode == 0, converted.stdout[-2000:]
    section = yaml.safe_load(new.read_text())["fsdp_config"]
    assert section["fsdp_version"] == 2 and section["fsdp_reshard_after_forward"] is True
    assert "fsdp_backward_prefetch" not in section and "fsdp_use_orig_params" not in section and "fsdp_sharding_strategy" not in section
    assert section["fsdp_transformer_layer_cls_to_wrap"] == "LlamaDecoderLayer"
    refused = _cli("to-fsdp2", "--config_file", str(old), "--output_file", str(new))
    assert refused.returncode != 0                                                                               # never overwrites silently
    assert _cli("to-fsdp2", "--config_file", str(old), "--output_file", str(new), "--overwrite").returncode == 0
