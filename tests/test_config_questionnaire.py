"""`accelerate config` questionnaire, config schema I/O, `config default` and `config update`.

The questionnaire is driven with scripted answers (stdin is not a TTY under pytest, so menus fall back to a numbered
prompt and every answer is one `input()` call).  An empty string accepts the default.
"""

import argparse
import builtins
import json

import pytest
import yaml

from accelerate_hpc_test_b200.commands.config import cluster, get_config_parser
from accelerate_hpc_test_b200.commands.config import config_args as ca
from accelerate_hpc_test_b200.commands.config.config_utils import (
    _ask_field,
    _convert_yes_no_to_bool,
    pick_from,
)
from accelerate_hpc_test_b200.commands.config.default import write_basic_config
from accelerate_hpc_test_b200.commands.config.update import update_config
from accelerate_hpc_test_b200.utils.dataclasses import DistributedType


def script(monkeypatch, answers):
    it = iter(answers)
    asked = []

    def fake_input(prompt=""):
        asked.append(prompt)
        try:
            return next(it)
        except StopIteration:
            raise AssertionError(f"questionnaire asked more than scripted; last prompt: {prompt!r}")

    monkeypatch.setattr(builtins, "input", fake_input)
    return asked, it


def leftovers(it):
    return list(it)


# --------------------------------------------------------------------------- prompt helpers
def test_yes_no_parser_accepts_short_forms_and_rejects_garbage():
    assert _convert_yes_no_to_bool("YES") is True and _convert_yes_no_to_bool("n") is False
    with pytest.raises(ValueError):
        _convert_yes_no_to_bool("maybe")


def test_ask_field_repeats_until_parse_succeeds(monkeypatch, capsys):
    script(monkeypatch, ["abc", "", "7"])
    assert _ask_field("n? ", int, error_message="need int") == 7
    assert capsys.readouterr().out.count("need int") == 2  # "abc" and "" (no default given) both fail


def test_ask_field_empty_returns_default(monkeypatch):
    script(monkeypatch, [""])
    assert _ask_field("n? ", int, default=3) == 3


def test_pick_from_wraps_values():
    assert pick_from(["a", "b"])(1) == "b"
    assert pick_from(["NO", "MULTI_GPU"], DistributedType)("1") == DistributedType.MULTI_GPU


# --------------------------------------------------------------------------- questionnaire branches
def test_single_process_gpu_bf16(monkeypatch):
    #           machine  cpu-only dynamo deepspeed gpu_ids numa precision
    _, it = script(monkeypatch, ["0", "", "", "", "", "", "2"])
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.distributed_type == DistributedType.NO and cfg.use_cpu is False
    assert cfg.mixed_precision == "bf16" and cfg.num_processes == 1 and cfg.gpu_ids == "all"
    assert cfg.fsdp_config == {} and cfg.deepspeed_config == {} and cfg.dynamo_config == {}


def test_cpu_only_skips_gpu_questions(monkeypatch):
    asked, it = script(monkeypatch, ["0", "yes", ""])
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.use_cpu is True and cfg.mixed_precision == "no" and cfg.gpu_ids is None
    assert not any("DeepSpeed" in p or "GPU(s)" in p for p in asked)


def test_multi_node_multi_gpu_asks_rendezvous(monkeypatch):
    answers = ["2",  # multi-GPU
               "2",  # machines
               "1",  # rank of this machine
               "10.0.0.1", "29500",
               "no", "c10d",  # not same network -> rendezvous backend
               "yes",  # debug
               "", "", "", "",  # dynamo, deepspeed, fsdp, megatron: no
               "16", "", "",  # processes, gpu ids, numa
               "0"]  # precision no
    _, it = script(monkeypatch, answers)
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert (cfg.num_machines, cfg.machine_rank, cfg.main_process_ip, cfg.main_process_port) == (2, 1, "10.0.0.1", 29500)
    assert cfg.same_network is False and cfg.rdzv_backend == "c10d" and cfg.debug is True
    assert cfg.distributed_type == DistributedType.MULTI_GPU and cfg.num_processes == 16


def test_fsdp2_with_parallelism_config(monkeypatch):
    answers = ["2", "", "",  # multi-GPU, 1 machine, debug no
               "",  # dynamo no
               "",  # deepspeed no
               "yes",  # fsdp
               "",  # version -> default index 1 == version 2
               "",  # reshard after forward: YES
               "",  # offload: no
               "0",  # transformer based wrap
               "",  # use _no_split_modules: no
               "LlamaDecoderLayer",
               "",  # state dict type -> SHARDED
               "",  # cpu ram efficient loading: yes
               "yes",  # activation checkpointing
               "yes",  # parallelism config
               "1", "4", "2", "1",  # dp_replicate, dp_shard, tp, cp
               "1",  # sp
               "8", ""]  # processes, gpu ids (numa is not asked for FSDP)
    answers.append("2")  # bf16
    _, it = script(monkeypatch, answers)
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.distributed_type == DistributedType.FSDP
    f = cfg.fsdp_config
    assert f["fsdp_version"] == 2 and f["fsdp_reshard_after_forward"] is True and f["fsdp_offload_params"] is False
    assert f["fsdp_auto_wrap_policy"] == "TRANSFORMER_BASED_WRAP" and f["fsdp_transformer_layer_cls_to_wrap"] == "LlamaDecoderLayer"
    assert f["fsdp_state_dict_type"] == "SHARDED_STATE_DICT" and f["fsdp_activation_checkpointing"] is True
    assert "fsdp_backward_prefetch" not in f and "fsdp_use_orig_params" not in f  # FSDP1-only questions
    p = cfg.parallelism_config
    assert (p["parallelism_config_dp_shard_size"], p["parallelism_config_tp_size"], p["parallelism_config_cp_size"]) == (4, 2, 1)
    assert "parallelism_config_cp_comm_strategy" not in p
    assert cfg.num_processes == 8 and cfg.mixed_precision == "bf16"


def test_fsdp1_asks_legacy_questions(monkeypatch):
    answers = ["2", "", "", "", "",
               "yes", "0",  # fsdp, version 1
               "0",  # FULL_SHARD
               "", "1",  # offload no, SIZE_BASED_WRAP
               "2000",  # min params
               "0",  # BACKWARD_PRE
               "",  # state dict type default (SHARDED)
               "", "",  # forward prefetch no, use_orig_params yes
               "no",  # ram efficient loading off -> sync_module_states is asked
               "",  # sync module states yes
               "",  # activation checkpointing no
               "2", "", "1"]
    _, it = script(monkeypatch, answers)
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    f = cfg.fsdp_config
    assert f["fsdp_version"] == 1 and f["fsdp_reshard_after_forward"] == "FULL_SHARD" and f["fsdp_min_num_params"] == 2000
    assert f["fsdp_backward_prefetch"] == "BACKWARD_PRE" and f["fsdp_state_dict_type"] == "SHARDED_STATE_DICT"
    assert f["fsdp_use_orig_params"] is True and f["fsdp_sync_module_states"] is True and f["fsdp_cpu_ram_efficient_loading"] is False
    assert cfg.parallelism_config == {} and cfg.mixed_precision == "fp16"


def test_deepspeed_zero3_with_offload(monkeypatch):
    answers = ["2", "", "", "",
               "yes",  # deepspeed
               "",  # no json file
               "3",  # stage 3
               "1", "2",  # optimizer -> cpu, params -> nvme
               "/mnt/nvme",  # nvme path for params
               "4",  # grad accumulation
               "yes", "0.5",  # clipping
               "yes",  # save 16 bit
               "yes",  # zero3 init
               "",  # moe no
               "8", "",  # processes, gpu ids
               "2"]
    _, it = script(monkeypatch, answers)
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    d = cfg.deepspeed_config
    assert cfg.distributed_type == DistributedType.DEEPSPEED
    assert d["zero_stage"] == 3 and d["offload_optimizer_device"] == "cpu" and d["offload_param_device"] == "nvme"
    assert d["offload_param_nvme_path"] == "/mnt/nvme" and "offload_optimizer_nvme_path" not in d
    assert d["gradient_accumulation_steps"] == 4 and d["gradient_clipping"] == 0.5
    assert d["zero3_save_16bit_model"] is True and d["zero3_init_flag"] is True


def test_deepspeed_config_file_leaves_precision_to_the_json(monkeypatch):
    answers = ["2", "", "", "", "yes", "yes", "ds.json", "", "", "4", ""]
    _, it = script(monkeypatch, answers)
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.deepspeed_config == {"deepspeed_config_file": "ds.json", "zero3_init_flag": False}
    assert cfg.mixed_precision is None


def test_megatron_branch(monkeypatch):
    answers = ["2", "", "", "", "", "",  # ... deepspeed no, fsdp no
               "yes", "2", "", "2", "8", "", "", "0",  # megatron: tp 2, seq-parallel yes, pp 2, 8 micro-batches, recompute, dist-opt, clip 0
               "4", "", "2"]
    _, it = script(monkeypatch, answers)
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    m = cfg.megatron_lm_config
    assert cfg.distributed_type == DistributedType.MEGATRON_LM
    assert m == {"megatron_lm_tp_degree": 2, "megatron_lm_sequence_parallelism": True, "megatron_lm_pp_degree": 2,
                 "megatron_lm_num_micro_batches": 8, "megatron_lm_recompute_activations": True,
                 "megatron_lm_use_distributed_optimizer": True, "megatron_lm_gradient_clipping": 0.0}


@pytest.mark.parametrize("backend_idx,extra,expect", [
    ("0", ["0"], {"backend": "NATIVE", "recipe": "mxfp8"}),
    ("2", ["", "no"], {"backend": "AO", "enable_fsdp_float8_all_gather": True, "pad_inner_dim": False}),
    ("3", [""], {"backend": "MSAMP", "optimization_level": "O2"}),
])
def test_fp8_backends(monkeypatch, backend_idx, extra, expect):
    _, it = script(monkeypatch, ["0", "", "", "", "", "", "3", backend_idx, *extra])
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.mixed_precision == "fp8" and cfg.fp8_config == expect


def test_fp8_transformer_engine_recipe(monkeypatch):
    te = ["1",  # TE
          "", "2", "", "1", "16", "1",  # eval autocast no, margin 2, interval 1, E4M3, history 16, most_recent
          "yes", "yes", "", "yes",  # override: fprop, (dgrad no), wgrad
          ""]  # mxfp8 yes
    _, it = script(monkeypatch, ["0", "", "", "", "", "", "3", *te])
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.fp8_config == {"backend": "TE", "use_autocast_during_eval": False, "margin": 2, "interval": 1, "fp8_format": "E4M3",
                              "amax_history_length": 16, "amax_compute_algorithm": "most_recent",
                              "override_linear_precision": (True, False, True), "use_mxfp8_block_scaling": True}


def test_dynamo_customisation(monkeypatch):
    _, it = script(monkeypatch, ["0", "", "yes", "", "yes", "2", "yes", "", "yes", "", "", "", "0"])
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.dynamo_config == {"dynamo_backend": "INDUCTOR", "dynamo_mode": "max-autotune", "dynamo_use_fullgraph": True,
                                 "dynamo_use_dynamic": False, "dynamo_use_regional_compilation": True}


def test_multi_cpu_with_mpirun(monkeypatch):
    _, it = script(monkeypatch, ["1", "2", "0", "host0", "29400", "", "", "yes", "~/hosts", "2", "", "4"])
    cfg = cluster.get_cluster_input()
    assert leftovers(it) == []
    assert cfg.distributed_type == DistributedType.MULTI_CPU and cfg.use_cpu is True and cfg.num_processes == 4
    assert cfg.mpirun_config["mpirun_hostfile"].endswith("/hosts") and "~" not in cfg.mpirun_config["mpirun_hostfile"]
    assert cfg.mpirun_config["mpirun_ccl"] == "2" and cfg.mixed_precision == "no"


# --------------------------------------------------------------------------- schema I/O
def test_yaml_and_json_round_trip(tmp_path):
    cfg = ca.ClusterConfig(compute_environment="LOCAL_MACHINE", distributed_type="FSDP", mixed_precision="bf16", use_cpu=False,
                           debug=False, num_processes=8, fsdp_config={"fsdp_version": 2})
    for name in ("c.yaml", "c.json"):
        path = tmp_path / name
        cfg.save(path)
        back = ca.load_config_from_file(str(path))
        assert back.distributed_type == DistributedType.FSDP and back.fsdp_config == {"fsdp_version": 2} and back.num_processes == 8
    raw = yaml.safe_load(open(tmp_path / "c.yaml"))
    assert raw["distributed_type"] == "FSDP" and "deepspeed_config" not in raw  # enums as strings, empty sections dropped
    assert json.load(open(tmp_path / "c.json"))["mixed_precision"] == "bf16"


def test_legacy_keys_are_upgraded(tmp_path):
    path = tmp_path / "old.yaml"
    path.write_text(yaml.safe_dump({"distributed_type": "NO", "fp16": True, "dynamo_backend": "INDUCTOR"}))
    cfg = ca.load_config_from_file(str(path))
    assert cfg.mixed_precision == "fp16" and cfg.dynamo_config == {"dynamo_backend": "INDUCTOR"}
    assert cfg.num_processes == 1 and cfg.use_cpu is False and cfg.debug is False and cfg.enable_cpu_affinity is False


def test_missing_distributed_type_and_unknown_keys(tmp_path):
    p = tmp_path / "bad.yaml"
    p.write_text(yaml.safe_dump({"mixed_precision": "no"}))
    with pytest.raises(ValueError, match="distributed_type"):
        ca.load_config_from_file(str(p))
    p.write_text(yaml.safe_dump({"distributed_type": "NO", "mixed_precision": "no", "zzz": 1, "aaa": 2}))
    with pytest.raises(ValueError, match=r"\['aaa', 'zzz'\]"):
        ca.load_config_from_file(str(p))
    with pytest.raises(FileNotFoundError):
        ca.load_config_from_file(str(tmp_path / "nope.yaml"))


def test_write_basic_config_never_overwrites(tmp_path, capsys):
    target = tmp_path / "sub" / "default.yaml"
    assert write_basic_config("bf16", str(target)) == target
    assert write_basic_config("no", str(target)) is False
    assert "will not override" in capsys.readouterr().out
    assert ca.load_config_from_file(str(target)).mixed_precision == "bf16"
    with pytest.raises(ValueError):
        write_basic_config("int4", str(tmp_path / "x.yaml"))


def test_config_update_rewrites_in_place(tmp_path):
    path = tmp_path / "cfg.yaml"
    path.write_text(yaml.safe_dump({"distributed_type": "MULTI_GPU", "fp16": False, "num_processes": 2}))
    assert update_config(argparse.Namespace(config_file=str(path))) == str(path)
    raw = yaml.safe_load(path.read_text())
    assert "fp16" not in raw and raw["use_cpu"] is False and raw["compute_environment"] == "LOCAL_MACHINE"
    with pytest.raises(ValueError):
        update_config(argparse.Namespace(config_file=str(tmp_path / "missing.yaml")))


def test_config_parser_has_default_and_update_subcommands(tmp_path):
    parser = get_config_parser()
    ns = parser.parse_args(["default", "--config_file", str(tmp_path / "d.yaml"), "--mixed_precision", "bf16"])
    ns.func(ns)
    assert (tmp_path / "d.yaml").exists()
    ns = parser.parse_args(["update", "--config_file", str(tmp_path / "d.yaml")])
    ns.func(ns)
