"""The example scripts run, and the "complete" example contains what the by-feature examples add
(reference tests/test_examples.py: ExampleDifferenceTests :70-155, FeatureExamplesTests :157-330).
BASELINE config 1: BERT sequence classification (nlp_example.py shape, synthetic MRPC) on 2 CPU / gloo processes."""

import os
import subprocess
import sys

import pytest

from accelerate_hpc_test_b200.test_utils.examples import compare_against_test

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EX = os.path.join(ROOT, "examples")
SMALL = ["--cpu", "--model", "tiny", "--num_epochs", "1", "--train_size", "64", "--eval_size", "41"]


def _run(cmd, cwd=None, timeout=420):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, cwd=cwd or ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout


def _run_inproc(script, *args):
    """Run an example's ``main()`` in this process (fresh state singletons), returning what it printed."""
    import contextlib
    import io
    import runpy

    from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState

    for st in (AcceleratorState, GradientState, PartialState):
        st._reset_state(True) if st is AcceleratorState else st._reset_state()
    path = os.path.join(EX, script)
    old_argv, old_path, old_env = sys.argv, list(sys.path), dict(os.environ)
    sys.argv = [path, *args]
    sys.path.insert(0, os.path.dirname(path))  # what `python script.py` does
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            runpy.run_path(path, run_name="__main__")
    finally:
        sys.argv, sys.path[:] = old_argv, old_path
        os.environ.clear()
        os.environ.update(old_env)  # e.g. ACCELERATE_USE_DEEPSPEED, which an Accelerator with a DeepSpeed plugin exports
        for mod in ("_common", "nlp_example", "cv_example"):
            sys.modules.pop(mod, None)
    return buf.getvalue()


def _launch(nproc, script, *args):
    return _run([sys.executable, "-m", "accelerate_hpc_test_b200.commands.accelerate_cli", "launch", "--num_processes", str(nproc), "--cpu",
                 "--main_process_port", str(29650 + (os.getpid() + hash(script)) % 200), os.path.join(EX, script), *args])


def test_nlp_example_two_cpu_processes():
    out = _launch(2, "nlp_example.py", *SMALL)
    assert "epoch 0:" in out and "'eval_samples': 41" in out  # gather_for_metrics dropped the padding duplicates


@pytest.mark.parametrize("script,extra", [
    ("by_feature/gradient_accumulation.py", ["--gradient_accumulation_steps", "2"]),
    ("by_feature/memory.py", []),
    ("by_feature/early_stopping.py", []),
    ("by_feature/profiler.py", []),
    ("by_feature/automatic_gradient_accumulation.py", []),
])
def test_by_feature_single_process(script, extra):
    assert "epoch 0" in _run_inproc(script, *SMALL, *extra)


def test_more_by_feature_examples_single_process():
    out = _run_inproc("by_feature/cross_validation.py", *SMALL, "--num_folds", "2")
    assert "fold 1 epoch 0" in out and "average of 2 folds on the test set" in out and "'eval_samples': 41" in out
    assert "epoch 0" in _run_inproc("by_feature/schedule_free.py", *SMALL)
    out = _run_inproc("by_feature/fsdp_with_peak_mem_tracking.py", *SMALL, "--use_fsdp")  # (two ranks: test_launched_scripts' memory script)
    assert "train: device memory used" in out and "epoch 0" in out
    out = _run_inproc("by_feature/deepspeed_with_config_support.py", *SMALL,
                      "--deepspeed_config", os.path.join(EX, "deepspeed_config_templates", "zero_stage3_config.json"))
    assert "optimizer: FusedAdamW, ZeRO stage 3" in out and "epoch 0" in out
    out = _run_inproc("by_feature/gradient_accumulation_for_autoregressive_models.py", "--cpu", "--model", "tiny", "--train_size", "20",
                      "--eval_size", "8", "--max_length", "24", "--gradient_accumulation_steps", "2")
    assert "perplexity" in out
    out = _run_inproc("by_feature/megatron_lm_gpt_pretraining.py", "--cpu", "--model", "tiny", "--block_size", "16", "--train_size", "16",
                      "--eval_size", "8", "--max_train_steps", "3")
    assert "steps: 3" in out


def test_cv_examples(tmp_path):
    small = ["--cpu", "--train_size", "128", "--eval_size", "37", "--image_size", "16", "--num_epochs", "1"]
    assert "'eval_samples': 37" in _run_inproc("cv_example.py", *small)
    ck = str(tmp_path / "cv")
    _run_inproc("complete_cv_example.py", *small, "--checkpointing_steps", "1", "--output_dir", ck, "--with_tracking",
                "--project_dir", str(tmp_path / "logs"))
    assert sorted(os.listdir(ck)) == ["step_1", "step_2"]
    out = _run_inproc("complete_cv_example.py", *small, "--resume_from_checkpoint", os.path.join(ck, "step_1"))
    assert "Resumed from checkpoint" in out and "epoch 0" in out


def test_deepspeed_templates_and_slurm_scripts_are_well_formed():
    import json

    from accelerate_hpc_test_b200.utils import DeepSpeedPlugin

    folder = os.path.join(EX, "deepspeed_config_templates")
    stages = {}
    for name in sorted(os.listdir(folder)):
        cfg = json.load(open(os.path.join(folder, name)))
        plugin = DeepSpeedPlugin(hf_ds_config=os.path.join(folder, name))
        stages[name] = plugin.get_value("zero_optimization.stage")
        assert plugin.is_auto("train_micro_batch_size_per_gpu") and cfg["optimizer"]["type"] == "AdamW"
    assert stages == {"zero_stage1_config.json": 1, "zero_stage2_config.json": 2, "zero_stage2_offload_config.json": 2,
                      "zero_stage3_config.json": 3, "zero_stage3_offload_config.json": 3}
    for name in os.listdir(os.path.join(EX, "slurm")):
        if name.endswith(".sh"):
            assert subprocess.run(["bash", "-n", os.path.join(EX, "slurm", name)]).returncode == 0, name


@pytest.mark.parametrize("script,extra", [("by_feature/local_sgd.py", ["--local_sgd_steps", "2"]), ("by_feature/multi_process_metrics.py", []),
                                          ("by_feature/ddp_comm_hook.py", ["--ddp_comm_hook", "bf16"])])
def test_by_feature_two_processes(script, extra):
    assert "'eval_samples': 41" in _launch(2, script, *SMALL, *extra)


def test_checkpointing_and_tracking_examples(tmp_path):
    ck = str(tmp_path / "ck")
    _run_inproc("by_feature/checkpointing.py", *SMALL, "--checkpointing_steps", "2", "--output_dir", ck)
    assert os.path.isdir(os.path.join(ck, "step_2"))
    out = _run_inproc("by_feature/checkpointing.py", *SMALL, "--resume_from_checkpoint", os.path.join(ck, "step_2"))
    assert "epoch 0" in out
    logs = str(tmp_path / "logs")
    _run_inproc("by_feature/tracking.py", *SMALL, "--with_tracking", "--project_dir", logs)
    assert os.path.isfile(os.path.join(logs, "tracking", "metrics.jsonl"))
    out = _run_inproc("complete_nlp_example.py", *SMALL, "--checkpointing_steps", "epoch", "--output_dir", ck,
                      "--with_tracking", "--project_dir", logs)
    assert os.path.isfile(os.path.join(ck, "epoch_0", "model.safetensors"))


def test_native_parallelism_examples_on_cpu():
    out = _launch(4, "torch_native_parallelism/nd_parallel.py", "--cpu", "--model", "tiny", "--tp-size", "2", "--dp-shard-size", "2",
                  "--sequence-length", "64", "--num-steps", "3")
    assert "tokens/s over 4 processes" in out
    out = _launch(2, "alst_ulysses_sequence_parallelism/sp_alst.py", "--cpu", "--model", "tiny", "--sp-size", "2", "--sequence-length", "128",
                  "--num-steps", "2")
    assert "sp=2" in out
    out = _launch(2, "inference/distributed_split.py", "--cpu")
    assert out.count("generated") == 5


@pytest.mark.parametrize("feature", ["checkpointing.py", "tracking.py"])
def test_complete_example_covers_feature(feature, monkeypatch):
    """Everything a by-feature script adds to its training function is also in the complete example."""
    monkeypatch.chdir(ROOT)
    complete = os.path.join(EX, "complete_nlp_example.py")
    diff = compare_against_test(complete, os.path.join(EX, "by_feature", feature), parser_only=False)
    # the by-feature scripts are written against the shared helper module, so only feature-API lines are compared
    api = [ln for ln in diff if any(k in ln for k in ("save_state", "load_state", "skip_first_batches", "init_trackers", "accelerator.log(",
                                                      "end_training"))]
    missing = []
    text = open(complete).read()
    for ln in api:
        call = ln.strip().split("(")[0].split("=")[-1].strip()
        if call and call not in text:
            missing.append(ln)
    assert not missing, missing


def test_example_side_files_are_well_formed():
    """Launcher configurations, the ZeRO configuration and the helper module shipped next to the examples load and mean what their
    READMEs say."""
    import importlib.util

    from accelerate_hpc_test_b200.commands.launch import _validate_launch_command, launch_command_parser
    from accelerate_hpc_test_b200.models import LlamaConfig
    from accelerate_hpc_test_b200.utils import DeepSpeedPlugin

    def merged(path):
        return _validate_launch_command(launch_command_parser().parse_args(["--config_file", os.path.join(ROOT, "examples", path), "train.py"]))[0]

    cp = merged("torch_native_parallelism/configs/cp.yaml")
    assert cp.use_fsdp and cp.parallelism_config_cp_size == 4 and cp.num_processes == 4
    hsdp = merged("torch_native_parallelism/configs/tp_hsdp.yaml")
    assert (hsdp.parallelism_config_dp_replicate_size, hsdp.parallelism_config_dp_shard_size, hsdp.parallelism_config_tp_size) == (2, 2, 2)
    sp = merged("alst_ulysses_sequence_parallelism/sp_alst.accelerate_config.yml")
    assert sp.parallelism_config_sp_size == 8 and sp.num_processes == 8 and sp.mixed_precision == "bf16"
    zero3 = DeepSpeedPlugin(hf_ds_config=os.path.join(ROOT, "examples", "alst_ulysses_sequence_parallelism", "sp_alst.ds_config.json"))
    assert zero3.zero_stage == 3 and zero3.stated_precision() == "bf16"
    assert os.access(os.path.join(ROOT, "examples", "alst_ulysses_sequence_parallelism", "sp_alst.sh"), os.X_OK)

    spec = importlib.util.spec_from_file_location("nd_utils", os.path.join(ROOT, "examples", "torch_native_parallelism", "utils.py"))
    helpers = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(helpers)
    config = LlamaConfig.llama3_8b()
    flops = helpers.get_model_flops_per_token(config, 2048)
    matmul_weights = config.num_parameters() - config.vocab_size * config.hidden_size          # the embedding table does no matmul
    assert 6 * matmul_weights < flops < 6.4 * matmul_weights                                   # 6N plus a few percent of attention
    tracker = helpers.PerformanceTracker(warmup_steps=2, flops_per_token=flops, device="cpu")
    for _ in range(5):
        tracker.step(4096)
    measured = tracker.metrics()
    assert measured["steps"] == 3 and measured["total_tokens"] == 3 * 4096 and measured["tflops_per_device"] > 0
    batch = helpers.collate(helpers.synthetic_tokens(3, 16, 100))
    assert batch["input_ids"].shape == (3, 16) and batch["labels"] is not batch["input_ids"]


def test_config_template_reporter_runs():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "config_yaml_templates", "run_me.py")], cwd=ROOT,
                         env=dict(os.environ, ACCELERATE_USE_CPU="true"), capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and "Accelerator state from the current environment" in res.stdout, res.stdout + res.stderr[-2000:]
