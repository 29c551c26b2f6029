"""The in-package launched scripts (``test_utils/scripts`` and ``scripts/external_deps``), each run the way a user would run it:
through ``accelerate launch`` on two CPU / gloo processes (reference tests/test_multigpu.py, tests/test_metrics.py,
tests/deepspeed/*, tests/test_cli.py launch the same-named scripts on GPUs)."""

import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPTS = "accelerate_hpc_test_b200.test_utils.scripts"


def _env():
    return dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), OMP_NUM_THREADS="1")


def _launch(module, *args, nproc=2, timeout=600):
    port = 29800 + (os.getpid() * 7 + abs(hash(module))) % 150
    cmd = [sys.executable, "-m", "accelerate_hpc_test_b200.commands.accelerate_cli", "launch", "--num_processes", str(nproc), "--cpu",
           "--main_process_port", str(port), "-m", module, *args]
    done = subprocess.run(cmd, cwd=ROOT, env=_env(), capture_output=True, text=True, timeout=timeout)
    assert done.returncode == 0, done.stdout[-1500:] + done.stderr[-3000:]
    return done.stdout


@pytest.mark.parametrize("script,args", [
    ("test_distributed_data_loop", []),
    ("test_ddp_comm_hook", []),
    ("test_merge_weights", []),
    ("external_deps.test_metrics", []),
    ("external_deps.test_checkpointing", ["--num_epochs", "2"]),
    ("external_deps.test_performance", ["--performance_lower_bound", "0.8", "--num_epochs", "6"]),
    ("external_deps.test_peak_memory_usage", ["--use_fsdp", "--peak_memory_upper_bound", "4000"]),
    ("external_deps.test_pippy", []),
    ("external_deps.test_zero3_integration", []),
    ("external_deps.test_ds_multiple_model", []),
    ("external_deps.test_ds_alst_ulysses_sp", []),
])
def test_script_on_two_cpu_processes(script, args):
    out = _launch(f"{SCRIPTS}.{script}", *args)
    assert f"{script.split('.')[-1]}: " in out and "all passed" in out


def test_cli_script_reports_every_rank():
    out = _launch(f"{SCRIPTS}.test_cli")
    assert "rank 0 of 2" in out and "rank 1 of 2" in out and "Successfully ran on 2 processes" in out


def test_performance_script_fails_below_the_bound():
    cmd = [sys.executable, "-m", f"{SCRIPTS}.external_deps.test_performance", "--performance_lower_bound", "1.01", "--num_epochs", "1",
           "--train_size", "64", "--eval_size", "32"]
    done = subprocess.run(cmd, cwd=ROOT, env=_env(), capture_output=True, text=True, timeout=300)
    assert done.returncode != 0 and "is lower than the lower bound" in done.stderr


def test_notebook_script():
    done = subprocess.run([sys.executable, "-m", f"{SCRIPTS}.test_notebook"], cwd=ROOT, env=_env(), capture_output=True, text=True, timeout=600)
    assert done.returncode == 0, done.stdout[-1500:] + done.stderr[-3000:]
    assert "test_notebook: all passed" in done.stdout
