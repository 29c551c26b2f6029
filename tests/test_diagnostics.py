"""Diagnostics: device timer, fault injector parsing, and the elastic-recovery drill (kill a rank mid-run, restart, resume
from the last checkpoint, same final weights as an uninterrupted run).  SURVEY.md §5.1-5.3."""

import json
import os
import subprocess
import sys

import pytest
import torch

from accelerate_hpc_test_b200.utils.diagnostics import DeviceTimer, FaultInjector, shadow_check

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_device_timer_cpu():
    t = DeviceTimer()
    for _ in range(3):
        with t.region("work"):
            torch.randn(64, 64) @ torch.randn(64, 64)
    s = t.summary(reduce=False)
    assert s["work"]["calls"] == 3 and s["work"]["ms"] > 0 and t.summary(reduce=False) == {}


def test_fault_injector_spec(monkeypatch):
    inj = FaultInjector("raise:rank=1,step=3")
    inj.maybe_fail(3, rank=0)
    inj.maybe_fail(2, rank=1)
    with pytest.raises(RuntimeError, match="fault injection"):
        inj.maybe_fail(3, rank=1)
    monkeypatch.setenv("TORCHELASTIC_RESTART_COUNT", "1")
    inj.maybe_fail(3, rank=1)  # only the first attempt fails
    with pytest.raises(ValueError):
        FaultInjector("explode:rank=0")
    assert not FaultInjector("").armed


def test_shadow_check_reports_first_difference():
    a = torch.arange(8.0)
    shadow_check("same", a, a.clone())
    b = a.clone()
    b[5] = -1
    with pytest.raises(RuntimeError, match="flat index 5"):
        shadow_check("all_gather[x]", a, b)


def _drill(tmp_path, name, fault):
    out = tmp_path / name
    out.mkdir()
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), OMP_NUM_THREADS="1")
    env.pop("ACCELERATE_B200_FAULT_INJECT", None)
    if fault:
        env["ACCELERATE_B200_FAULT_INJECT"] = fault
    cmd = [sys.executable, "-m", "accelerate_hpc_test_b200.commands.accelerate_cli", "launch", "--num_processes", "2", "--cpu",
           "--max_restarts", "1", "--monitor_interval", "0.5", "--main_process_port", str(29900 + os.getpid() % 90),
           "-m", "accelerate_hpc_test_b200.test_utils.scripts.test_fault_recovery", str(out), "6"]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    return json.load(open(out / "final.json")), r.stdout + r.stderr


def test_kill_restart_resume_matches_uninterrupted_run(tmp_path):
    clean, _ = _drill(tmp_path, "clean", None)
    faulty, log = _drill(tmp_path, "faulty", "kill:rank=1,step=3")
    assert clean["attempt"] == 0 and faulty["attempt"] == 1, (clean, faulty)
    assert "injected `kill` at optimizer step 3" in log and "resumed after step 2" in log
    assert abs(clean["a"] - faulty["a"]) < 1e-6 and abs(clean["b"] - faulty["b"]) < 1e-6, (clean, faulty)
