"""The parity benchmarks assert before they report; their CPU forms run here so they cannot rot (SURVEY §4.4 "benchmarks as tests":
the reference's benchmarks/fsdp2 asserts equal loss curves, benchmarks/fp8 equal metrics)."""

import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))


def _last_json(text):
    rows = [json.loads(line) for line in text.splitlines() if line.startswith("{")]
    assert rows, text[-3000:]
    return rows


def test_fsdp2_parity_benchmark_two_cpu_processes(tmp_path):
    out = tmp_path / "parity.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port",
           "29647", os.path.join(ROOT, "benchmarks", "fsdp2_parity.py"), "--cpu", "--model", "tiny", "--steps", "8", "--out", str(out)]
    res = subprocess.run(cmd, cwd=ROOT, env=ENV, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    record = _last_json(res.stdout)[-1]
    assert record["world"] == 2 and record["max_rel_loss_diff"] <= record["rtol"] and record["last_loss"] < record["first_loss"]
    assert record["accelerate"]["engine"] == "B200ShardedModule" and record["accelerate"]["ms_per_step"] > 0
    saved = json.loads(out.read_text())
    assert len(saved["losses_accelerate"]) == len(saved["losses_torch_fsdp2"]) == 8


def test_big_model_inference_benchmark_same_tokens_for_every_placement(tmp_path):
    out = tmp_path / "bmi.json"
    res = subprocess.run([sys.executable, os.path.join(ROOT, "benchmarks", "big_model_inference.py"), "--model", "tiny", "--cpu",
                          "--new-tokens", "4", "--out", str(out)], cwd=ROOT, env=ENV, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    rows = {row["placement"]: row for row in _last_json(res.stdout)}
    assert {"one_device", "half_on_disk"} <= set(rows) and all(row["same_tokens"] for row in rows.values())
    assert all(row["load_s"] > 0 and row["per_token_s"] > 0 for row in rows.values())
    assert json.loads(out.read_text())["new_tokens"] == 4


def test_regional_compilation_benchmark_structure():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "benchmarks", "regional_compilation.py"), "--backend", "eager", "--cpu", "--layers",
                          "3", "--hidden", "64", "--iters", "2", "--seq-len", "16"], cwd=ROOT, env=ENV, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    record = _last_json(res.stdout)[-1]
    assert record["regional_keeps_original"] and record["regional_max_abs_err"] < 1e-3
    assert {"eager", "full", "regional"} <= set(record)
