"""Launches the in-package multi-GPU check script under torchrun when >= 2 GPUs are visible."""

import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_multigpu_collectives_and_engines():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", "-m", "accelerate_hpc_test_b200.test_utils.scripts.test_comm_gpu"]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "ALL MULTI-GPU CHECKS PASSED" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


@pytest.mark.gpu
def test_transformers_llama_under_tensor_parallelism_on_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29534", "-m", "accelerate_hpc_test_b200.test_utils.scripts.test_tp_adoption_gpu"]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "TP ADOPTION GPU CHECK PASSED" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]
