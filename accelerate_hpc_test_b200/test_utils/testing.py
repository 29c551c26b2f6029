"""Helpers for writing tests against the framework: skip decorators, state-resetting test cases, subprocess runners
(reference src/accelerate/test_utils/testing.py).  Decorators for accelerators this build does not support (XPU, NPU,
MLU, MUSA, SDAA, HPU, MPS, TPU) always skip."""

from __future__ import annotations

import asyncio
import inspect
import io
import os
import shutil
import subprocess
import sys
import tempfile
import unittest
from contextlib import redirect_stdout
from functools import partial
from pathlib import Path
from typing import Optional, Union
from unittest import mock

import torch

from ..state import AcceleratorState, GradientState, PartialState
from ..utils import imports as _imp
from ..utils.versions import is_torch_version


def get_backend():
    """``(device_type, device_count, memory_allocated_fn)`` of the accelerator under test."""
    if torch.cuda.is_available():
        return "cuda", torch.cuda.device_count(), torch.cuda.memory_allocated
    return "cpu", 1, lambda: 0


torch_device, device_count, memory_allocated_func = get_backend()
DEFAULT_LAUNCH_COMMAND = [sys.executable, "-m", "accelerate_hpc_test_b200.commands.launch"]


def get_launch_command(**kwargs) -> list:
    """``accelerate launch`` prefix with ``--key value`` (or bare ``--flag`` for ``True``) arguments."""
    cmd = list(DEFAULT_LAUNCH_COMMAND)
    for k, v in kwargs.items():
        if isinstance(v, bool):
            if v:
                cmd.append(f"--{k}")
        elif v is not None:
            cmd.append(f"--{k}={v}")
    return cmd


def parse_flag_from_env(key, default=False):
    value = os.environ.get(key)
    if value is None:
        return default
    v = value.strip().lower()
    if v in ("1", "true", "yes", "y", "on", "t"):
        return True
    if v in ("0", "false", "no", "n", "off", "f", ""):
        return False
    raise ValueError(f"If set, {key} must be yes or no.")


_run_slow_tests = parse_flag_from_env("RUN_SLOW", default=False)


def skip(test_case):
    return unittest.skip("Test was skipped")(test_case)


def slow(test_case):
    return unittest.skipUnless(_run_slow_tests, "test is slow")(test_case)


def _require(cond, reason):
    def deco(test_case):
        return unittest.skipUnless(bool(cond() if callable(cond) else cond), reason)(test_case)

    return deco


def _never(reason):
    return _require(False, reason)


require_cpu = _require(lambda: torch_device == "cpu", "test requires only a CPU")
require_non_cpu = _require(lambda: torch_device != "cpu", "test requires a GPU")
require_cuda = _require(lambda: torch.cuda.is_available(), "test requires a GPU")
require_cuda_or_hpu = require_cuda
require_cuda_or_xpu = require_cuda
require_xpu = _never("test requires an XPU (not supported by this build)")
require_non_xpu = _require(True, "")
require_non_hpu = _require(True, "")
require_mlu = _never("test requires MLU (not supported by this build)")
require_sdaa = _never("test requires SDAA (not supported by this build)")
require_musa = _never("test requires MUSA (not supported by this build)")
require_npu = _never("test requires NPU (not supported by this build)")
require_mps = _never("test requires MPS (not supported by this build)")
require_tpu = _never("test requires TPU (not supported by this build)")
require_non_torch_xla = _require(True, "")
require_fp16 = _require(lambda: torch.cuda.is_available(), "test requires fp16 support")
require_fp8 = _require(lambda: torch.cuda.is_available() and torch.cuda.get_device_capability() >= (8, 9), "test requires fp8 support")
require_fsdp2 = _require(lambda: is_torch_version(">=", "2.5.0"), "test requires FSDP2")
require_transformers = _require(lambda: _imp.is_transformers_available(), "test requires the transformers library")
require_huggingface_suite = _require(lambda: _imp.is_transformers_available() and _imp.is_datasets_available(), "test requires the Hugging Face suite")
require_timm = _require(lambda: _imp.is_timm_available(), "test requires the timm library")
require_torchvision = _require(lambda: _imp.is_torchvision_available(), "test requires the torchvision library")
require_triton = _require(lambda: _imp.is_triton_available(), "test requires the triton library")
require_schedulefree = _require(lambda: _imp.is_schedulefree_available(), "test requires the schedulefree library")
require_bnb = _require(lambda: _imp.is_bnb_available(), "test requires the bitsandbytes library")
require_single_device = _require(lambda: torch_device != "cpu" and device_count == 1, "test requires a single device accelerator")
require_single_gpu = _require(lambda: torch.cuda.device_count() == 1, "test requires a GPU")
require_single_xpu = _never("test requires an XPU")
require_multi_device = _require(lambda: torch_device != "cpu" and device_count > 1, "test requires multiple hardware accelerators")
require_multi_gpu = _require(lambda: torch.cuda.device_count() > 1, "test requires multiple GPUs")
require_multi_xpu = _never("test requires multiple XPUs")
require_multi_gpu_or_xpu = require_multi_gpu
require_deepspeed = _require(lambda: _imp.is_deepspeed_available(), "test requires DeepSpeed")
require_tp = _require(lambda: is_torch_version(">=", "2.3.0"), "test requires torch tensor parallelism")
require_tensorboard = _require(lambda: _imp.is_tensorboard_available(), "test requires Tensorboard")
require_wandb = _require(lambda: _imp.is_wandb_available(), "test requires wandb")
require_trackio = _require(lambda: _imp.is_trackio_available(), "test requires trackio")
require_comet_ml = _require(lambda: _imp.is_comet_ml_available(), "test requires comet_ml")
require_aim = _require(lambda: _imp.is_aim_available(), "test requires aim")
require_clearml = _require(lambda: _imp.is_clearml_available(), "test requires clearml")
require_dvclive = _require(lambda: _imp.is_dvclive_available(), "test requires dvclive")
require_swanlab = _require(lambda: _imp.is_swanlab_available(), "test requires swanlab")
require_pandas = _require(lambda: _imp.is_pandas_available(), "test requires pandas")
require_mlflow = _require(lambda: _imp.is_mlflow_available(), "test requires mlflow")
require_pippy = _require(lambda: _imp.is_pippy_available(), "test requires torch.distributed.pipelining")
require_import_timer = _require(lambda: _imp.is_import_timer_available(), "test requires tuna interpreter")
require_transformer_engine = _require(lambda: _imp.is_transformer_engine_available(), "test requires transformers engine")
require_transformer_engine_mxfp8 = _require(lambda: _imp.is_transformer_engine_available() and torch.cuda.is_available()
                                            and torch.cuda.get_device_capability() >= (10, 0), "test requires TE MXFP8 support")
require_torchao = _require(lambda: _imp.is_torchao_available(), "test requires torchao")
require_matplotlib = _require(lambda: _imp.is_matplotlib_available(), "test requires matplotlib")
require_torchdata_stateful_dataloader = _require(lambda: _imp.is_torchdata_stateful_dataloader_available(),
                                                 "test requires torchdata.stateful_dataloader")


def require_torch_min_version(test_case=None, version=None):
    if test_case is None:
        return partial(require_torch_min_version, version=version)
    return unittest.skipUnless(is_torch_version(">=", version), f"test requires torch version >= {version}")(test_case)


def require_trackers(test_case):
    have = any(f() for f in (_imp.is_tensorboard_available, _imp.is_wandb_available)) and not _imp.is_comet_ml_available()
    return unittest.skipUnless(have, "test requires at least one tracker to be available and for `comet_ml` to not be installed")(test_case)


def run_first(test_case):
    """Mark a test to run before the rest of the suite (pytest-order, when installed)."""
    try:
        import pytest

        return pytest.mark.order("first")(test_case)
    except Exception:  # pragma: no cover
        return test_case


class TempDirTestCase(unittest.TestCase):
    """A class-wide temporary directory, emptied before every test unless ``clear_on_setup = False``."""

    clear_on_setup = True

    @classmethod
    def setUpClass(cls):
        cls.tmpdir = Path(tempfile.mkdtemp())

    @classmethod
    def tearDownClass(cls):
        if os.path.exists(cls.tmpdir):
            shutil.rmtree(cls.tmpdir)

    def setUp(self):
        if self.clear_on_setup:
            for path in self.tmpdir.glob("**/*"):
                if path.is_file():
                    path.unlink()
                elif path.is_dir():
                    shutil.rmtree(path)


class AccelerateTestCase(unittest.TestCase):
    """Resets the state singletons after every test."""

    def tearDown(self):
        super().tearDown()
        AcceleratorState._reset_state(True)
        PartialState._reset_state()
        GradientState._reset_state()


class MockingTestCase(unittest.TestCase):
    """``self.add_mocks(mock or [mocks])`` starts them and stops them automatically at the end of the test."""

    def add_mocks(self, mocks: Union[mock.Mock, list]):
        self.mocks = mocks if isinstance(mocks, (tuple, list)) else [mocks]
        for m in self.mocks:
            m.start()
            self.addCleanup(m.stop)


def are_the_same_tensors(tensor):
    """True when ``tensor`` holds the same values on every process."""
    from ..utils.operations import gather

    state = AcceleratorState()
    t = tensor[None].clone().to(state.device)
    gathered = gather(t).cpu()
    t0 = t[0].cpu()
    return all(torch.equal(gathered[i], t0) for i in range(gathered.shape[0]))


class _RunOutput:
    def __init__(self, returncode, stdout, stderr):
        self.returncode, self.stdout, self.stderr = returncode, stdout, stderr


async def _stream(cmd, env, stdin, timeout, quiet, echo):
    if echo:
        print("\nRunning: ", " ".join(cmd))
    proc = await asyncio.create_subprocess_exec(cmd[0], *cmd[1:], stdin=stdin, stdout=asyncio.subprocess.PIPE,
                                                stderr=asyncio.subprocess.PIPE, env=env)
    out, err = [], []

    async def pump(stream, sink, pipe, label):
        while True:
            line = await stream.readline()
            if not line:
                break
            text = line.decode("utf-8").rstrip()
            sink.append(text)
            if not quiet:
                print(label, text, file=pipe)

    await asyncio.wait([asyncio.create_task(pump(proc.stdout, out, sys.stdout, "stdout:")),
                        asyncio.create_task(pump(proc.stderr, err, sys.stderr, "stderr:"))], timeout=timeout)
    return _RunOutput(await proc.wait(), out, err)


def execute_subprocess_async(cmd: list, env=None, stdin=None, timeout=180, quiet=False, echo=True) -> _RunOutput:
    """Run ``cmd`` streaming its output; raise with the captured stderr when it fails."""
    cmd = [str(c) for c in cmd]
    result = asyncio.run(_stream(cmd, env, stdin, timeout, quiet, echo))
    if result.returncode > 0:
        stderr = "\n".join(result.stderr)
        raise RuntimeError(f"'{' '.join(cmd)}' failed with returncode {result.returncode}\n\nThe combined stderr from workers follows:\n{stderr}")
    return result


def pytest_xdist_worker_id():
    worker = os.environ.get("PYTEST_XDIST_WORKER", "gw0")
    return int("".join(ch for ch in worker if ch.isdigit()) or 0)


def get_torch_dist_unique_port():
    """A rendezvous port that differs between pytest-xdist workers."""
    return 29500 + pytest_xdist_worker_id()


class SubprocessCallException(Exception):
    pass


def run_command(command: list, return_stdout=False, env=None):
    """``subprocess.check_output`` with stderr folded into the exception message."""
    command = [str(c) for c in command]
    if env is None:
        env = os.environ.copy()
    try:
        output = subprocess.check_output(command, stderr=subprocess.STDOUT, env=env)
        if return_stdout:
            return output.decode("utf-8") if hasattr(output, "decode") else output
    except subprocess.CalledProcessError as e:
        raise SubprocessCallException(f"Command `{' '.join(command)}` failed with the following error:\n\n{e.output.decode()}") from e


def path_in_accelerate_package(*components: str) -> Path:
    """Path of a resource inside the installed package."""
    import accelerate_hpc_test_b200

    return Path(inspect.getfile(accelerate_hpc_test_b200)).parent.joinpath(*components)


class assert_exception:
    """``with assert_exception(ValueError, "msg"):`` — the block must raise that exception (containing ``msg``)."""

    def __init__(self, exception_class, msg: Optional[str] = None):
        self.exception_class, self.msg = exception_class, msg

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc, tb):
        if exc_type is None:
            raise AssertionError(f"{self.exception_class.__name__} was not raised")
        if not issubclass(exc_type, self.exception_class):
            return False
        if self.msg is not None and self.msg not in str(exc):
            raise AssertionError(f"expected message containing {self.msg!r}, got {str(exc)!r}")
        return True


def capture_call_output(func, *args, **kwargs):
    """stdout printed by ``func(*args, **kwargs)``."""
    buf = io.StringIO()
    with redirect_stdout(buf):
        func(*args, **kwargs)
    return buf.getvalue()
