"""Helpers for testing code written against this package: skip decorators keyed on hardware and optional libraries, launch-command
builders, subprocess runners that surface the child's stderr, tiny regression workloads, and the launched scripts ``accelerate test``
runs (``scripts/``).  The names ``from accelerate.test_utils import ...`` offers in the reference are offered here."""

from .testing import (
    DEFAULT_LAUNCH_COMMAND,
    are_the_same_tensors,
    assert_exception,
    capture_call_output,
    device_count,
    execute_subprocess_async,
    get_launch_command,
    get_torch_dist_unique_port,
    memory_allocated_func,
    path_in_accelerate_package,
    pytest_xdist_worker_id,
    require_bnb,
    require_cpu,
    require_cuda,
    require_cuda_or_hpu,
    require_cuda_or_xpu,
    require_fp8,
    require_fp16,
    require_huggingface_suite,
    require_mlu,
    require_mps,
    require_multi_device,
    require_multi_gpu,
    require_multi_gpu_or_xpu,
    require_multi_xpu,
    require_musa,
    require_non_cpu,
    require_non_hpu,
    require_non_torch_xla,
    require_non_xpu,
    require_npu,
    require_pippy,
    require_sdaa,
    require_single_device,
    require_single_gpu,
    require_single_xpu,
    require_torch_min_version,
    require_torchao,
    require_torchvision,
    require_tpu,
    require_transformer_engine,
    require_xpu,
    run_first,
    skip,
    slow,
    torch_device,
)
from .training import RegressionDataset, RegressionModel


def __getattr__(name):
    """``test_utils.test_script`` / ``test_sync`` / ``test_ops`` (the launched scripts) load on first use: importing them with the
    package would put them in ``sys.modules`` before ``python -m ...scripts.test_script`` runs them."""
    if name in ("test_script", "test_sync", "test_ops"):
        import importlib

        return importlib.import_module(f"{__name__}.scripts.{name}")
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
