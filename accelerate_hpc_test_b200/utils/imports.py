"""Optional-dependency and capability probes.

The reference probes ~60 packages and a dozen device vendors (reference src/accelerate/utils/
imports.py); this build targets CUDA on B200 plus CPU/gloo for tests, so vendor probes return
False and the optional packages are the ones a Blackwell training job can actually meet.
"""

from __future__ import annotations

import importlib
import importlib.metadata
import importlib.util
import os
from functools import lru_cache

import torch

from .environment import parse_flag_from_env, str_to_bool
from .versions import compare_versions, is_torch_version


def _is_package_available(pkg_name: str, metadata_name: str | None = None) -> bool:
    if importlib.util.find_spec(pkg_name) is None:
        return False
    try:
        importlib.metadata.metadata(pkg_name if metadata_name is None else metadata_name)
        return True
    except importlib.metadata.PackageNotFoundError:
        return False


def is_torch_distributed_available() -> bool:
    return torch.distributed.is_available()


def is_cuda_available() -> bool:
    """CUDA availability that respects ``ACCELERATE_USE_CPU`` (no CUDA init side effects beyond torch's own)."""
    return torch.cuda.is_available()


def is_b200_kernels_available() -> bool:
    """True when the in-tree sm_100a kernel library can run on this machine."""
    from ..ops import _lib

    return _lib.available()


def is_bf16_available(ignore_tpu: bool = False) -> bool:
    if torch.cuda.is_available():
        return torch.cuda.is_bf16_supported()
    return True


def is_fp16_available() -> bool:
    return True


def is_fp8_available() -> bool:
    """Native MXFP8 path (own kernels) or one of the library backends."""
    return is_b200_kernels_available() or is_transformer_engine_available() or is_torchao_available()


def is_ccl_available() -> bool:
    return False


def is_deepspeed_available() -> bool:
    return _is_package_available("deepspeed")


def is_megatron_lm_available() -> bool:
    return str_to_bool(os.environ.get("ACCELERATE_USE_MEGATRON_LM", "False")) == 1 and _is_package_available("megatron")


def is_transformer_engine_available() -> bool:
    return _is_package_available("transformer_engine", "transformer-engine")


def is_transformer_engine_mxfp8_available() -> bool:
    if not is_transformer_engine_available():
        return False
    try:
        import transformer_engine.pytorch as te

        return bool(te.fp8.check_mxfp8_support()[0])
    except Exception:
        return False


def is_torchao_available() -> bool:
    return _is_package_available("torchao") and compare_versions("torchao", ">=", "0.6.1")


def is_msamp_available() -> bool:
    return _is_package_available("msamp", "ms-amp")


def is_bnb_available(min_version: str | None = None) -> bool:
    ok = _is_package_available("bitsandbytes")
    if ok and min_version is not None:
        return compare_versions("bitsandbytes", ">=", min_version)
    return ok


def is_4bit_bnb_available() -> bool:
    return is_bnb_available("0.39.0")


def is_8bit_bnb_available() -> bool:
    return is_bnb_available("0.37.2")


def is_transformers_available() -> bool:
    return _is_package_available("transformers")


def is_datasets_available() -> bool:
    return _is_package_available("datasets")


def is_peft_available() -> bool:
    return _is_package_available("peft")


def is_timm_available() -> bool:
    return _is_package_available("timm")


def is_safetensors_available() -> bool:
    return _is_package_available("safetensors")


def is_tensorboard_available() -> bool:
    return _is_package_available("tensorboard") or _is_package_available("tensorboardX")


def is_wandb_available() -> bool:
    return _is_package_available("wandb")


def is_comet_ml_available() -> bool:
    return _is_package_available("comet_ml")


def is_aim_available() -> bool:
    return _is_package_available("aim")


def is_mlflow_available() -> bool:
    return _is_package_available("mlflow")


def is_clearml_available() -> bool:
    return _is_package_available("clearml")


def is_dvclive_available() -> bool:
    return _is_package_available("dvclive")


def is_swanlab_available() -> bool:
    return _is_package_available("swanlab")


def is_trackio_available() -> bool:
    return _is_package_available("trackio")


def is_rich_available() -> bool:
    return _is_package_available("rich") and parse_flag_from_env("ACCELERATE_ENABLE_RICH", False)


def is_tqdm_available() -> bool:
    return _is_package_available("tqdm")


def is_pandas_available() -> bool:
    return _is_package_available("pandas")


def is_pynvml_available() -> bool:
    return _is_package_available("pynvml") or _is_package_available("pynvml", "nvidia-ml-py")


def is_pytest_available() -> bool:
    return _is_package_available("pytest")


def is_boto3_available() -> bool:
    return _is_package_available("boto3")


def is_sagemaker_available() -> bool:
    return False  # SageMaker launch is out of scope for a B200 node


def is_torchdata_available() -> bool:
    return _is_package_available("torchdata")


def is_torchdata_stateful_dataloader_available() -> bool:
    return _is_package_available("torchdata") and compare_versions("torchdata", ">=", "0.8.0")


def is_lomo_available() -> bool:
    return _is_package_available("lomo_optim")


def is_schedulefree_available() -> bool:
    return _is_package_available("schedulefree")


def is_matplotlib_available() -> bool:
    return _is_package_available("matplotlib")


def is_import_timer_available() -> bool:
    return _is_package_available("import_timer")


def is_pippy_available() -> bool:
    return is_torch_version(">=", "2.4.0")


def is_weights_only_available() -> bool:
    return is_torch_version(">=", "2.4.0")


def is_torch_xla_available(check_is_tpu: bool = False, check_is_gpu: bool = False) -> bool:
    return False


def is_mps_available(min_version: str = "1.12") -> bool:
    return False


def is_xpu_available(check_device: bool = False) -> bool:
    return False


def is_npu_available(check_device: bool = False) -> bool:
    return False


def is_mlu_available(check_device: bool = False) -> bool:
    return False


def is_musa_available(check_device: bool = False) -> bool:
    return False


def is_sdaa_available(check_device: bool = False) -> bool:
    return False


def is_hpu_available(init_hccl: bool = False) -> bool:
    return False


def is_habana_gaudi1() -> bool:
    return False


def is_xccl_available() -> bool:
    return False


def is_mpi_available() -> bool:
    return torch.distributed.is_available() and torch.distributed.is_mpi_available()


def is_symmetric_memory_available() -> bool:
    """torch.distributed._symmetric_memory is importable (used only for NVLS multicast mappings)."""
    try:
        import torch.distributed._symmetric_memory  # noqa: F401

        return True
    except Exception:
        return False


@lru_cache
def is_notebook() -> bool:
    try:
        from IPython import get_ipython

        return get_ipython() is not None and "IPKernelApp" in get_ipython().config
    except Exception:
        return False


def torchao_required(func):
    """Decorator: raise a clear ImportError when torchao is missing (reference utils/imports.py:473-487)."""
    import functools

    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        if not is_torchao_available():
            raise ImportError("`torchao` is not available, please install it before calling this function via `pip install torchao`.")
        return func(*args, **kwargs)

    return wrapper


def deepspeed_required(func):
    """Decorator: the wrapped function only makes sense for a DeepSpeed-configured run (reference utils/imports.py:490-504)."""
    import functools

    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        from ..state import AcceleratorState
        from .dataclasses import DistributedType

        if AcceleratorState._shared_state != {} and AcceleratorState().distributed_type != DistributedType.DEEPSPEED:
            raise ValueError("DeepSpeed is not enabled, please make sure that an `Accelerator` is configured for `deepspeed` before calling this function.")
        return func(*args, **kwargs)

    return wrapper


def is_torchvision_available():
    return _is_package_available("torchvision") if "_is_package_available" in globals() else importlib.util.find_spec("torchvision") is not None


def is_triton_available():
    return importlib.util.find_spec("triton") is not None


def is_bitsandbytes_multi_backend_available():
    if not is_bnb_available():
        return False
    import bitsandbytes as bnb

    return "multi_backend" in getattr(bnb, "features", set())


def is_numpy_available(min_version: str = "1.25.0") -> bool:
    """numpy is a hard dependency; the question is whether it is at least ``min_version``."""
    import importlib.metadata

    from packaging.version import parse

    return parse(importlib.metadata.version("numpy")) >= parse(min_version)
