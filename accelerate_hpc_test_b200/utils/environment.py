"""Environment / capability probing (L1 of SURVEY.md §1).

Behavioural parity with reference src/accelerate/utils/environment.py (flag parsing :59-91,
``patch_environment`` / ``clear_environment`` / ``purge_accelerate_environment`` :341-471, MPI rank
discovery :263-280, NUMA affinity :283-338, old-driver P2P guard :197-222) — written for a CUDA-only
B200 node, so the multi-vendor ladders are gone.
"""

from __future__ import annotations

import logging
import math
import os
import platform
import subprocess
import sys
from contextlib import contextmanager
from dataclasses import dataclass, field
from functools import lru_cache, wraps
from shutil import which
from typing import Optional

import torch
from packaging.version import parse

logger = logging.getLogger(__name__)

_TRUE = {"1", "true", "t", "yes", "y", "on"}
_FALSE = {"0", "false", "f", "no", "n", "off"}


def convert_dict_to_env_variables(current_env: dict) -> list[str]:
    """``{"A": "1"}`` -> ``["A=1\\n"]`` skipping entries that would break a shell env file."""
    forbidden = [";", "\n", "<", ">", " "]
    out = []
    for key, value in current_env.items():
        if all(c not in (key + value) for c in forbidden) and len(key) >= 1 and len(value) >= 1:
            out.append(f"{key}={value}\n")
        else:
            logger.warning(f"WARNING: Skipping {key}={value} as it contains forbidden characters or missing values.")
    return out


def str_to_bool(value, to_bool: bool = False):
    """'yes'/'true'/'1'... -> 1 (or True), 'no'/'false'/'0'... -> 0 (or False); anything else raises."""
    v = str(value).lower()
    if v in _TRUE:
        return True if to_bool else 1
    if v in _FALSE:
        return False if to_bool else 0
    raise ValueError(f"invalid truth value {value}")


def get_int_from_env(env_keys, default):
    """First non-negative int found among ``env_keys``, else ``default``."""
    for e in env_keys:
        val = int(os.environ.get(e, -1))
        if val >= 0:
            return val
    return default


def parse_flag_from_env(key, default=False):
    return str_to_bool(os.environ.get(key, str(default))) == 1


def parse_choice_from_env(key, default="no"):
    return os.environ.get(key, str(default))


def are_libraries_initialized(*library_names: str) -> list[str]:
    return [lib for lib in library_names if lib in sys.modules.keys()]


def _nvidia_smi():
    if platform.system() == "Windows":
        command = which("nvidia-smi")
        if command is None:
            command = f"{os.environ['systemdrive']}\\Program Files\\NVIDIA Corporation\\NVSMI\\nvidia-smi.exe"
    else:
        command = "nvidia-smi"
    return command


def get_gpu_info():
    """``(list of GPU names, count)`` without initialising CUDA."""
    output = subprocess.check_output([_nvidia_smi(), "--query-gpu=count,name", "--format=csv,noheader"], universal_newlines=True)
    output = output.strip()
    gpus = output.split(os.linesep)
    gpu_count = len(gpus)
    gpu_names = [gpu.split(",")[1].strip() for gpu in gpus]
    return gpu_names, gpu_count


def get_driver_version():
    output = subprocess.check_output([_nvidia_smi(), "--query-gpu=driver_version", "--format=csv,noheader"], universal_newlines=True)
    return output.strip().split(os.linesep)[0]


def check_cuda_p2p_ib_support() -> bool:
    """False for consumer Ada boards on drivers that mis-handle P2P/IB; always True on B200."""
    try:
        device_names, device_count = get_gpu_info()
        unsupported = {"RTX 40"}
        if device_count > 1 and any(d in name for d in unsupported for name in device_names):
            if parse(get_driver_version()) < parse("550.40.07"):
                return False
    except Exception:
        pass
    return True


@lru_cache
def check_cuda_fp8_capability() -> bool:
    """fp8 tensor cores need compute capability >= 8.9 (B200 is 10.0)."""
    try:
        return torch.cuda.get_device_capability() >= (8, 9)
    except Exception:
        return False


@lru_cache
def check_blackwell() -> bool:
    """True when the current device has tcgen05 / MXFP8 block-scaled tensor cores (sm_100+)."""
    try:
        return torch.cuda.get_device_capability()[0] >= 10
    except Exception:
        return False


@dataclass
class CPUInformation:
    """Rank / world-size facts of an mpirun-launched CPU job, read from the MPI flavours' env vars."""

    rank: int = field(default=0)
    world_size: int = field(default=1)
    local_rank: int = field(default=0)
    local_world_size: int = field(default=1)


def get_cpu_distributed_information() -> CPUInformation:
    info = {
        "rank": get_int_from_env(["RANK", "PMI_RANK", "OMPI_COMM_WORLD_RANK", "MV2_COMM_WORLD_RANK"], 0),
        "world_size": get_int_from_env(["WORLD_SIZE", "PMI_SIZE", "OMPI_COMM_WORLD_SIZE", "MV2_COMM_WORLD_SIZE"], 1),
        "local_rank": get_int_from_env(["LOCAL_RANK", "MPI_LOCALRANKID", "OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK"], 0),
        "local_world_size": get_int_from_env(["LOCAL_WORLD_SIZE", "MPI_LOCALNRANKS", "OMPI_COMM_WORLD_LOCAL_SIZE", "MV2_COMM_WORLD_LOCAL_SIZE"], 1),
    }
    return CPUInformation(**info)


def override_numa_affinity(local_process_index: int, verbose: Optional[bool] = None) -> None:
    """Pin this process to the CPU cores of the NUMA node its GPU hangs off (needs pynvml)."""
    if verbose is None:
        verbose = parse_flag_from_env("ACCELERATE_DEBUG_MODE", False)
    if not torch.cuda.is_available():
        return
    try:
        import pynvml as nvml
    except ImportError:
        return
    nvml.nvmlInit()
    num_elements = math.ceil(os.cpu_count() / 64)
    handle = nvml.nvmlDeviceGetHandleByIndex(local_process_index)
    affinity_string = ""
    for j in nvml.nvmlDeviceGetCpuAffinity(handle, num_elements):
        affinity_string = f"{j:064b}{affinity_string}"
    affinity_list = [int(x) for x in affinity_string]
    affinity_list.reverse()
    affinity_to_set = [i for i, e in enumerate(affinity_list) if e != 0]
    os.sched_setaffinity(0, affinity_to_set)
    if verbose:
        cpu_cores = os.sched_getaffinity(0)
        logger.info(f"Assigning {len(cpu_cores)} cpu cores to process {local_process_index}: {cpu_cores}")


@lru_cache
def set_numa_affinity(local_process_index: int, verbose: Optional[bool] = None) -> None:
    override_numa_affinity(local_process_index=local_process_index, verbose=verbose)


@contextmanager
def clear_environment():
    """Empty ``os.environ`` inside the block and restore it afterwards (same dict object)."""
    _old = os.environ.copy()
    os.environ.clear()
    try:
        yield
    finally:
        os.environ.clear()
        os.environ.update(_old)


@contextmanager
def patch_environment(**kwargs):
    """Temporarily set (upper-cased) env vars; previous values come back on exit."""
    existing = {}
    for key, value in kwargs.items():
        key = key.upper()
        if key in os.environ:
            existing[key] = os.environ[key]
        os.environ[key] = str(value)
    try:
        yield
    finally:
        for key in kwargs:
            key = key.upper()
            if key in existing:
                os.environ[key] = existing[key]
            else:
                os.environ.pop(key, None)


def purge_accelerate_environment(func_or_cls):
    """Decorator (function or TestCase class): ``ACCELERATE_*`` variables set inside do not leak out."""
    prefix = "ACCELERATE_"

    @contextmanager
    def env_var_context():
        existing = {k: v for k, v in os.environ.items() if k.startswith(prefix)}
        try:
            yield
        finally:
            for k in [k for k in os.environ if k.startswith(prefix)]:
                if k in existing:
                    os.environ[k] = existing[k]
                else:
                    os.environ.pop(k, None)
            for k, v in existing.items():
                os.environ[k] = v

    def wrap_function(func):
        @wraps(func)
        def wrapper(*args, **kwargs):
            with env_var_context():
                return func(*args, **kwargs)

        wrapper._accelerate_is_purged_environment_wrapped = True
        return wrapper

    if not isinstance(func_or_cls, type):
        return wrap_function(func_or_cls)
    cls = func_or_cls
    for name in dir(cls):
        attr = getattr(cls, name, None)
        if name.startswith("test") and callable(attr) and not getattr(attr, "_accelerate_is_purged_environment_wrapped", False):
            setattr(cls, name, wrap_function(attr))
    return cls


def get_current_device_type():
    """``(device_type, distributed_type)`` of the accelerator this process would use (reference utils/environment.py:101-130)."""
    import torch

    if torch.cuda.is_available():
        return "cuda", "MULTI_GPU"
    return "cpu", "MULTI_CPU"
