"""What the process can learn from its environment: flags in ``os.environ``, the GPUs and driver underneath, MPI rank variables, the
NUMA node a GPU hangs off — plus the context managers tests and launchers use to edit ``os.environ`` and put it back.

Behavioural parity: reference ``src/accelerate/utils/environment.py`` (flag parsing :59-91, driver P2P guard :197-222, MPI rank discovery
:263-280, NUMA affinity :283-338, ``clear_environment`` / ``patch_environment`` / ``purge_accelerate_environment`` :341-471).  Written for
CUDA nodes only: the reference's per-vendor ladders have no counterpart here.
"""

from __future__ import annotations

import functools
import logging
import os
import platform
import shutil
import subprocess
import sys
from contextlib import contextmanager
from dataclasses import dataclass
from typing import Iterable, Optional

import torch
from packaging.version import parse

logger = logging.getLogger(__name__)

_TRUTH = {**dict.fromkeys(("1", "true", "t", "yes", "y", "on"), True), **dict.fromkeys(("0", "false", "f", "no", "n", "off"), False)}
_ENV_FILE_BREAKERS = frozenset(";\n<> ")


# ------------------------------------------------------------------------------------------------- flags
def str_to_bool(value, to_bool: bool = False):
    """The usual spellings of yes and no (case-insensitive) as 1 / 0 — or True / False with ``to_bool``; anything else raises."""
    try:
        verdict = _TRUTH[str(value).lower()]
    except KeyError:
        raise ValueError(f"invalid truth value {value}") from None
    return verdict if to_bool else int(verdict)


def parse_flag_from_env(key, default=False) -> bool:
    return bool(str_to_bool(os.environ.get(key, str(default))))


def parse_choice_from_env(key, default="no"):
    return os.environ.get(key, str(default))


def get_int_from_env(env_keys: Iterable[str], default):
    """The first of ``env_keys`` that holds a non-negative integer, else ``default``."""
    for key in env_keys:
        found = int(os.environ.get(key, -1))
        if found >= 0:
            return found
    return default


def convert_dict_to_env_variables(current_env: dict) -> list[str]:
    """``KEY=value`` lines for an env file; pairs that are empty or contain characters a shell would choke on are skipped (and logged)."""
    lines = []
    for key, value in current_env.items():
        if key and value and not _ENV_FILE_BREAKERS & set(key + value):
            lines.append(f"{key}={value}\n")
        else:
            logger.warning(f"WARNING: Skipping {key}={value} as it contains forbidden characters or missing values.")
    return lines


def are_libraries_initialized(*library_names: str) -> list[str]:
    """Which of the named libraries have been imported already (some initialise CUDA on import, which forbids forking)."""
    return [name for name in library_names if name in sys.modules]


# ------------------------------------------------------------------------------------------------- GPUs underneath
def _nvidia_smi() -> str:
    if platform.system() != "Windows":
        return "nvidia-smi"
    return shutil.which("nvidia-smi") or f"{os.environ['systemdrive']}\\Program Files\\NVIDIA Corporation\\NVSMI\\nvidia-smi.exe"


def _query_gpus(fields: str) -> list[list[str]]:
    """One list of fields per GPU, from ``nvidia-smi`` — which, unlike ``torch.cuda``, does not create a CUDA context."""
    text = subprocess.check_output([_nvidia_smi(), f"--query-gpu={fields}", "--format=csv,noheader"], universal_newlines=True)
    return [[cell.strip() for cell in line.split(",")] for line in text.strip().splitlines()]


def get_gpu_info():
    """``([name of every GPU], how many)``."""
    rows = _query_gpus("count,name")
    return [row[1] for row in rows], len(rows)


def get_driver_version() -> str:
    return _query_gpus("driver_version")[0][0]


def check_cuda_p2p_ib_support() -> bool:
    """False only for multi-GPU RTX 40 boxes on drivers older than 550.40.07, whose peer-to-peer / InfiniBand paths hang; the caller
    then disables them for NCCL.  Any probing failure counts as "supported"."""
    try:
        names, count = get_gpu_info()
        if count > 1 and any("RTX 40" in name for name in names):
            return parse(get_driver_version()) >= parse("550.40.07")
    except Exception:
        pass
    return True


def _compute_capability() -> tuple:
    try:
        return tuple(torch.cuda.get_device_capability())
    except Exception:
        return (0, 0)


@functools.lru_cache
def check_cuda_fp8_capability() -> bool:
    """fp8 tensor cores exist from compute capability 8.9 on (B200 is 10.0)."""
    return _compute_capability() >= (8, 9)


@functools.lru_cache
def check_blackwell() -> bool:
    """tcgen05 tensor cores with tensor memory and block-scaled MXFP8 operands: sm_100 and later."""
    return _compute_capability() >= (10, 0)


def get_current_device_type():
    """``(device type, distributed type)`` this process would use — reference utils/environment.py:101-130 minus the other vendors."""
    return ("cuda", "MULTI_GPU") if torch.cuda.is_available() else ("cpu", "MULTI_CPU")


# ------------------------------------------------------------------------------------------------- MPI launches
@dataclass
class CPUInformation:
    """Who this process is in a job started by ``mpirun`` rather than ``torchrun``."""

    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    local_world_size: int = 1


# torchrun's name first, then Intel MPI / MPICH, Open MPI, MVAPICH2
_MPI_VARIABLES = {
    "rank": ("RANK", "PMI_RANK", "OMPI_COMM_WORLD_RANK", "MV2_COMM_WORLD_RANK"),
    "world_size": ("WORLD_SIZE", "PMI_SIZE", "OMPI_COMM_WORLD_SIZE", "MV2_COMM_WORLD_SIZE"),
    "local_rank": ("LOCAL_RANK", "MPI_LOCALRANKID", "OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK"),
    "local_world_size": ("LOCAL_WORLD_SIZE", "MPI_LOCALNRANKS", "OMPI_COMM_WORLD_LOCAL_SIZE", "MV2_COMM_WORLD_LOCAL_SIZE"),
}


def get_cpu_distributed_information() -> CPUInformation:
    fallback = CPUInformation()
    return CPUInformation(**{name: get_int_from_env(keys, getattr(fallback, name)) for name, keys in _MPI_VARIABLES.items()})


# ------------------------------------------------------------------------------------------------- NUMA
def override_numa_affinity(local_process_index: int, verbose: Optional[bool] = None) -> None:
    """Restrict this process to the CPU cores NVML reports as closest to GPU ``local_process_index``, so that pinned-memory copies
    and kernel launches do not cross the socket interconnect.  A no-op without CUDA or without ``pynvml``."""
    if not torch.cuda.is_available():
        return
    try:
        import pynvml
    except ImportError:
        return
    if verbose is None:
        verbose = parse_flag_from_env("ACCELERATE_DEBUG_MODE", False)
    pynvml.nvmlInit()
    words = -(-os.cpu_count() // 64)  # NVML hands the mask back in 64-bit words, least significant first
    mask = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_process_index), words)
    cores = [64 * word + bit for word, bits in enumerate(mask) for bit in range(64) if bits >> bit & 1]
    os.sched_setaffinity(0, cores)
    if verbose:
        now = os.sched_getaffinity(0)
        logger.info(f"Assigning {len(now)} cpu cores to process {local_process_index}: {now}")


@functools.lru_cache
def set_numa_affinity(local_process_index: int, verbose: Optional[bool] = None) -> None:
    """``override_numa_affinity`` once per (process, GPU)."""
    override_numa_affinity(local_process_index=local_process_index, verbose=verbose)


# ------------------------------------------------------------------------------------------------- editing os.environ and putting it back
@contextmanager
def _environment_restored(only_prefix: str = ""):
    """On exit every variable starting with ``only_prefix`` has the value it had on entry (variables born inside are removed)."""
    before = {key: value for key, value in os.environ.items() if key.startswith(only_prefix)}
    try:
        yield
    finally:
        for key in [key for key in os.environ if key.startswith(only_prefix) and key not in before]:
            del os.environ[key]
        os.environ.update(before)


@contextmanager
def clear_environment():
    """An empty ``os.environ`` (the same mapping object, emptied) for the duration of the block."""
    with _environment_restored():
        os.environ.clear()
        yield


@contextmanager
def patch_environment(**kwargs):
    """Set the given variables (names upper-cased, values stringified) for the duration of the block; other variables changed inside
    the block keep their new values."""
    names = [name.upper() for name in kwargs]
    before = {name: os.environ.get(name) for name in names}
    os.environ.update({name: str(value) for name, value in zip(names, kwargs.values())})
    try:
        yield
    finally:
        for name, old in before.items():
            if old is None:
                os.environ.pop(name, None)
            else:
                os.environ[name] = old


def purge_accelerate_environment(func_or_cls):
    """Decorator for a test function, or a test class (its ``test*`` methods): ``ACCELERATE_*`` variables set while it runs do not
    leak into the tests after it."""
    marker = "_accelerate_is_purged_environment_wrapped"

    def guarded(func):
        if getattr(func, marker, False):
            return func

        @functools.wraps(func)
        def run(*args, **kwargs):
            with _environment_restored("ACCELERATE_"):
                return func(*args, **kwargs)

        setattr(run, marker, True)
        return run

    if not isinstance(func_or_cls, type):
        return guarded(func_or_cls)
    for name in dir(func_or_cls):
        member = getattr(func_or_cls, name, None)
        if name.startswith("test") and callable(member):
            setattr(func_or_cls, name, guarded(member))
    return func_or_cls
