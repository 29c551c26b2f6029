"""ctypes binding of ``lib/libb200.so`` (the sm_100a kernel library, C ABI in ``include/b200_ops.h``).

The library is loaded lazily on first use.  On a machine with a GPU a missing library is a hard
error (the CUDA path must never silently fall back to eager PyTorch); on a CPU-only machine the
ops in :mod:`.functional` run their plain-PyTorch restatement instead, which is also the oracle
the GPU tests compare against.
"""

from __future__ import annotations

import ctypes
import os
import threading
from ctypes import c_float, c_int, c_longlong, c_void_p

import torch

from .._build import LIB_PATH, build

_lock = threading.Lock()
_lib = None

P = c_void_p
I = c_int
L = c_longlong
F = c_float
U = ctypes.c_uint

# name -> argtypes (restype is always int unless listed in _RESTYPES)
_SIGNATURES = {
    "b200_gemm_bf16": [P, P, P, P, I, I, I, L, L, L, I, I, I, I, I, P],
    "b200_seq_head_exchange": [P, I, I, L, P, I, I, I, L, L, P, I, I, I, U, P, I, P],
    "b200_gemm2_mxfp8": [P, P, P, P, P, P, I, I, I, L, L, L, I, I, P],
    "b200_gemm_mxfp8": [P, P, P, P, P, P, I, I, I, L, L, L, I, I, P],
    "b200_quantize_mx_rowwise": [P, L, I, I, P, L, P, P],
    "b200_quantize_mx_colwise": [P, L, I, I, P, L, P, P],
    "b200_gemm2_e4m3": [P, P, P, P, P, P, I, I, I, L, L, L, I, I, P],
    "b200_amax_bf16": [P, L, P, P],
    "b200_quantize_e4m3": [P, L, I, I, P, P, L, P, L, P, P],
    "b200_gemm2_bf16": [P, P, P, P, I, I, I, L, L, L, I, I, I, I, P],
    "b200_gemm2_bf16_allgather_a": [P, I, I, L, P, I, P, P, I, I, L, L, I, P, ctypes.c_ulonglong, P],
    "b200_gemm2_bf16_gather": [P, P, P, P, I, I, I, L, L, L, I, I, I, I, P, I, I, L, P, L, I, I, P, ctypes.c_ulonglong, P],
    "b200_gemm2_bf16_ulysses_scatter": [P, P, P, I, I, L, I, I, P, I, I, L, L, I, P],
    "b200_gemm2_bf16_scatter": [P, P, P, I, I, L, I, I, I, L, L, I, I, P],
    "b200_gemm_bf16_allgather_a": [P, I, I, L, P, I, P, P, I, I, L, L, I, P, ctypes.c_ulonglong, P],
    "b200_gemm_bf16_scatter": [P, P, P, I, I, L, I, I, I, L, L, I, I, P],
    "b200_reduce_slots": [P, P, I, L, P],
    "b200_rmsnorm_fwd": [P, P, P, P, P, P, L, I, F, P],
    "b200_rmsnorm_bwd_parts": [L],
    "b200_rmsnorm_bwd": [P, P, P, P, P, P, P, P, L, I, I, P],
    "b200_rmsnorm_bwd_noalloc": [P, P, P, P, P, P, P, L, I, I, P],
    "b200_rope_inplace": [P, P, P, P, L, L, I, I, I, P],
    "b200_swiglu_fwd": [P, P, L, I, P],
    "b200_swiglu_bwd": [P, P, P, L, I, P],
    "b200_embedding_fwd": [P, P, P, L, I, L, P],
    "b200_embedding_bwd": [P, P, P, L, I, L, P],
    "b200_scale_cast": [P, P, L, F, I, P],
    "b200_add_inplace_bf16": [P, P, L, P],
    "b200_scale_by_scalar_bf16": [P, P, L, P, I, P],
    "b200_adamw_flat": [P, P, P, P, P, L, I, F, F, F, F, F, I, F, P, P, P],
    "b200_adamw_multi": [P, I, L, I, F, F, F, F, F, I, F, P, P, P],
    "b200_sumsq": [P, L, I, P, P],
    "b200_clip_coef": [P, F, P, P, P],
    "b200_ce_fwd_bwd": [P, P, P, P, L, L, I, P, F, L, I, P],
    "b200_count_valid": [P, L, L, P, P, P],
    "b200_loss_reduce": [P, L, P, F, P, I, P],
    "b200_flash_attn_fwd": [P, P, P, P, P, I, I, I, I, I, L, L, L, L, F, I, P],
    "b200_flash_attn_bwd_prep": [P, P, P, I, I, I, L, L, P],
    "b200_flash_attn_bwd": [P, P, P, P, P, P, P, P, P, I, I, I, I, I, L, L, L, L, L, F, I, P],
    "b200_dq_finalize": [P, P, I, I, I, L, F, P],
    "b200_launch_count": [],
    "b200_reset_launch_count": [],
    "b200_version": [],
    "b200_device_info": [P, P, P, P],
    "b200_error_string": [I],
    "b200_symm_alloc": [P, L],
    "b200_symm_free": [P],
    "b200_ipc_get_handle": [P, P],
    "b200_ipc_open_handle": [P, P],
    "b200_ipc_close_handle": [P],
    "b200_memset_async": [P, I, L, P],
    "b200_memcpy_async": [P, P, L, P],
    "b200_comm_limits": [P, P, P],
    "b200_comm_set_timeout_ms": [L],
    "b200_barrier_all": [P, I, I, I, U, P],
    "b200_allgather_pull": [P, I, I, L, L, P, I, U, I, P],
    "b200_allgather_pull_segs": [P, I, I, L, P, I, P, I, U, I, P],
    "b200_reduce_scatter_adamw_segs": [P, I, I, I, L, P, I, P, P, P, L, P, F, F, F, F, F, I, F, I, P, I, I, U, P, I, P],
    "b200_reduce_scatter_adamw_nvls": [P, P, I, I, I, L, P, I, P, P, P, L, P, F, F, F, F, F, I, F, I, P, I, I, U, P, I, P],
    "b200_allreduce_nvls": [P, P, I, I, L, L, F, I, I, U, P, I, P],
    "b200_reduce_scatter_adamw": [P, I, I, L, L, P, P, P, L, P, F, F, F, F, F, I, F, I, P, I, I, U, P, I, P],
    "b200_allreduce": [P, I, I, L, L, I, F, I, I, U, P, I, P],
    "b200_allreduce_tma": [P, I, I, L, L, F, I, I, I, U, P, I, P],
    "b200_alltoall_pull": [P, I, I, L, L, P, I, I, U, P, I, P],
}
_RESTYPES = {
    "b200_launch_count": ctypes.c_ulonglong,
    "b200_reset_launch_count": None,
    "b200_error_string": ctypes.c_char_p,
}


class B200KernelError(RuntimeError):
    pass


def lib_path() -> str:
    return str(LIB_PATH)


def available() -> bool:
    """True when the kernel library can be used in this process (built and a CUDA device exists)."""
    return torch.cuda.is_available() and (LIB_PATH.exists() or _try_build())


def _try_build() -> bool:
    try:
        build()
        return True
    except Exception:
        return False


def get_lib():
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not LIB_PATH.exists():
            build()
        lib = ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)
        for name, argtypes in _SIGNATURES.items():
            try:
                fn = getattr(lib, name)
            except AttributeError:
                continue  # entry point not built (optional component)
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, c_int)
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = get_lib().b200_error_string(rc)
        msg = msg.decode() if isinstance(msg, bytes) else str(msg)
        raise B200KernelError(f"{what or 'b200 kernel'} failed with code {rc}: {msg}")


_cuda_checked = False


def require_gpu_lib():
    """Return the library or raise loudly: used by ops when handed CUDA tensors."""
    global _cuda_checked
    if not _cuda_checked:  # once: is_available() re-reads the environment on every call and this sits on every op's path
        if not torch.cuda.is_available():
            raise B200KernelError("CUDA tensors passed to a b200 op but no CUDA device is available")
        _cuda_checked = True
    return _lib if _lib is not None else get_lib()


def ptr(t) -> c_void_p:
    return c_void_p(0) if t is None else c_void_p(t.data_ptr())


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def cur_stream() -> c_void_p:
    """Handle of torch's current stream on the current device.  Called once per kernel launch: the raw accessor skips the
    ``torch.cuda.Stream`` object (and the device-index / availability checks) that ``torch.cuda.current_stream()`` builds."""
    if _raw_stream is not None:
        return c_void_p(_raw_stream(torch._C._cuda_getDevice()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count() -> int:
    if _lib is None:
        return 0
    return int(_lib.b200_launch_count())


def reset_launch_count() -> None:
    if _lib is not None:
        _lib.b200_reset_launch_count()


def loaded() -> bool:
    return _lib is not None


if os.environ.get("B200_EAGER_LOAD", "0") == "1" and torch.cuda.is_available():  # pragma: no cover
    get_lib()
