"""The allocation-free RMSNorm backward, called straight through the C ABI.

Some hosts may not allocate device memory at the point where they need this kernel: the device heap
is owned by another allocator with a fixed budget, or the call comes from inside that allocator's own
callback, where a nested cudaMalloc / tensor allocation deadlocks or is refused.
``b200_rmsnorm_bwd_noalloc`` therefore keeps its scratch in a fixed array inside the library.  This
test drives it with ctypes (no Python wrapper in between), twice in a row on different inputs, and
compares with the regular entry point, which takes its scratch from the caller, and with the fp32
formula of the op (``ops/functional.py`` CPU body).  That nothing was allocated is checked on the driver's
own counter (``cudaMemGetInfo`` through ``torch.cuda.mem_get_info``), which also sees a ``cudaMalloc`` made
inside the library — ``torch.cuda.memory_allocated`` would not.  The file name sorts last on purpose: a
`-x` run has been through every other test before it gets here.
"""

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from accelerate_hpc_test_b200._build import LIB_PATH, build

    if not LIB_PATH.exists():
        build()
    lib = ctypes.CDLL(str(LIB_PATH))
    P, L, I = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
    lib.b200_rmsnorm_bwd_noalloc.argtypes = [P, P, P, P, P, P, P, L, I, I, P]
    lib.b200_rmsnorm_bwd_noalloc.restype = I
    lib.b200_rmsnorm_bwd.argtypes = [P, P, P, P, P, P, P, P, L, I, I, P]
    lib.b200_rmsnorm_bwd.restype = I
    lib.b200_rmsnorm_bwd_parts.argtypes = [L]
    lib.b200_rmsnorm_bwd_parts.restype = I
    return lib


def _p(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp(min=1e-12)).item()


def _fp32_formula(dy, x, w, rstd, dres):
    """dx = rstd * (g - xh * mean(g * xh)) (+ dres),  dw = sum_t dy * xh,  with g = dy * w and xh = x * rstd."""
    xh = x.float() * rstd[:, None]
    g = dy.float() * w.float()
    dx = rstd[:, None] * (g - xh * (g * xh).mean(-1, keepdim=True))
    if dres is not None:
        dx = dx + dres.float()
    return dx, (dy.float() * xh).sum(0)


def test_noalloc_rmsnorm_bwd_matches_regular_entry_point_twice():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # second call: fewer rows and a narrower hidden size than the first, so anything the first call
    # left behind in the shared array would show up if it were read
    for seed, (T, H, with_res) in enumerate([(1000, 4096, True), (37, 1024, False)]):
        g = torch.Generator(device="cuda").manual_seed(seed)
        dy = torch.randn(T, H, device=dev, generator=g).to(torch.bfloat16)
        x = torch.randn(T, H, device=dev, generator=g).to(torch.bfloat16)
        w = (1 + 0.1 * torch.randn(H, device=dev, generator=g)).to(torch.bfloat16)
        rstd = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5)
        dres = torch.randn(T, H, device=dev, generator=g).to(torch.bfloat16) if with_res else None
        dx_ref, dw_ref = torch.empty_like(x), torch.empty_like(w)
        scratch = torch.empty(lib.b200_rmsnorm_bwd_parts(T), H, device=dev, dtype=torch.float32)
        rc = lib.b200_rmsnorm_bwd(_p(dy), _p(x), _p(w), _p(rstd), _p(dres), _p(dx_ref), _p(dw_ref), _p(scratch), T, H, 0, stream)
        assert rc == 0, rc
        dx, dw = torch.empty_like(x), torch.empty_like(w)
        dx_f, dw_f = _fp32_formula(dy, x, w, rstd, dres)
        if seed == 0:
            # the very first call on a device resolves the array's address (cudaGetSymbolAddress): do it before measuring
            lib.b200_rmsnorm_bwd_noalloc(_p(dy), _p(x), _p(w), _p(rstd), _p(dres), _p(dx), _p(dw), T, H, 0, stream)
        torch.cuda.synchronize()
        before_torch = torch.cuda.memory_allocated()
        free_before, _ = torch.cuda.mem_get_info(dev)
        rc = lib.b200_rmsnorm_bwd_noalloc(_p(dy), _p(x), _p(w), _p(rstd), _p(dres), _p(dx), _p(dw), T, H, 0, stream)
        assert rc == 0, rc
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info(dev)
        assert torch.cuda.memory_allocated() == before_torch
        assert free_after >= free_before, ("device memory was allocated inside the call", free_before - free_after)
        # against the entry point that takes caller scratch (same kernels: tight), and against the fp32 formula
        assert _rel(dx, dx_ref) < 1e-2, ("dx", T, H, _rel(dx, dx_ref))
        assert _rel(dw, dw_ref) < 2e-2, ("dw", T, H, _rel(dw, dw_ref))
        assert _rel(dx, dx_f) < 1e-2 and (dx.float() - dx_f).abs().max().item() < 0.1, ("dx vs fp32", T, H, _rel(dx, dx_f))
        assert _rel(dw, dw_f) < 2e-2, ("dw vs fp32", T, H, _rel(dw, dw_f))


def test_noalloc_rejects_sizes_beyond_its_limit():
    lib = _lib()
    t = torch.zeros(8, device="cuda", dtype=torch.bfloat16)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b200_rmsnorm_bwd_noalloc(_p(t), _p(t), _p(t), _p(t), None, _p(t), _p(t), 1, 8192 + 8, 0, stream)
    assert rc == -3  # B200_ERR_TOO_LARGE
