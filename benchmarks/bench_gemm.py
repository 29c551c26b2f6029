"""Micro-benchmark: tcgen05 GEMM vs cuBLASLt (torch.matmul) on the Llama-3-8B shapes. CUDA-event timed."""
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200.ops import functional as F


class Ms(float):
    """Milliseconds that remember the clock record of the run they came from."""

    def __new__(cls, value, clocks=None):
        self = super().__new__(cls, value)
        self.clocks = clocks
        return self


_PROBE = None
LAST_CLOCKS = None  # clocks / throttle reasons sampled while the kernels of the last timeit() call were in flight


def timeit(fn, iters=10, warm=3):
    """Median ms over ``iters`` calls: CUDA events on the launching stream, L2 flushed (256 MiB memset) before every call, NVML
    clock samples taken while each call is in flight (``LAST_CLOCKS``)."""
    global _PROBE, LAST_CLOCKS
    from accelerate_hpc_test_b200.utils.diagnostics import ClockProbe

    if _PROBE is None:
        _PROBE = ClockProbe()
    _PROBE.reset()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        _PROBE.sample()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    LAST_CLOCKS = _PROBE.summary()
    return Ms(ts[len(ts) // 2], LAST_CLOCKS)


def main():
    T = int(os.environ.get("TOKENS", 8192))
    shapes = [("qkv", T, 6144, 4096), ("o", T, 4096, 4096), ("gate_up", T, 28672, 4096), ("down", T, 4096, 14336)]
    rows = []
    for name, M, N, K in shapes:
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
        dy = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        for kind, ours, ref in [
            ("fwd", lambda: F.gemm(x, w), lambda: x @ w.t()),
            ("dgrad", lambda: F.gemm(dy, w, b_mn_major=True), lambda: dy @ w),
            ("wgrad", lambda: F.gemm(dy, x, a_mn_major=True, b_mn_major=True), lambda: dy.t() @ x),
        ]:
            t_r = timeit(ref)
            t_o = timeit(ours)
            rows.append(dict(shape=name, kind=kind, M=M, N=N, K=K, ours_ms=float(t_o), cublas_ms=float(t_r),
                             ours_tflops=fl / t_o / 1e9, cublas_tflops=fl / t_r / 1e9, clocks=t_o.clocks, cublas_clocks=t_r.clocks))
            print(json.dumps(rows[-1]), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/bench_gemm.json", "w"), indent=1)


if __name__ == "__main__":
    main()
