"""Micro-benchmark: fp8 (e4m3, kind::f8f6f4) 2-CTA GEMM vs the bf16 one on Llama-3-8B shapes, plus the quantise kernels."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200.ops import functional as F
from accelerate_hpc_test_b200.parallel.precision import gemm_e4m3, quantize_e4m3
from benchmarks.bench_gemm import timeit


def main():
    T = 8192
    for name, M, N, K in [("qkv", T, 6144, 4096), ("gate_up", T, 28672, 4096), ("down", T, 4096, 14336)]:
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
        xq, _, dx = quantize_e4m3(x)
        wq, _, dw = quantize_e4m3(w)
        fl = 2.0 * M * N * K
        t8 = timeit(lambda: gemm_e4m3(xq, wq, dx, dw))
        t16 = timeit(lambda: F.gemm(x, w))
        tq = timeit(lambda: quantize_e4m3(x))
        tqw = timeit(lambda: quantize_e4m3(w, transpose=True))
        print(json.dumps(dict(shape=name, M=M, N=N, K=K, fp8_ms=t8, bf16_ms=t16, fp8_tflops=fl / t8 / 1e9, bf16_tflops=fl / t16 / 1e9,
                              quant_x_ms=tqw and tq, quant_w_t_ms=tqw)), flush=True)


if __name__ == "__main__":
    main()


def main_mx():
    from accelerate_hpc_test_b200.parallel.precision import gemm_mxfp8, quantize_mx

    T = 8192
    for name, M, N, K in [("qkv", T, 6144, 4096), ("gate_up", T, 28672, 4096), ("down", T, 4096, 14336)]:
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
        xq, wq = quantize_mx(x), quantize_mx(w)
        fl = 2.0 * M * N * K
        t8 = timeit(lambda: gemm_mxfp8(xq, wq))
        tq = timeit(lambda: quantize_mx(x))
        tqt = timeit(lambda: quantize_mx(w, transpose=True))
        print(json.dumps(dict(shape=name, kind="mxfp8", M=M, N=N, K=K, ms=t8, tflops=fl / t8 / 1e9, quant_x_ms=tq, quant_w_t_ms=tqt)), flush=True)


if __name__ == "__main__":
    if os.environ.get("MX", "1") == "1":
        main_mx()
