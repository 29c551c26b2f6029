"""AdamW kernel bandwidth: 28 bytes of HBM traffic per element (fp32 master/m/v read+write, bf16 grad read, bf16 param write)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200.ops import functional as F
from benchmarks.bench_gemm import timeit

n = 4096 * 14336 * 4
master = torch.randn(n, device="cuda")
m = torch.zeros(n, device="cuda")
v = torch.zeros(n, device="cuda")
g = torch.randn(n, device="cuda").to(torch.bfloat16)
sh = torch.empty(n, device="cuda", dtype=torch.bfloat16)
t = timeit(lambda: F.adamw_flat_(master, m, v, g, sh, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=3))
print(f"adamw {n/1e6:.0f}M elems: {t:.3f} ms, {28 * n / t / 1e9:.2f} TB/s")
out2 = torch.zeros(2, device="cuda")
t = timeit(lambda: F.sumsq_(g, out2))
print(f"sumsq: {t:.3f} ms, {2 * n / t / 1e9:.2f} TB/s")
