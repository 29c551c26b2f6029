"""Flash attention fwd / bwd timing at the Llama-3-8B bench shape (B=2, S=2048, 32/8 heads, D=128, causal)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200.ops import functional as F
from benchmarks.bench_gemm import timeit

B, S, Hq, Hkv, D = int(os.environ.get("B", 2)), int(os.environ.get("S", 2048)), 32, 8, 128
qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device="cuda") * 0.5).to(torch.bfloat16)
scale = D ** -0.5
o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale)
do = torch.randn_like(o)
fl_fwd = 4.0 * B * Hq * S * S * D / 2
t = timeit(lambda: F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale))
print(f"flash fwd: {t*1e3:.0f} us  {fl_fwd / t / 1e9:.0f} TFLOP/s")
t = timeit(lambda: F.flash_attn_bwd(do, qkv, o, lse, B, S, Hq, Hkv, D, True, scale))
print(f"flash bwd (prep + main + finalize + zero): {t*1e3:.0f} us  {2.5 * fl_fwd / t / 1e9:.0f} TFLOP/s")
