"""Small single-GPU program that launches each hot kernel a few times (the target of ``ncu --set full``)."""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200.ops import functional as F

T, H, I = 4096, 4096, 14336
dev = "cuda"
x = torch.randn(T, H, device=dev, dtype=torch.bfloat16)
w = torch.randn(2 * I, H, device=dev, dtype=torch.bfloat16) * 0.02
dy = torch.randn(T, 2 * I, device=dev, dtype=torch.bfloat16)
B, S, Hq, Hkv, D = 2, 2048, 32, 8, 128
qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device=dev) * 0.5).to(torch.bfloat16)
for _ in range(3):
    y = F.gemm(x, w)                                        # fwd  gate_up
    dx = F.gemm(dy, w, b_mn_major=True)                     # dgrad
    dw = F.gemm(dy, x, a_mn_major=True, b_mn_major=True)    # wgrad
    o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, 1 / math.sqrt(D))
    dqkv = F.flash_attn_bwd(torch.randn_like(o), qkv, o, lse, B, S, Hq, Hkv, D, True, 1 / math.sqrt(D))
    act = F.swiglu_fwd(y)
# optimizer + fp8 path
n = 4096 * 14336
master = torch.randn(n, device=dev); m1 = torch.zeros(n, device=dev); v1 = torch.zeros(n, device=dev)
g = torch.randn(n, device=dev).to(torch.bfloat16); sh = torch.empty(n, device=dev, dtype=torch.bfloat16)
from accelerate_hpc_test_b200.parallel.precision import gemm_e4m3, quantize_e4m3
for _ in range(2):
    F.adamw_flat_(master, m1, v1, g, sh, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=3)
    xq, _, dxs = quantize_e4m3(x)
    wq, wqt, dws = quantize_e4m3(w, transpose=True)
    y8 = gemm_e4m3(xq, wq, dxs, dws)
torch.cuda.synchronize()
print("ok")
