"""Small single-GPU run of every kernel family — the target of
    compute-sanitizer --tool memcheck|racecheck|synccheck python benchmarks/sanitizer_targets.py
(the reference has no sanitizer story; kernels that own shared memory rings, TMEM and mbarriers need one).
Shapes are tiny so that the sanitizer's slowdown stays in seconds; every result is also checked against fp32 PyTorch."""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.ops import functional as F
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.parallel.precision import gemm_e4m3, gemm_mxfp8, quantize_e4m3, quantize_mx

dev = "cuda"
torch.manual_seed(0)


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp(min=1e-6)).item()


M, N, K = 384, 512, 256
x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
dy = torch.randn(M, N, device=dev, dtype=torch.bfloat16)
for two_cta in ("1", "0"):
    F._GEMM_2CTA = two_cta == "1"
    assert rel(F.gemm(x, w), x.float() @ w.float().t()) < 1e-2
    assert rel(F.gemm(dy, w, b_mn_major=True), dy.float() @ w.float()) < 1e-2
    assert rel(F.gemm(dy, x, a_mn_major=True, b_mn_major=True), dy.float().t() @ x.float()) < 1e-2
F._GEMM_2CTA = True
xq, _, dx = quantize_e4m3(x)
wq, wqt, dw = quantize_e4m3(w, transpose=True)
assert rel(gemm_e4m3(xq, wq, dx, dw), x.float() @ w.float().t()) < 0.08
for env in ("1", "0"):
    os.environ["ACCELERATE_B200_GEMM_2CTA"] = env
    assert rel(gemm_mxfp8(quantize_mx(x), quantize_mx(w)), x.float() @ w.float().t()) < 0.08
os.environ["ACCELERATE_B200_GEMM_2CTA"] = "1"
quantize_mx(w.t().contiguous()[:256], transpose=True)

B, S, Hq, Hkv, D = 1, 320, 4, 2, 128
qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device=dev) * 0.5).to(torch.bfloat16)
o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, 1 / math.sqrt(D))
ref = F._attn_reference(qkv, B, S, Hq, Hkv, D, True, 1 / math.sqrt(D))
assert rel(o, ref) < 2e-2
dqkv = F.flash_attn_bwd(torch.randn_like(o), qkv, o, lse, B, S, Hq, Hkv, D, True, 1 / math.sqrt(D))
assert torch.isfinite(dqkv.float()).all()

cfg = LlamaConfig.tiny(num_hidden_layers=2)
model = LlamaForCausalLM(cfg, device=dev)
opt = FusedAdamW(model.parameters(), lr=1e-3)
ids = torch.randint(0, cfg.vocab_size, (2, 128), device=dev)
for _ in range(2):
    loss = model(ids, labels=ids).loss  # embedding, rmsnorm, rope, flash, swiglu, fused linear + cross entropy
    loss.backward()
    opt.step()
    opt.zero_grad()
torch.cuda.synchronize()
assert torch.isfinite(loss)
print("sanitizer targets ok")
