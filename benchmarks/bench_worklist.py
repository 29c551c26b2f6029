"""SURVEY.md §8 rows W1-W6 on one B200: our kernel vs the library baseline the reference would run, at Llama-3-8B shapes,
with the achieved fraction of the roofline term of each row.  CUDA-event timed, L2 flushed between iterations.
Writes gpurun_out/worklist_single_gpu.json (copied to profiles/)."""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as TF

from accelerate_hpc_test_b200.ops import functional as F
from accelerate_hpc_test_b200.parallel.precision import gemm_mxfp8, quantize_mx
from benchmarks.bench_gemm import timeit

# roofline denominators: the MEASURED numbers of this box (driver-written MEASURED_PEAKS.json), nominal values as fallback
PEAKS = {"bf16_tflops": 2250.0, "fp8_tflops": 4500.0, "hbm_tbs": 7.7, "source": "nominal"}
try:
    mp = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    PEAKS = {"bf16_tflops": float(mp["bf16_tflops"]), "fp8_tflops": 2.0 * float(mp["bf16_tflops"]), "hbm_tbs": float(mp["hbm_gbs"]) / 1000.0,
             "source": "MEASURED_PEAKS.json (cuBLAS bf16 burst, copy bandwidth; fp8 = 2 x bf16)"}
except Exception:
    pass
rows = []
dev = "cuda"
T, H, I, V = 4096, 4096, 14336, 128256


def add(row, name, ours_ms, base_ms, base_name, work, unit, peak):
    achieved = work / ours_ms / 1e9  # TFLOP/s or TB/s
    rows.append(dict(row=row, op=name, ours_ms=round(ours_ms, 4), baseline=base_name, baseline_ms=round(base_ms, 4),
                     speedup_vs_baseline=round(base_ms / ours_ms, 3), achieved=round(achieved, 2), unit=unit,
                     roofline_fraction=round(achieved / peak, 3), clocks=getattr(ours_ms, "clocks", None), baseline_clocks=getattr(base_ms, "clocks", None)))
    print(json.dumps(rows[-1]), flush=True)


# ---- W1: bf16 GEMM (gate_up shape, fwd / dgrad / wgrad)
x = torch.randn(T, H, device=dev, dtype=torch.bfloat16)
w = torch.randn(2 * I, H, device=dev, dtype=torch.bfloat16) * 0.02
dy = torch.randn(T, 2 * I, device=dev, dtype=torch.bfloat16)
fl = 2.0 * T * 2 * I * H
for kind, ours, base in [("fwd", lambda: F.gemm(x, w), lambda: x @ w.t()),
                         ("dgrad", lambda: F.gemm(dy, w, b_mn_major=True), lambda: dy @ w),
                         ("wgrad", lambda: F.gemm(dy, x, a_mn_major=True, b_mn_major=True), lambda: dy.t() @ x)]:
    add("W1", f"bf16 GEMM gate_up {kind}", timeit(ours), timeit(base), "cuBLASLt (torch.matmul)", fl, "TFLOP/s", PEAKS["bf16_tflops"])

# ---- W2: MXFP8 quantize + block-scaled GEMM vs torch._scaled_mm (tensorwise fp8, cuBLASLt)
xq, wq = quantize_mx(x), quantize_mx(w)
t_ours = timeit(lambda: gemm_mxfp8(xq, wq))
try:
    x8, w8 = x.to(torch.float8_e4m3fn), w.to(torch.float8_e4m3fn)
    one = torch.ones((), device=dev)
    t_base = timeit(lambda: torch._scaled_mm(x8, w8.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
    bname = "torch._scaled_mm (cuBLASLt fp8, tensorwise)"
except Exception as e:  # noqa: BLE001
    t_base, bname = float("nan"), f"torch._scaled_mm unavailable: {type(e).__name__}"
add("W2", "MXFP8 block-scaled GEMM gate_up fwd", t_ours, t_base, bname, fl, "TFLOP/s", PEAKS["fp8_tflops"])
tq = timeit(lambda: quantize_mx(x))
add("W2", "MX quantize activations [4096,4096]", tq, float("nan"), "-", x.numel() * 3.03, "TB/s", PEAKS["hbm_tbs"])

# ---- W3: flash attention fwd / fwd+bwd vs SDPA
B, S, Hq, Hkv, D = 2, 2048, 32, 8, 128
qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device=dev) * 0.5).to(torch.bfloat16)
scale = 1 / math.sqrt(D)
o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale)
do = torch.randn_like(o)
q = qkv[:, : Hq * D].view(B, S, Hq, D).transpose(1, 2).contiguous().requires_grad_(True)
k = qkv[:, Hq * D : (Hq + Hkv) * D].view(B, S, Hkv, D).transpose(1, 2).contiguous().requires_grad_(True)
v = qkv[:, (Hq + Hkv) * D :].view(B, S, Hkv, D).transpose(1, 2).contiguous().requires_grad_(True)
dob = do.view(B, S, Hq, D).transpose(1, 2).contiguous()
afl = 4.0 * B * Hq * S * S * D / 2


def sdpa_fwd():
    with torch.no_grad():
        return TF.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True)


def sdpa_fwd_bwd():
    out = TF.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True)
    out.backward(dob)
    q.grad = k.grad = v.grad = None


def ours_fwd_bwd():
    o2, l2 = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale)
    F.flash_attn_bwd(do, qkv, o2, l2, B, S, Hq, Hkv, D, True, scale)


add("W3", "flash attention fwd (B2 S2048 32/8 heads)", timeit(lambda: F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale)), timeit(sdpa_fwd),
    "F.scaled_dot_product_attention", afl, "TFLOP/s", PEAKS["bf16_tflops"])
add("W3", "flash attention fwd+bwd", timeit(ours_fwd_bwd), timeit(sdpa_fwd_bwd), "F.scaled_dot_product_attention + autograd", 3.5 * afl,
    "TFLOP/s", PEAKS["bf16_tflops"])

# ---- W4: RMSNorm / SwiGLU / RoPE vs eager ATen sequences
xs = torch.randn(T, H, device=dev, dtype=torch.bfloat16)
wn = torch.ones(H, device=dev, dtype=torch.bfloat16)


def eager_rms():
    xf = xs.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)).to(torch.bfloat16) * wn


add("W4", "RMSNorm fwd [4096,4096]", timeit(lambda: F.rmsnorm_fwd(xs, wn, 1e-5)), timeit(eager_rms), "eager ATen (transformers LlamaRMSNorm)",
    xs.numel() * 4.0, "TB/s", PEAKS["hbm_tbs"])
gu = torch.randn(T, 2 * I, device=dev, dtype=torch.bfloat16)
add("W4", "SwiGLU fwd [4096,2x14336]", timeit(lambda: F.swiglu_fwd(gu)), timeit(lambda: TF.silu(gu[:, :I]) * gu[:, I:]), "eager ATen silu*mul",
    gu.numel() * 2.0 * 1.5, "TB/s", PEAKS["hbm_tbs"])
cos, sin = F.rope_tables(S, D, 500000.0, dev)
pos = torch.arange(S, device=dev, dtype=torch.int32).repeat(B)
qk = qkv.clone()


def eager_rope():
    t = qk[:, : (Hq + Hkv) * D].view(B * S, Hq + Hkv, D).float()
    c, s_ = cos[pos.long()][:, None, :], sin[pos.long()][:, None, :]
    lo, hi = t[..., : D // 2], t[..., D // 2 :]
    return torch.cat([lo * c - hi * s_, hi * c + lo * s_], -1).to(torch.bfloat16)


add("W4", "RoPE in place on fused QKV", timeit(lambda: F.rope_inplace(qk, cos, sin, pos, Hq + Hkv, D)), timeit(eager_rope), "eager ATen rotate_half",
    B * S * (Hq + Hkv) * D * 4.0, "TB/s", PEAKS["hbm_tbs"])

# ---- W5: fused lm_head + cross-entropy vs ATen linear + cross_entropy (fwd+bwd)
h = (torch.randn(T, H, device=dev) * 0.5).to(torch.bfloat16).requires_grad_(True)
wh = (torch.randn(V, H, device=dev) * 0.02).to(torch.bfloat16).requires_grad_(True)
labels = torch.randint(0, V, (T,), device=dev)


def ours_ce():
    loss = F.fused_linear_cross_entropy(h, wh, labels, -100, 4096)
    loss.backward()
    h.grad = wh.grad = None


def aten_ce():
    loss = TF.cross_entropy((h @ wh.t()).float(), labels)
    loss.backward()
    h.grad = wh.grad = None


add("W5", "lm_head + cross-entropy fwd+bwd (4096 tokens, vocab 128256)", timeit(ours_ce, iters=5), timeit(aten_ce, iters=5),
    "ATen linear + fp32 cross_entropy", 6.0 * T * V * H, "TFLOP/s", PEAKS["bf16_tflops"])

# ---- W6: fused AdamW vs torch.optim.AdamW(fused=True)
n = 4096 * 14336
p_ = torch.randn(n, device=dev)
g_ = torch.randn(n, device=dev).to(torch.bfloat16)
m_, v_ = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
sh = torch.empty(n, device=dev, dtype=torch.bfloat16)
tp = torch.nn.Parameter(torch.randn(n, device=dev))
tp.grad = torch.randn(n, device=dev)
topt = torch.optim.AdamW([tp], lr=1e-3, fused=True)
topt.step()
add("W6", "AdamW, 58.7M params (fp32 master+moments, bf16 grad in, bf16 param out)",
    timeit(lambda: F.adamw_flat_(p_, m_, v_, g_, sh, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=3)),
    timeit(lambda: topt.step()), "torch.optim.AdamW(fused=True) on fp32 params+grads (no bf16 copy)", n * 28.0, "TB/s", PEAKS["hbm_tbs"])

os.makedirs("gpurun_out", exist_ok=True)
json.dump({"peaks": PEAKS, "rows": rows}, open("gpurun_out/worklist_single_gpu.json", "w"), indent=1)
