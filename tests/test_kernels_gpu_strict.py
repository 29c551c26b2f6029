"""Element-level checks of the sm_100a kernels: a relative Frobenius norm can hide one wrong row or a mis-masked diagonal
tile in a large output, so these tests bound the worst row and the worst element, and use shapes that are not multiples of
any tile (ragged sequence lengths for attention, token counts that do not divide the loss chunk)."""

import math

import pytest
import torch

from accelerate_hpc_test_b200.ops import functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF = torch.bfloat16


def worst_row_rel(a, b):
    """max over rows of ||a_r - b_r|| / (||b_r|| + eps * ||b||_rms)."""
    a, b = a.float().reshape(-1, a.shape[-1]), b.float().reshape(-1, b.shape[-1])
    floor = b.norm() / math.sqrt(b.shape[0]) * 1e-2 + 1e-12
    return ((a - b).norm(dim=1) / (b.norm(dim=1) + floor)).max().item()


def max_abs(a, b):
    return (a.float() - b.float()).abs().max().item()


@pytest.mark.parametrize("M,N,K", [(257, 200, 136), (1000, 712, 520), (4096, 384, 4096), (130, 72, 264)])
def test_gemm_every_row_and_element(M, N, K):
    torch.manual_seed(1)
    a = torch.randn(M, K, device=DEV, dtype=BF)
    b = torch.randn(N, K, device=DEV, dtype=BF)
    ref = a.float() @ b.float().t()
    out = F.gemm(a, b)
    torch.cuda.synchronize()
    assert worst_row_rel(out, ref) < 1.5e-2
    # fp32 accumulation, one bf16 rounding at the end: every element within one bf16 ulp of the fp32 value
    assert ((out.float() - ref).abs() <= 2.0 ** -7 * ref.abs() + 0.02).all()
    # transposed views (dgrad / wgrad layouts) on the same ragged shapes
    out2 = F.gemm(a.t().contiguous(), b.t().contiguous(), a_mn_major=True, b_mn_major=True)
    torch.cuda.synchronize()
    assert worst_row_rel(out2, ref) < 1.5e-2


def _attn_ref(qkv, B, S, Hq, Hkv, D, causal, scale):
    return F._attn_reference(qkv.float(), B, S, Hq, Hkv, D, causal, scale)


@pytest.mark.parametrize("B,S,Hq,Hkv,causal", [(1, 65, 2, 1, True), (2, 129, 2, 2, True), (1, 191, 4, 1, True), (1, 333, 2, 1, False),
                                               (1, 1000, 4, 2, True), (3, 77, 2, 1, True)])
def test_flash_attention_ragged_lengths_row_by_row(B, S, Hq, Hkv, causal):
    """Sequence lengths that end in the middle of a 128-row Q tile and of a 64-row KV block: the masked diagonal tile and the
    out-of-range tail rows are where a wrong mask would show, and only in a few rows."""
    torch.manual_seed(12)
    D = 128
    W = (Hq + 2 * Hkv) * D
    qkv = (torch.randn(B * S, W, device=DEV) * 0.7).to(BF)
    scale = 1.0 / math.sqrt(D)
    o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, causal, scale)
    x = qkv.detach().float().requires_grad_(True)
    ref = _attn_ref(x, B, S, Hq, Hkv, D, causal, scale)
    do = torch.randn_like(o)
    (ref_d,) = torch.autograd.grad(ref, x, do.float())
    torch.cuda.synchronize()
    assert torch.isfinite(o.float()).all() and torch.isfinite(lse).all()
    assert worst_row_rel(o, ref) < 3e-2, ("o", worst_row_rel(o, ref))
    assert max_abs(o, ref) < 3e-2
    # log-sum-exp of every row against the fp32 value
    q = qkv[:, : Hq * D].float().view(B, S, Hq, D).transpose(1, 2)
    k = qkv[:, Hq * D : (Hq + Hkv) * D].float().view(B, S, Hkv, D).transpose(1, 2).repeat_interleave(Hq // Hkv, dim=1)
    s = (q @ k.transpose(-1, -2)) * scale
    if causal:
        s = s.masked_fill(~torch.ones(S, S, dtype=torch.bool, device=DEV).tril(), float("-inf"))
    assert max_abs(lse, torch.logsumexp(s, dim=-1)) < 2e-2
    dqkv = F.flash_attn_bwd(do, qkv, o, lse, B, S, Hq, Hkv, D, causal, scale)
    torch.cuda.synchronize()
    assert torch.isfinite(dqkv.float()).all()
    nq, nk = Hq * D, Hkv * D
    for name, lo, hi in (("dq", 0, nq), ("dk", nq, nq + nk), ("dv", nq + nk, nq + 2 * nk)):
        assert worst_row_rel(dqkv[:, lo:hi], ref_d[:, lo:hi]) < 6e-2, (name, worst_row_rel(dqkv[:, lo:hi], ref_d[:, lo:hi]))


@pytest.mark.parametrize("T,V,chunk", [(1000, 128256, 384), (300, 50257, 128), (4097, 32000, 4096)])
def test_fused_linear_cross_entropy_uneven_chunks(T, V, chunk):
    """Token counts that leave a short last chunk, a vocabulary that is not a multiple of anything, ignored labels inside."""
    torch.manual_seed(3)
    H = 256
    h = (torch.randn(T, H, device=DEV) * 0.5).to(BF).requires_grad_(True)
    w = (torch.randn(V, H, device=DEV) * 0.05).to(BF).requires_grad_(True)
    labels = torch.randint(0, V, (T,), device=DEV)
    labels[::7] = -100
    loss = F.fused_linear_cross_entropy(h, w, labels, -100, chunk)
    loss.backward()
    hf, wf = h.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(hf @ wf.t(), labels, ignore_index=-100)
    ref.backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - ref.item()) < 2e-3 * abs(ref.item())
    assert worst_row_rel(h.grad, hf.grad) < 5e-2 and max_abs(h.grad, hf.grad) < 5e-2 * hf.grad.abs().max().item() + 1e-6
    assert (h.grad[::7].float().abs().max().item()) == 0.0  # ignored tokens get exactly zero gradient
    assert ((w.grad.float() - wf.grad).norm() / wf.grad.norm()).item() < 3e-2


@pytest.mark.parametrize("T,H", [(1, 4096), (37, 1024), (1001, 4096), (513, 8192)])
def test_rmsnorm_rows(T, H):
    torch.manual_seed(4)
    x = torch.randn(T, H, device=DEV).to(BF)
    res = torch.randn(T, H, device=DEV).to(BF)
    w = (1 + 0.1 * torch.randn(H, device=DEV)).to(BF)
    y, s, rstd = F.rmsnorm_fwd(x, w, 1e-5, residual=res)
    sf = (x.float() + res.float()).to(BF).float()
    ref = sf * torch.rsqrt(sf.pow(2).mean(-1, keepdim=True) + 1e-5) * w.float()
    torch.cuda.synchronize()
    assert worst_row_rel(y, ref) < 1e-2 and max_abs(y, ref) < 0.06
    assert max_abs(rstd, torch.rsqrt(sf.pow(2).mean(-1) + 1e-5)) < 1e-3
