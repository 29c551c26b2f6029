"""Numerics of every sm_100a kernel against a plain PyTorch fp32 restatement of the same op."""

import math

import pytest
import torch

from accelerate_hpc_test_b200.ops import _lib
from accelerate_hpc_test_b200.ops import functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF = torch.bfloat16


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp(min=1e-12)).item()


def test_library_is_native_and_loaded():
    lib = _lib.get_lib()
    assert lib.b200_version() >= 100
    assert _lib.loaded()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (1000, 712, 520), (2048, 4096, 4096), (48, 8, 72)])
def test_gemm_layouts(a_mn, b_mn, M, N, K):
    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device=DEV, dtype=BF)
    b = torch.randn((K, N) if b_mn else (N, K), device=DEV, dtype=BF)
    out = F.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn)
    af = a.float().t() if a_mn else a.float()
    bf = b.float() if b_mn else b.float().t()
    ref = af @ bf
    torch.cuda.synchronize()
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)


def test_gemm_epilogues():
    torch.manual_seed(1)
    M, N, K = 384, 640, 320
    a = torch.randn(M, K, device=DEV, dtype=BF)
    b = torch.randn(N, K, device=DEV, dtype=BF)
    bias = torch.randn(N, device=DEV, dtype=BF)
    ref = a.float() @ b.float().t()
    o32 = F.gemm(a, b, out_dtype=torch.float32)
    assert rel_err(o32, ref) < 1e-3
    ob = F.gemm(a, b, bias=bias)
    assert rel_err(ob, ref + bias.float()) < 1e-2
    acc = torch.ones(M, N, device=DEV, dtype=torch.float32)
    F.gemm(a, b, out=acc, accumulate=True)
    assert rel_err(acc, ref + 1) < 1e-3
    accb = torch.ones(M, N, device=DEV, dtype=BF)
    F.gemm(a, b, out=accb, accumulate=True)
    assert rel_err(accb, ref + 1) < 1e-2
    # strided output (column slice of a wider buffer)
    wide = torch.zeros(M, 2 * N, device=DEV, dtype=BF)
    F.gemm(a, b, out=wide[:, N:])
    assert rel_err(wide[:, N:], ref) < 1e-2
    assert wide[:, :N].abs().max().item() == 0


def test_linear_autograd():
    torch.manual_seed(2)
    x = torch.randn(4, 96, 256, device=DEV, dtype=BF, requires_grad=True)
    w = torch.randn(512, 256, device=DEV, dtype=BF, requires_grad=True)
    y = F.linear(x, w)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().float().requires_grad_(True)
    yr = xr @ wr.t()
    yr.backward(dy.float())
    assert rel_err(y, yr) < 1e-2
    assert rel_err(x.grad, xr.grad) < 1e-2
    assert rel_err(w.grad, wr.grad) < 1e-2


@pytest.mark.parametrize("T,H", [(7, 256), (300, 4096), (64, 1024)])
def test_rmsnorm(T, H):
    torch.manual_seed(3)
    x = torch.randn(T, H, device=DEV, dtype=BF)
    r = torch.randn(T, H, device=DEV, dtype=BF)
    w = (1 + 0.1 * torch.randn(H, device=DEV)).to(BF)
    y, s, rstd = F.rmsnorm_fwd(x, w, 1e-5, residual=r)
    yc, sc, rc = F.rmsnorm_fwd(x.cpu(), w.cpu(), 1e-5, residual=r.cpu())
    assert rel_err(y.cpu(), yc) < 1e-2 and rel_err(s.cpu(), sc) < 1e-3 and rel_err(rstd.cpu(), rc) < 1e-3
    y2, s2, _ = F.rmsnorm_fwd(x, w, 1e-5)
    y2c, _, _ = F.rmsnorm_fwd(x.cpu(), w.cpu(), 1e-5)
    assert rel_err(y2.cpu(), y2c) < 1e-2
    dy = torch.randn(T, H, device=DEV, dtype=BF)
    dres = torch.randn(T, H, device=DEV, dtype=BF)
    dx, dw = F.rmsnorm_bwd(dy, s, w, rstd, dres=dres)
    dxc, dwc = F.rmsnorm_bwd(dy.cpu(), s.cpu(), w.cpu(), rstd.cpu(), dres=dres.cpu())
    assert rel_err(dx.cpu(), dxc) < 1e-2
    assert rel_err(dw.cpu(), dwc) < 2e-2


def test_rope_roundtrip_and_reference():
    torch.manual_seed(4)
    T, Hq, Hkv, D = 200, 4, 2, 128
    W = (Hq + 2 * Hkv) * D
    buf = torch.randn(T, W, device=DEV, dtype=BF)
    orig = buf.clone()
    cos, sin = F.rope_tables(512, D, 5e5, DEV)
    pos = torch.randint(0, 512, (T,), device=DEV, dtype=torch.int32)
    F.rope_inplace(buf, cos, sin, pos, Hq + Hkv, D)
    ref = orig.cpu().clone()
    F.rope_inplace(ref, cos.cpu(), sin.cpu(), pos.cpu(), Hq + Hkv, D)
    assert rel_err(buf.cpu(), ref) < 1e-2
    assert torch.equal(buf[:, (Hq + Hkv) * D :], orig[:, (Hq + Hkv) * D :])  # V untouched
    F.rope_inplace(buf, cos, sin, pos, Hq + Hkv, D, backward=True)
    assert rel_err(buf, orig) < 2e-2


def test_swiglu():
    torch.manual_seed(5)
    gu = torch.randn(130, 2 * 1024, device=DEV, dtype=BF)
    act = F.swiglu_fwd(gu)
    assert rel_err(act.cpu(), F.swiglu_fwd(gu.cpu())) < 1e-2
    d = torch.randn_like(act)
    dgu = F.swiglu_bwd(gu, d)
    assert rel_err(dgu.cpu(), F.swiglu_bwd(gu.cpu(), d.cpu())) < 1e-2


def test_embedding():
    torch.manual_seed(6)
    V, H = 1000, 256
    table = torch.randn(V, H, device=DEV, dtype=BF, requires_grad=True)
    ids = torch.randint(0, V, (4, 33), device=DEV)
    out = F.embedding(ids, table)
    assert torch.equal(out, table.detach()[ids])
    d = torch.randn_like(out)
    out.backward(d)
    ref = torch.zeros(V, H, dtype=torch.float32, device=DEV).index_add_(0, ids.reshape(-1), d.reshape(-1, H).float())
    assert rel_err(table.grad, ref) < 2e-2


@pytest.mark.parametrize("T,V", [(64, 1024), (33, 128256)])
def test_cross_entropy_chunk(T, V):
    torch.manual_seed(7)
    logits = (torch.randn(T, V, device=DEV) * 3).to(BF)
    labels = torch.randint(0, V, (T,), device=DEV)
    labels[::5] = -100
    inv = torch.tensor([1.0 / (labels != -100).sum().item()], device=DEV)
    ref_logits = logits.float().clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(ref_logits, labels, ignore_index=-100, reduction="sum") * inv
    ref.backward()
    rows = F.cross_entropy_chunk_(logits, labels, inv)
    assert abs((rows.sum() * inv).item() - ref.item()) < 2e-3 * abs(ref.item())
    assert rel_err(logits, ref_logits.grad) < 2e-2


def test_fused_linear_cross_entropy():
    torch.manual_seed(8)
    T, H, V = 300, 256, 2048
    h = (torch.randn(T, H, device=DEV) * 0.5).to(BF).requires_grad_(True)
    w = (torch.randn(V, H, device=DEV) * 0.05).to(BF).requires_grad_(True)
    labels = torch.randint(0, V, (T,), device=DEV)
    labels[:7] = -100
    loss = F.fused_linear_cross_entropy(h, w, labels, chunk_tokens=128)
    (loss * 0.5).backward()
    hr = h.detach().float().requires_grad_(True)
    wr = w.detach().float().requires_grad_(True)
    lr = torch.nn.functional.cross_entropy(hr @ wr.t(), labels, ignore_index=-100)
    (lr * 0.5).backward()
    assert abs(loss.item() - lr.item()) < 1e-2 * abs(lr.item())
    assert rel_err(h.grad, hr.grad) < 3e-2
    assert rel_err(w.grad, wr.grad) < 3e-2


def test_adamw_and_gradnorm():
    torch.manual_seed(9)
    n = 100_003
    p = torch.randn(n, device=DEV)
    g = torch.randn(n, device=DEV).to(BF)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    sh = torch.empty(n, device=DEV, dtype=BF)
    pr, mr, vr = p.cpu().clone(), m.cpu().clone(), v.cpu().clone()
    ss = torch.zeros(2, device=DEV)
    F.sumsq_(g, ss)
    norm = torch.zeros(1, device=DEV)
    coef = torch.zeros(1, device=DEV)
    F.clip_coef_(ss, 1.0, norm, coef)
    assert abs(norm.item() - g.float().norm().item()) < 1e-3 * norm.item()
    assert ss[1].item() == 0
    kw = dict(lr=1e-2, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1)
    for step in (1, 2, 3):
        F.adamw_flat_(p, m, v, g, sh, step=step, clip_coef=coef, **kw)
        F.adamw_flat_(pr, mr, vr, g.cpu(), None, step=step, clip_coef=coef.cpu(), **kw)
    assert rel_err(p.cpu(), pr) < 1e-5
    assert rel_err(sh.cpu(), pr) < 1e-2
    g[5] = float("inf")
    ss.zero_()
    F.sumsq_(g, ss)
    assert ss[1].item() == 1


def _attn_ref_grads(qkv, do, B, S, Hq, Hkv, D, causal, scale):
    x = qkv.detach().float().requires_grad_(True)
    o = F._attn_reference(x, B, S, Hq, Hkv, D, causal, scale)
    (dx,) = torch.autograd.grad(o, x, do.float())
    return o, dx


@pytest.mark.parametrize("B,S,Hq,Hkv,causal", [(1, 128, 1, 1, True), (2, 256, 4, 2, True), (1, 200, 2, 1, True),
                                               (1, 1024, 4, 1, True), (2, 192, 2, 2, False), (1, 2048, 8, 2, True)])
def test_flash_attention_fwd_bwd(B, S, Hq, Hkv, causal):
    torch.manual_seed(10)
    D = 128
    W = (Hq + 2 * Hkv) * D
    qkv = (torch.randn(B * S, W, device=DEV) * 0.7).to(BF)
    scale = 1.0 / math.sqrt(D)
    o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, causal, scale)
    do = torch.randn_like(o)
    ref_o, ref_dqkv = _attn_ref_grads(qkv, do, B, S, Hq, Hkv, D, causal, scale)
    torch.cuda.synchronize()
    assert rel_err(o, ref_o) < 2e-2, ("fwd", rel_err(o, ref_o))
    dqkv = F.flash_attn_bwd(do, qkv, o, lse, B, S, Hq, Hkv, D, causal, scale)
    torch.cuda.synchronize()
    nq, nk = Hq * D, Hkv * D
    assert rel_err(dqkv[:, nq + nk :], ref_dqkv[:, nq + nk :]) < 3e-2, ("dv", rel_err(dqkv[:, nq + nk :], ref_dqkv[:, nq + nk :]))
    assert rel_err(dqkv[:, nq : nq + nk], ref_dqkv[:, nq : nq + nk]) < 3e-2, ("dk", rel_err(dqkv[:, nq : nq + nk], ref_dqkv[:, nq : nq + nk]))
    assert rel_err(dqkv[:, :nq], ref_dqkv[:, :nq]) < 3e-2, ("dq", rel_err(dqkv[:, :nq], ref_dqkv[:, :nq]))


def test_rope_attention_autograd():
    torch.manual_seed(11)
    B, S, Hq, Hkv, D = 2, 160, 4, 2, 128
    W = (Hq + 2 * Hkv) * D
    base = (torch.randn(B * S, W, device=DEV) * 0.7).to(BF)
    cos, sin = F.rope_tables(S, D, 5e5, DEV)
    pos = torch.arange(S, device=DEV, dtype=torch.int32).repeat(B)
    x = base.clone().requires_grad_(True)
    o = F.rope_attention(x * 1.0, cos, sin, pos, B, S, Hq, Hkv, D)
    do = torch.randn_like(o)
    o.backward(do)
    xc = base.cpu().clone().requires_grad_(True)
    oc = F.rope_attention(xc * 1.0, cos.cpu(), sin.cpu(), pos.cpu(), B, S, Hq, Hkv, D)
    oc.backward(do.cpu())
    assert rel_err(o.cpu(), oc) < 2e-2
    assert rel_err(x.grad.cpu(), xc.grad) < 4e-2


@pytest.mark.gpu
def test_fused_adamw_update_in_backward_matches_plain_step():
    """FusedAdamW with the update fused into backward (side stream, grads released) == the ordinary step."""
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.optim import FusedAdamW

    cfg = LlamaConfig.tiny(num_hidden_layers=2)
    torch.manual_seed(0)
    a = LlamaForCausalLM(cfg, device="cuda")
    torch.manual_seed(0)
    b = LlamaForCausalLM(cfg, device="cuda")
    oa = FusedAdamW(a.parameters(), lr=1e-3, weight_decay=0.01)
    ob = FusedAdamW(b.parameters(), lr=1e-3, weight_decay=0.01)
    ob.enable_update_in_backward(mode="auto")
    ids = torch.randint(0, cfg.vocab_size, (2, 256), device="cuda")
    for step in range(4):
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            m(ids, labels=ids).loss.backward()
            o.step()
        if step == 0:
            assert ob._bwd_mode == "probing"  # one clean step is not enough: a hand-accumulating loop also starts with one
        elif step == 1:
            assert ob._bwd_mode == "active"  # two steps in a row with one backward each and no clipping: fusion switched on
        else:
            assert all(p.grad is None for p in b.parameters())
    torch.cuda.synchronize()
    # dQ and the embedding gradient are accumulated with (order-dependent) fp32 atomics and Adam turns a sign flip of a
    # near-zero gradient into a full +-lr step, so single elements may differ by a few lr; on average the runs coincide
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        d = (p.float() - q.float()).abs()
        assert d.max().item() <= 3 * 2 * 1e-3 + 1e-4, (n, d.max().item())
        assert d.mean().item() <= 0.05 * 1e-3, (n, d.mean().item())
    with pytest.raises(RuntimeError):
        b(ids, labels=ids).loss.backward()
        ob.set_clip_coef(torch.ones(1, device="cuda"))
