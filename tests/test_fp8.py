"""Native fp8 path: quantisation, fp8 GEMM and the fp8 linear layer (reference: tests/test_fp8.py, utils/ao.py,
utils/transformer_engine.py — there the check is "the model converted and trains"; here the kernels are compared
against an fp32 emulation of the same per-tensor-scaled e4m3 arithmetic)."""

import pytest
import torch

from accelerate_hpc_test_b200.parallel.precision import (FP8, convert_model_to_mxfp8, fp8_linear, gemm_e4m3, has_fp8_layers,
                                                         quantize_e4m3)


def _emulate_quant(x):
    amax = x.float().abs().max()
    scale = 448.0 / amax
    q = (x.float() * scale).clamp(-448, 448).to(FP8)
    return q, 1.0 / scale


def test_fp8_linear_cpu_matches_fp32_within_fp8_error():
    torch.manual_seed(0)
    x = torch.randn(64, 128, requires_grad=True)
    w = (torch.randn(96, 128) * 0.05).requires_grad_(True)
    y = fp8_linear(x, w)
    ref = x @ w.t()
    assert ((y - ref).norm() / ref.norm()).item() < 0.06
    dy = torch.randn_like(y)
    y.backward(dy)
    assert ((x.grad - dy @ w).norm() / (dy @ w).norm()).item() < 0.06
    assert torch.allclose(w.grad, dy.t() @ x.detach(), atol=1e-4)  # wgrad stays in high precision


def test_convert_model_keeps_first_and_last_linear():
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=128, num_hidden_layers=2)
    m = convert_model_to_mxfp8(LlamaForCausalLM(cfg, dtype=torch.float32))
    assert has_fp8_layers(m)
    assert not getattr(m.model.layers[0].self_attn.qkv_proj, "_b200_fp8", False)
    assert not getattr(m.lm_head, "_b200_fp8", False)
    assert m.model.layers[1].mlp.down_proj._b200_fp8
    ids = torch.randint(0, cfg.vocab_size, (2, 16))
    out = m(ids, labels=ids)
    out.loss.backward()
    assert torch.isfinite(out.loss) and m.model.layers[1].mlp.down_proj.weight.grad is not None


@pytest.mark.gpu
def test_fp8_quantize_matches_emulation():
    torch.manual_seed(0)
    for R, C in [(256, 512), (130, 200), (64, 4096)]:
        x = (torch.randn(R, C, device="cuda") * 3).to(torch.bfloat16)
        q, qt, deq = quantize_e4m3(x, transpose=True)
        qe, deqe = _emulate_quant(x)
        assert abs(deq.item() - deqe.item()) < 1e-6 * deqe.item() + 1e-12
        same = (q.float() == qe.float()).float().mean().item()
        assert same > 0.999, same  # ties in the scale multiply may round differently
        assert torch.equal(qt.float(), q.float().t())


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(512, 512, 256), (1000, 384, 4096), (4096, 6144, 4096)])
def test_fp8_gemm_matches_emulation(M, N, K):
    torch.manual_seed(1)
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda").to(torch.bfloat16)
    aq, _, da = quantize_e4m3(a)
    bq, _, db = quantize_e4m3(b)
    y = gemm_e4m3(aq, bq, da, db, bias=bias)
    ref = (aq.float() @ bq.float().t()) * (da * db) + bias.float()
    err = ((y.float() - ref).norm() / ref.norm()).item()
    assert err < 5e-3, err  # only the bf16 rounding of the output differs
    full = a.float() @ b.float().t() + bias.float()
    assert ((y.float() - full).norm() / full.norm()).item() < 0.06


@pytest.mark.gpu
def test_fp8_linear_gpu_backward():
    torch.manual_seed(2)
    x = torch.randn(4, 128, 512, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    w = (torch.randn(1024, 512, device="cuda") * 0.03).to(torch.bfloat16).requires_grad_(True)
    y = fp8_linear(x, w)
    dy = torch.randn_like(y)
    y.backward(dy)
    ref_y = x.float() @ w.float().t()
    ref_dx = dy.float() @ w.float()
    ref_dw = dy.float().reshape(-1, 1024).t() @ x.float().reshape(-1, 512)
    assert ((y.float() - ref_y).norm() / ref_y.norm()).item() < 0.06
    assert ((x.grad.float() - ref_dx).norm() / ref_dx.norm()).item() < 0.06
    assert ((w.grad.float() - ref_dw).norm() / ref_dw.norm()).item() < 0.01


def test_mxfp8_linear_cpu_emulation():
    from accelerate_hpc_test_b200.parallel.precision import mxfp8_linear, quantize_mx

    torch.manual_seed(0)
    x = torch.randn(64, 256, requires_grad=True)
    w = (torch.randn(128, 256) * 0.05).requires_grad_(True)
    t = quantize_mx(x.detach())
    assert t.q.shape == (64, 256) and t.sf.shape == (64, 8)
    assert ((t.dequantize() - x.detach()).norm() / x.detach().norm()).item() < 0.05
    y = mxfp8_linear(x, w)
    ref = x @ w.t()
    assert ((y - ref).norm() / ref.norm()).item() < 0.06
    dy = torch.randn_like(y)
    y.backward(dy)
    assert ((x.grad - dy @ w).norm() / (dy @ w).norm()).item() < 0.06


@pytest.mark.gpu
@pytest.mark.parametrize("R,C", [(128, 128), (300, 512), (4096, 1024)])
def test_mx_quantizers_match_emulation(R, C):
    from accelerate_hpc_test_b200.parallel.precision import _mx_emulate, quantize_mx

    torch.manual_seed(0)
    x = (torch.randn(R, C, device="cuda") * torch.rand(R, 1, device="cuda") * 4).to(torch.bfloat16)
    t = quantize_mx(x)
    qe, se = _mx_emulate(x)
    assert torch.equal(t.q.float(), qe.float())
    # un-swizzle the scale bytes: block (r/128, g/4), byte (r%32)*16 + ((r/32)%4)*4 + g%4
    sf = t.sf.view(-1, C // 128, 32, 4, 4)  # [mblk (padded to whole 256-row tiles), kblk, r%32, (r/32)%4, g%4]
    r = torch.arange(R, device="cuda")
    g = torch.arange(C // 32, device="cuda")
    got = sf[(r // 128)[:, None], (g // 4)[None, :], (r % 32)[:, None], ((r // 32) % 4)[:, None], (g % 4)[None, :]]
    assert torch.equal(torch.exp2(got.float() - 127.0), se)
    if R % 128 == 0:
        tt = quantize_mx(x, transpose=True)
        qte, _ = _mx_emulate(x.t().contiguous())
        assert torch.equal(tt.q.float(), qte.float())


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (256, 384, 512), (1000, 640, 1024), (4096, 6144, 4096)])
def test_mxfp8_gemm_matches_emulation(M, N, K):
    from accelerate_hpc_test_b200.parallel.precision import _mx_emulate, gemm_mxfp8, quantize_mx

    torch.manual_seed(1)
    a = (torch.randn(M, K, device="cuda") * (1 + 3 * torch.rand(M, 1, device="cuda"))).to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    y = gemm_mxfp8(quantize_mx(a), quantize_mx(b))
    qa, sa = _mx_emulate(a)
    qb, sb = _mx_emulate(b)
    da = (qa.float().view(M, K // 32, 32) * sa[..., None]).view(M, K)
    db = (qb.float().view(N, K // 32, 32) * sb[..., None]).view(N, K)
    ref = da @ db.t()
    err = ((y.float() - ref).norm() / ref.norm()).item()
    assert err < 5e-3, err
    full = a.float() @ b.float().t()
    assert ((y.float() - full).norm() / full.norm()).item() < 0.05


@pytest.mark.gpu
def test_mxfp8_linear_gpu_backward():
    from accelerate_hpc_test_b200.parallel.precision import mxfp8_linear

    torch.manual_seed(2)
    x = torch.randn(4, 128, 512, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    w = (torch.randn(1024, 512, device="cuda") * 0.03).to(torch.bfloat16).requires_grad_(True)
    y = mxfp8_linear(x, w)
    dy = torch.randn_like(y)
    y.backward(dy)
    ref_y = x.float() @ w.float().t()
    ref_dx = dy.float() @ w.float()
    ref_dw = dy.float().reshape(-1, 1024).t() @ x.float().reshape(-1, 512)
    assert ((y.float() - ref_y).norm() / ref_y.norm()).item() < 0.05
    assert ((x.grad.float() - ref_dx).norm() / ref_dx.norm()).item() < 0.05
    assert ((w.grad.float() - ref_dw).norm() / ref_dw.norm()).item() < 0.01
