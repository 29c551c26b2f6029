"""Native Llama model: CPU oracle path trains, GPU path matches the CPU fp32 oracle."""

import pytest
import torch

from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW, grad_global_norm


def _batch(cfg, B=2, S=64, device="cpu", seed=0):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, cfg.vocab_size, (B, S), generator=g)
    return ids.to(device)


def test_cpu_tiny_llama_trains():
    torch.manual_seed(0)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    opt = FusedAdamW(model.parameters(), lr=3e-3, weight_decay=0.0)
    ids = _batch(cfg, 2, 32)
    losses = []
    for _ in range(8):
        out = model(ids, labels=ids)
        out.loss.backward()
        norm, coef, inf = grad_global_norm(model.parameters(), 1.0)
        opt.set_clip_coef(coef)
        opt.step()
        opt.zero_grad()
        losses.append(out.loss.item())
    assert losses[-1] < losses[0] - 0.3, losses


def test_cpu_matches_plain_torch_llama():
    """The fused-parameter model equals a straightforward fp32 PyTorch restatement of Llama."""
    torch.manual_seed(1)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=96, vocab_size=96, num_attention_heads=4,
                           num_key_value_heads=2)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    ids = _batch(cfg, 2, 24, seed=3)
    out = model(ids, labels=ids, return_logits=True)
    sd = model.hf_state_dict()

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + cfg.rms_norm_eps) * w

    B, S = ids.shape
    D, Hq, Hkv = cfg.head_dim, cfg.num_attention_heads, cfg.num_key_value_heads
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, D, 2).float() / D))
    ang = torch.arange(S).float()[:, None] * inv[None]
    cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)

    def rot(x):
        h = x.shape[-1] // 2
        return torch.cat([-x[..., h:], x[..., :h]], -1)

    x = sd["model.embed_tokens.weight"][ids]
    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}."
        h = rms(x, sd[p + "input_layernorm.weight"])
        q = (h @ sd[p + "self_attn.q_proj.weight"].t()).view(B, S, Hq, D).transpose(1, 2)
        k = (h @ sd[p + "self_attn.k_proj.weight"].t()).view(B, S, Hkv, D).transpose(1, 2)
        v = (h @ sd[p + "self_attn.v_proj.weight"].t()).view(B, S, Hkv, D).transpose(1, 2)
        q, k = q * cos + rot(q) * sin, k * cos + rot(k) * sin
        k, v = k.repeat_interleave(Hq // Hkv, 1), v.repeat_interleave(Hq // Hkv, 1)
        a = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True)
        x = x + a.transpose(1, 2).reshape(B, S, Hq * D) @ sd[p + "self_attn.o_proj.weight"].t()
        h = rms(x, sd[p + "post_attention_layernorm.weight"])
        g, u = h @ sd[p + "mlp.gate_proj.weight"].t(), h @ sd[p + "mlp.up_proj.weight"].t()
        x = x + (torch.nn.functional.silu(g) * u) @ sd[p + "mlp.down_proj.weight"].t()
    logits = rms(x, sd["model.norm.weight"]) @ sd["lm_head.weight"].t()
    ref_loss = torch.nn.functional.cross_entropy(logits[:, :-1].reshape(-1, cfg.vocab_size), ids[:, 1:].reshape(-1))
    torch.testing.assert_close(out.logits, logits, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(out.loss, ref_loss, rtol=1e-4, atol=1e-4)


@pytest.mark.gpu
def test_gpu_llama_matches_cpu_oracle():
    torch.manual_seed(2)
    cfg = LlamaConfig.tiny()
    ref = LlamaForCausalLM(cfg, dtype=torch.float32)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
            q.copy_(p.float().cpu())
    ids = _batch(cfg, 2, 192, seed=5)
    out_ref = ref(ids, labels=ids)
    out_ref.loss.backward()
    out = model(ids.cuda(), labels=ids.cuda())
    out.loss.backward()
    torch.cuda.synchronize()
    assert abs(out.loss.item() - out_ref.loss.item()) < 3e-2 * abs(out_ref.loss.item())
    for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
        err = ((p.grad.float().cpu() - q.grad).norm() / q.grad.norm().clamp(min=1e-9)).item()
        assert err < 8e-2, (n, err)


@pytest.mark.gpu
def test_gpu_llama_trains_and_counts_launches():
    from accelerate_hpc_test_b200.ops import _lib

    torch.manual_seed(3)
    cfg = LlamaConfig.tiny()
    model = LlamaForCausalLM(cfg, device="cuda")
    opt = FusedAdamW(model.parameters(), lr=2e-3, weight_decay=0.0)
    ids = _batch(cfg, 4, 256, seed=6).cuda()
    losses = []
    _lib.get_lib()
    _lib.reset_launch_count()
    for _ in range(10):
        out = model(ids, labels=ids)
        out.loss.backward()
        norm, coef, inf = grad_global_norm(model.parameters(), 1.0)
        opt.set_clip_coef(coef)
        opt.step()
        opt.zero_grad(set_to_none=True)
        losses.append(out.loss.item())
    assert losses[-1] < losses[0] - 0.5, losses
    assert _lib.launch_count() > 100
