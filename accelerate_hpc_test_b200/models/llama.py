"""Native Llama-3 family decoder built directly on the sm_100a kernels.

Design (B200-first, not a translation of transformers' ``LlamaForCausalLM``):

* Q/K/V and gate/up projections are stored *fused* (one GEMM each), RoPE is applied in place on
  the fused QKV GEMM output, attention reads Q/K/V straight out of that buffer through TMA
  descriptors, and the residual add is fused into the following RMSNorm.
* The LM head and the loss are one op (chunked GEMM + cross-entropy + both gradient GEMMs), so
  ``[tokens, vocab]`` logits never exist.  The reference instead materialises logits and then
  up-casts them to fp32 (``convert_outputs_to_fp32``, reference src/accelerate/accelerator.py:1797-1802).
* Parameters are bf16; fp32 master weights live in the fused optimizer (:mod:`..optim`).

The BASELINE.json configs name Llama-3-8B (hidden 4096, FFN 14336, 32 layers, 32/8 heads,
head_dim 128, vocab 128256, rope theta 5e5, eps 1e-5).  ``hf_state_dict`` /
``load_hf_state_dict`` convert to and from the transformers parameter naming so checkpoints
interchange.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as F


@dataclass
class LlamaConfig:
    vocab_size: int = 128256
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: int = 128
    rope_theta: float = 500000.0
    rms_norm_eps: float = 1e-5
    max_position_embeddings: int = 8192
    tie_word_embeddings: bool = False
    initializer_range: float = 0.02
    loss_chunk_tokens: int = 4096
    model_type: str = "llama"

    @classmethod
    def llama3_8b(cls, **kw) -> "LlamaConfig":
        return cls(**kw)

    @classmethod
    def llama3_1b(cls, **kw) -> "LlamaConfig":
        base = dict(hidden_size=2048, intermediate_size=8192, num_hidden_layers=16, num_attention_heads=16,
                    num_key_value_heads=8)
        base.update(kw)
        return cls(**base)

    @classmethod
    def tiny(cls, **kw) -> "LlamaConfig":
        base = dict(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                    num_attention_heads=2, num_key_value_heads=1, max_position_embeddings=512, loss_chunk_tokens=256)
        base.update(kw)
        return cls(**base)

    @property
    def qkv_width(self) -> int:
        return (self.num_attention_heads + 2 * self.num_key_value_heads) * self.head_dim

    def num_parameters(self) -> int:
        h, i = self.hidden_size, self.intermediate_size
        per_layer = h * self.qkv_width + self.num_attention_heads * self.head_dim * h + 3 * h * i + 2 * h
        head = 0 if self.tie_word_embeddings else self.vocab_size * h
        return self.num_hidden_layers * per_layer + self.vocab_size * h + head + h

    def flops_per_token(self, seq_len: int) -> float:
        """Training FLOPs per token (fwd + bwd = 3x fwd), causal attention counted at half."""
        h = self.hidden_size
        matmul_params = self.num_parameters() - (0 if self.tie_word_embeddings else self.vocab_size * h)
        lin = 2.0 * matmul_params  # the embedding gather is not a matmul; the LM head is
        attn = 2.0 * self.num_hidden_layers * self.num_attention_heads * self.head_dim * seq_len
        return 3.0 * (lin + attn)


@dataclass
class CausalLMOutput:
    loss: Optional[torch.Tensor] = None
    logits: Optional[torch.Tensor] = None
    hidden_states: Optional[torch.Tensor] = None

    def __getitem__(self, k):
        if isinstance(k, str):
            return getattr(self, k)
        return tuple(v for v in (self.loss, self.logits) if v is not None)[k]


class B200Linear(nn.Module):
    """Bias-free (by default) linear layer on the tcgen05 GEMM; weight is ``[out, in]`` like nn.Linear."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features, dtype=dtype, device=device))
        self.bias = nn.Parameter(torch.zeros(out_features, dtype=dtype, device=device)) if bias else None

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"


class B200RMSNorm(nn.Module):
    """RMSNorm with the residual add fused in: ``forward(x, residual) -> (normed, new_residual)`` on the sm_100a kernel."""

    def __init__(self, hidden_size: int, eps: float = 1e-5, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x, residual=None):
        if residual is None:
            return F.rmsnorm(x, self.weight, self.eps), x
        return F.add_rmsnorm(x, residual, self.weight, self.eps)


class LlamaAttention(nn.Module):
    """Grouped-query causal self-attention: one fused QKV projection, RoPE + flash attention as one op (Ulysses / context-parallel variants when those axes are on), output projection."""

    def __init__(self, cfg: LlamaConfig, dtype, device=None):
        super().__init__()
        self.cfg = cfg
        self.qkv_proj = B200Linear(cfg.hidden_size, cfg.qkv_width, dtype=dtype, device=device)
        self.o_proj = B200Linear(cfg.num_attention_heads * cfg.head_dim, cfg.hidden_size, dtype=dtype, device=device)

    def forward(self, x, rope, B, S):
        c = self.cfg
        cos, sin, pos = rope[:3]
        hq, hkv = getattr(self, "tp_heads", (c.num_attention_heads, c.num_key_value_heads))  # local heads under TP
        sp = getattr(self, "sp_comm", None)
        cp = getattr(self, "cp_comm", None)
        if sp is not None and sp.world > 1 and len(rope) > 3 and self.qkv_proj.bias is None and self.o_proj.bias is None \
                and not getattr(self.qkv_proj, "_b200_fp8", False):
            from ..parallel.ulysses import ulysses_attention_block

            # projections and all-to-alls are fused: QKV GEMM -> head-sharded buffers, attention, heads -> output GEMM
            return ulysses_attention_block(x, self.qkv_proj.weight, self.o_proj.weight, cos, sin, rope[3], B, S, hq, hkv,
                                           c.head_dim, sp, causal=True)
        qkv = self.qkv_proj(x)
        if sp is not None and sp.world > 1:
            from ..parallel.ulysses import ulysses_attention

            attn = ulysses_attention(qkv, cos, sin, pos, B, S, hq, hkv, c.head_dim, sp, causal=True)
        elif cp is not None and cp.world > 1:
            from ..parallel.context_parallel import cp_attention

            attn = cp_attention(qkv, cos, sin, pos, B, S, hq, hkv, c.head_dim, cp)
        else:
            attn = F.rope_attention(qkv, cos, sin, pos, B, S, hq, hkv, c.head_dim, causal=True)
        return self.o_proj(attn)


class LlamaMLP(nn.Module):
    """Gated MLP: one fused gate/up projection, SwiGLU kernel, down projection."""

    def __init__(self, cfg: LlamaConfig, dtype, device=None):
        super().__init__()
        self.gate_up_proj = B200Linear(cfg.hidden_size, 2 * cfg.intermediate_size, dtype=dtype, device=device)
        self.down_proj = B200Linear(cfg.intermediate_size, cfg.hidden_size, dtype=dtype, device=device)

    def forward(self, x):
        return self.down_proj(F.swiglu(self.gate_up_proj(x)))


class LlamaDecoderLayer(nn.Module):
    """Pre-norm decoder block (attention, MLP) that passes the un-normalised residual stream along with the activations."""

    def __init__(self, cfg: LlamaConfig, layer_idx: int, dtype, device=None):
        super().__init__()
        self.layer_idx = layer_idx
        self.input_layernorm = B200RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, dtype, device)
        self.self_attn = LlamaAttention(cfg, dtype, device)
        self.post_attention_layernorm = B200RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, dtype, device)
        self.mlp = LlamaMLP(cfg, dtype, device)

    def forward(self, x, residual, rope, B, S):
        """``x``: previous sub-layer output, ``residual``: residual stream (None for the first layer)."""
        normed, residual = self.input_layernorm(x, residual)
        a = self.self_attn(normed, rope, B, S)
        normed, residual = self.post_attention_layernorm(a, residual)
        return self.mlp(normed), residual


class LlamaModel(nn.Module):
    """Token embedding, the stack of decoder layers and the final norm; activations are ``[tokens, hidden]`` throughout."""

    def __init__(self, cfg: LlamaConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.cfg = cfg
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size, dtype=dtype, device=device)
        self.layers = nn.ModuleList([LlamaDecoderLayer(cfg, i, dtype, device) for i in range(cfg.num_hidden_layers)])
        self.norm = B200RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, dtype, device)
        self.gradient_checkpointing = False
        self._rope_cache = None

    def rope(self, position_ids: torch.Tensor, max_pos: int):
        dev = position_ids.device
        c = self._rope_cache
        if c is None or c[0].device != dev or c[0].shape[0] < max_pos:
            n = max(max_pos, self.cfg.max_position_embeddings)
            self._rope_cache = F.rope_tables(n, self.cfg.head_dim, self.cfg.rope_theta, dev)
        cos, sin = self._rope_cache
        return cos, sin, position_ids.reshape(-1).to(torch.int32).contiguous()

    def forward(self, input_ids: torch.Tensor, position_ids: Optional[torch.Tensor] = None):
        B, S = input_ids.shape
        sp = getattr(self, "sp_comm", None)
        cp = getattr(self, "cp_comm", None)
        if position_ids is None and sp is not None and sp.world > 1:
            # `input_ids` is this rank's contiguous S/sp slice of the sequence
            position_ids = (torch.arange(S, device=input_ids.device) + sp.rank * S).unsqueeze(0).expand(B, S)
            max_pos = S * sp.world
        elif position_ids is None and cp is not None and cp.world > 1:
            position_ids = cp.local_positions(S, input_ids.device).unsqueeze(0).expand(B, S)
            max_pos = S * cp.world
        elif position_ids is None:
            position_ids = torch.arange(S, device=input_ids.device).unsqueeze(0).expand(B, S)
            max_pos = S
        else:
            max_pos = self.cfg.max_position_embeddings
        rope = self.rope(position_ids, max_pos)
        if sp is not None and sp.world > 1:
            rope = rope + (sp.gather_positions(rope[2], B, S),)  # positions of the whole sequence, head-sharded row order
        x = F.embedding(input_ids, self.embed_tokens.weight).view(B * S, -1)
        tp = getattr(self, "tp_comm", None)
        if tp is not None and tp.world > 1:
            from ..parallel.tensor_parallel import scatter_tokens

            x = scatter_tokens(x, tp)  # activations between blocks are token-sharded (sequence parallel)
        residual = None
        for layer in self.layers:
            if self.gradient_checkpointing and self.training and torch.is_grad_enabled():
                x, residual = torch.utils.checkpoint.checkpoint(layer, x, residual, rope, B, S, use_reentrant=False)
            else:
                x, residual = layer(x, residual, rope, B, S)
        h, _ = self.norm(x, residual)
        if tp is not None and tp.world > 1:
            return h  # [B*S / tp, H]: the caller works on this rank's token shard
        return h.view(B, S, -1)


class LlamaForCausalLM(nn.Module):
    """Causal LM with the fused LM-head + cross-entropy loss.

    ``forward(input_ids, labels=..., position_ids=..., shift_labels=...)`` mirrors the transformers
    call signature the reference's examples use (reference examples/torch_native_parallelism/
    nd_parallel.py:57-80): ``labels`` are shifted inside, ``shift_labels`` (already shifted, used
    by the CP / SP paths) are consumed as they are.  ``return_logits=True`` also materialises logits.
    """

    _no_split_modules = ["LlamaDecoderLayer"]
    _tied_weights_keys = ["lm_head.weight"]

    def __init__(self, cfg: LlamaConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.config = cfg
        self.model = LlamaModel(cfg, dtype, device)
        self.lm_head = B200Linear(cfg.hidden_size, cfg.vocab_size, dtype=dtype, device=device)
        if cfg.tie_word_embeddings:
            self.lm_head.weight = self.model.embed_tokens.weight
        if device is None or torch.device(device).type != "meta":
            self.reset_parameters()

    @torch.no_grad()
    def reset_parameters(self, seed: Optional[int] = None):
        g = None
        std = self.config.initializer_range
        if seed is not None:
            g = torch.Generator(device=self.lm_head.weight.device).manual_seed(seed)
        for name, p in self.named_parameters():
            if p.dim() >= 2:
                # fill through an fp32 staging chunk so bf16 init matches an fp32 init rounded once
                p.normal_(0.0, std, generator=g)
            else:
                p.fill_(1.0)

    def gradient_checkpointing_enable(self, **_):
        self.model.gradient_checkpointing = True

    def gradient_checkpointing_disable(self):
        self.model.gradient_checkpointing = False

    def tie_weights(self):
        if self.config.tie_word_embeddings:
            self.lm_head.weight = self.model.embed_tokens.weight

    def get_input_embeddings(self):
        return self.model.embed_tokens

    def forward(self, input_ids, labels=None, position_ids=None, shift_labels=None, attention_mask=None,
                return_logits: bool = False, **_):
        h = self.model(input_ids, position_ids)
        loss = logits = None
        given_shifted = shift_labels is not None
        if shift_labels is None and labels is not None:
            shift_labels = torch.full_like(labels, -100)
            shift_labels[:, :-1] = labels[:, 1:]
        tp = getattr(self, "tp_comm", None)
        if tp is not None and tp.world > 1:
            return self._tp_loss(h, shift_labels, tp)
        seq = getattr(self, "sp_comm", None) or getattr(self, "cp_comm", None)
        if seq is not None and seq.world > 1 and (shift_labels is not None or labels is not None):
            return self._seq_sharded_loss(h, labels, shift_labels, seq, given_shifted=given_shifted)
        if shift_labels is not None:
            loss = F.fused_linear_cross_entropy(h.reshape(-1, h.shape[-1]), self.lm_head.weight,
                                                shift_labels.reshape(-1), -100, self.config.loss_chunk_tokens)
        if return_logits or shift_labels is None:
            logits = self.lm_head(h)
        return CausalLMOutput(loss=loss, logits=logits)

    def _seq_sharded_loss(self, h, labels, shift_labels, comm, given_shifted: bool):
        """Loss when the sequence is sharded over an ``sp`` / ``cp`` group (each rank holds a slice of the tokens).

        ``shift_labels`` supplied by the caller (dataloader adapter / ``maybe_context_parallel``) were shifted
        before the sequence was cut.  Plain ``labels`` are shifted here, fetching the label that follows this
        rank's last token from the rank holding the next slice."""
        from ..parallel.ulysses import sp_loss

        if not given_shifted:
            shift_labels = comm.shift_labels_across_ranks(labels)
        lab = shift_labels.reshape(-1)
        local = F.fused_linear_cross_entropy(h.reshape(-1, h.shape[-1]), self.lm_head.weight, lab, -100,
                                             self.config.loss_chunk_tokens)
        return CausalLMOutput(loss=sp_loss(local, (lab != -100).sum(), comm))

    def _tp_loss(self, h_shard, shift_labels, tp):
        """Loss over this rank's token shard, combined over the TP group into the global token mean."""
        if shift_labels is None:
            from ..parallel.tensor_parallel import gather_tokens

            return CausalLMOutput(loss=None, logits=self.lm_head(gather_tokens(h_shard, tp)))
        lab = shift_labels.reshape(-1)
        t = lab.numel() // tp.world
        lab = lab[tp.rank * t : (tp.rank + 1) * t]
        local = F.fused_linear_cross_entropy(h_shard, self.lm_head.weight, lab, -100, self.config.loss_chunk_tokens)
        cnt = (lab != -100).sum().to(torch.float32)
        tot = tp.all_reduce_(cnt.clone())
        part = local * (cnt / tot.clamp(min=1.0))
        whole = tp.all_reduce_(part.detach().clone())
        # value = global mean; gradient flows through the local part only (the other ranks own theirs)
        return CausalLMOutput(loss=part + (whole - part.detach()))

    # ------------------------------------------------------------------ pipeline-parallel inference
    def pipeline_blocks(self):
        """The model as an ordered list of blocks for ``prepare_pippy`` (inference.py): token embedding, one block per decoder
        layer, then final norm + LM head.  What travels between blocks is ``(x, residual, meta)`` with ``meta = [B, S]``, so a
        stage that starts in the middle can rebuild positions and the RoPE tables on its own device."""
        from ..inference import Block

        core = self.model

        def embed(input_ids, **_):
            B, S = input_ids.shape
            x = F.embedding(input_ids, core.embed_tokens.weight).view(B * S, -1)
            meta = torch.tensor([B, S], dtype=torch.int64, device=x.device)
            return x, torch.zeros_like(x), meta

        def layer_fn(layer, first):
            def run(x, residual, meta):
                B, S = int(meta[0]), int(meta[1])
                pos = torch.arange(S, device=x.device).unsqueeze(0).expand(B, S)
                rope = core.rope(pos, S)
                y, res = layer(x, None if first else residual, rope, B, S)
                return y, res, meta

            return run

        def head(x, residual, meta):
            B, S = int(meta[0]), int(meta[1])
            h, _ = core.norm(x, residual)
            return (self.lm_head(h).view(B, S, -1),)

        blocks = [Block("model.embed_tokens", embed, core.embed_tokens)]
        for i, layer in enumerate(core.layers):
            blocks.append(Block(f"model.layers.{i}", layer_fn(layer, i == 0), layer))
        blocks.append(Block("model.norm", head, [core.norm, self.lm_head]))
        return blocks

    # ------------------------------------------------------------------ transformers interchange
    def hf_state_dict(self) -> dict:
        """State dict under transformers' parameter names (q/k/v and gate/up un-fused)."""
        c = self.config
        nq, nk = c.num_attention_heads * c.head_dim, c.num_key_value_heads * c.head_dim
        out = {"model.embed_tokens.weight": self.model.embed_tokens.weight, "model.norm.weight": self.model.norm.weight}
        if not c.tie_word_embeddings:
            out["lm_head.weight"] = self.lm_head.weight
        for i, l in enumerate(self.model.layers):
            pre = f"model.layers.{i}."
            w = l.self_attn.qkv_proj.weight
            out[pre + "self_attn.q_proj.weight"] = w[:nq]
            out[pre + "self_attn.k_proj.weight"] = w[nq : nq + nk]
            out[pre + "self_attn.v_proj.weight"] = w[nq + nk :]
            out[pre + "self_attn.o_proj.weight"] = l.self_attn.o_proj.weight
            gu = l.mlp.gate_up_proj.weight
            out[pre + "mlp.gate_proj.weight"] = gu[: c.intermediate_size]
            out[pre + "mlp.up_proj.weight"] = gu[c.intermediate_size :]
            out[pre + "mlp.down_proj.weight"] = l.mlp.down_proj.weight
            out[pre + "input_layernorm.weight"] = l.input_layernorm.weight
            out[pre + "post_attention_layernorm.weight"] = l.post_attention_layernorm.weight
        return {k: v.detach() for k, v in out.items()}

    @classmethod
    def from_transformers(cls, hf_model, dtype: Optional[torch.dtype] = None, device=None) -> "LlamaForCausalLM":
        """The native counterpart of a ``transformers`` ``LlamaForCausalLM`` (recognised by its config, not by import), with its
        weights: what ``Accelerator.prepare`` builds when tensor / sequence / context parallelism is asked for a transformers Llama.
        Refuses variants the native kernels do not implement (attention / MLP biases, scaled RoPE, sliding windows)."""
        hc = hf_model.config
        if getattr(hc, "model_type", None) != "llama":
            raise NotImplementedError(f"only Llama-architecture models can be adopted, got model_type={getattr(hc, 'model_type', None)!r}")
        unsupported = []
        if getattr(hc, "attention_bias", False) or getattr(hc, "mlp_bias", False):
            unsupported.append("attention_bias / mlp_bias")
        scaling = getattr(hc, "rope_scaling", None) or getattr(hc, "rope_parameters", None)
        if isinstance(scaling, dict) and scaling.get("rope_type", scaling.get("type", "default")) not in ("default", None):
            unsupported.append(f"rope scaling {scaling.get('rope_type', scaling.get('type'))!r}")
        if getattr(hc, "sliding_window", None):
            unsupported.append("sliding_window")
        if unsupported:
            raise NotImplementedError("the native Llama does not implement: " + ", ".join(unsupported))
        head_dim = getattr(hc, "head_dim", None) or hc.hidden_size // hc.num_attention_heads
        theta = getattr(hc, "rope_theta", None)
        if theta is None and isinstance(scaling, dict):
            theta = scaling.get("rope_theta")
        cfg = LlamaConfig(vocab_size=hc.vocab_size, hidden_size=hc.hidden_size, intermediate_size=hc.intermediate_size,
                          num_hidden_layers=hc.num_hidden_layers, num_attention_heads=hc.num_attention_heads,
                          num_key_value_heads=getattr(hc, "num_key_value_heads", None) or hc.num_attention_heads, head_dim=head_dim,
                          rope_theta=float(theta or 10000.0), rms_norm_eps=hc.rms_norm_eps,
                          max_position_embeddings=hc.max_position_embeddings, tie_word_embeddings=bool(getattr(hc, "tie_word_embeddings", False)))
        first = next(hf_model.parameters())
        native = cls(cfg, dtype=dtype or first.dtype, device=device if device is not None else first.device)
        sd = dict(hf_model.state_dict())
        if cfg.tie_word_embeddings:
            sd.pop("lm_head.weight", None)
        native.load_hf_state_dict(sd)
        if cfg.tie_word_embeddings:
            native.tie_weights()
        native.train(hf_model.training)
        return native

    @torch.no_grad()
    def load_hf_state_dict(self, sd: dict):
        mine = self.hf_state_dict()
        missing = [k for k in mine if k not in sd]
        if missing:
            raise KeyError(f"missing keys in transformers state dict: {missing[:5]}...")
        for k, dst in mine.items():
            dst.copy_(sd[k].to(dst.dtype))
        return self
