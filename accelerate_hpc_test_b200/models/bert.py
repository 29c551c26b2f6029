"""BERT encoder for sequence classification (BASELINE.json config 1: bert-base-cased on MRPC-shaped
pairs, CPU / gloo, world_size 2 — the shape of reference examples/nlp_example.py:99-147).

Hard-coded configuration (no hub access): 12 layers, hidden 768, 12 heads, FFN 3072, vocab 28996,
2 labels.  Linear layers go through :func:`..ops.functional.linear`, i.e. the tcgen05 GEMM (with the
bias epilogue) on CUDA and the fp32 oracle on CPU; attention / LayerNorm / GELU use PyTorch — this
model family is the framework-level oracle workload, not a hot path.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as F


@dataclass
class BertConfig:
    vocab_size: int = 28996
    hidden_size: int = 768
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    intermediate_size: int = 3072
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    layer_norm_eps: float = 1e-12
    hidden_dropout_prob: float = 0.1
    attention_probs_dropout_prob: float = 0.1
    num_labels: int = 2
    initializer_range: float = 0.02
    model_type: str = "bert"

    @classmethod
    def base_cased(cls, **kw):
        return cls(**kw)

    @classmethod
    def tiny(cls, **kw):
        base = dict(vocab_size=1000, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128,
                    max_position_embeddings=128)
        base.update(kw)
        return cls(**base)


@dataclass
class SequenceClassifierOutput:
    loss: Optional[torch.Tensor] = None
    logits: Optional[torch.Tensor] = None

    def __getitem__(self, k):
        if isinstance(k, str):
            return getattr(self, k)
        return tuple(v for v in (self.loss, self.logits) if v is not None)[k]


class _Linear(nn.Module):
    def __init__(self, i, o, dtype=torch.float32):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(o, i, dtype=dtype))
        self.bias = nn.Parameter(torch.zeros(o, dtype=dtype))

    def forward(self, x):
        if x.is_cuda and x.dtype != torch.bfloat16:
            return torch.nn.functional.linear(x, self.weight, self.bias)
        return F.linear(x, self.weight, self.bias)


class BertLayer(nn.Module):
    """Post-norm encoder block: fused QKV projection, scaled-dot-product attention with a key padding mask, GELU feed-forward."""

    def __init__(self, c: BertConfig, dtype):
        super().__init__()
        self.c = c
        self.qkv = _Linear(c.hidden_size, 3 * c.hidden_size, dtype)
        self.attn_out = _Linear(c.hidden_size, c.hidden_size, dtype)
        self.attn_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps, dtype=dtype)
        self.ffn_in = _Linear(c.hidden_size, c.intermediate_size, dtype)
        self.ffn_out = _Linear(c.intermediate_size, c.hidden_size, dtype)
        self.ffn_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps, dtype=dtype)
        self.drop = nn.Dropout(c.hidden_dropout_prob)

    def forward(self, x, mask):
        B, S, H = x.shape
        nh = self.c.num_attention_heads
        q, k, v = self.qkv(x).view(B, S, 3, nh, H // nh).permute(2, 0, 3, 1, 4)
        a = torch.nn.functional.scaled_dot_product_attention(
            q, k, v, attn_mask=mask, dropout_p=self.c.attention_probs_dropout_prob if self.training else 0.0)
        a = a.transpose(1, 2).reshape(B, S, H)
        x = self.attn_norm(x + self.drop(self.attn_out(a)))
        h = self.ffn_out(torch.nn.functional.gelu(self.ffn_in(x)))
        return self.ffn_norm(x + self.drop(h))


class BertModel(nn.Module):
    """Word + position + token-type embeddings, the encoder stack and the ``[CLS]`` pooler."""

    def __init__(self, c: BertConfig, dtype=torch.float32):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.hidden_size, dtype=dtype)
        self.position_embeddings = nn.Embedding(c.max_position_embeddings, c.hidden_size, dtype=dtype)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, c.hidden_size, dtype=dtype)
        self.emb_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps, dtype=dtype)
        self.drop = nn.Dropout(c.hidden_dropout_prob)
        self.layers = nn.ModuleList([BertLayer(c, dtype) for _ in range(c.num_hidden_layers)])
        self.pooler = _Linear(c.hidden_size, c.hidden_size, dtype)

    def forward(self, input_ids, attention_mask=None, token_type_ids=None):
        B, S = input_ids.shape
        pos = torch.arange(S, device=input_ids.device)[None]
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        x = self.word_embeddings(input_ids) + self.position_embeddings(pos) + self.token_type_embeddings(token_type_ids)
        x = self.drop(self.emb_norm(x))
        mask = None
        if attention_mask is not None:
            mask = attention_mask[:, None, None, :].to(torch.bool)
        for layer in self.layers:
            x = layer(x, mask)
        return x, torch.tanh(self.pooler(x[:, 0]))


class BertForSequenceClassification(nn.Module):
    """BERT encoder with a classification head on the pooled ``[CLS]`` state (the MRPC-style example workload)."""

    _no_split_modules = ["BertLayer"]

    def __init__(self, config: BertConfig, dtype=torch.float32):
        super().__init__()
        self.config = config
        self.bert = BertModel(config, dtype)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = _Linear(config.hidden_size, config.num_labels, dtype)
        self.reset_parameters()

    @torch.no_grad()
    def reset_parameters(self):
        std = self.config.initializer_range
        for m in self.modules():
            if isinstance(m, _Linear):
                m.weight.normal_(0.0, std)
                m.bias.zero_()
            elif isinstance(m, nn.Embedding):
                m.weight.normal_(0.0, std)
            elif isinstance(m, nn.LayerNorm):
                m.weight.fill_(1.0)
                m.bias.zero_()

    def forward(self, input_ids, attention_mask=None, token_type_ids=None, labels=None, **_):
        _, pooled = self.bert(input_ids, attention_mask, token_type_ids)
        logits = self.classifier(self.dropout(pooled))
        loss = None
        if labels is not None:
            loss = torch.nn.functional.cross_entropy(logits.float().view(-1, self.config.num_labels), labels.view(-1))
        return SequenceClassifierOutput(loss=loss, logits=logits)


    # ------------------------------------------------------------------ pipeline-parallel inference
    def pipeline_blocks(self):
        """The model as an ordered list of blocks for ``prepare_pippy``: embeddings, one block per encoder layer, pooler +
        classifier.  ``(hidden states, key mask)`` travel between blocks; the mask goes as a float tensor (1 = attend) so that
        it can cross a process boundary like any activation."""
        from ..inference import Block

        core = self.bert

        def embed(input_ids, attention_mask=None, token_type_ids=None, **_):
            B, S = input_ids.shape
            positions = torch.arange(S, device=input_ids.device)[None]
            if token_type_ids is None:
                token_type_ids = torch.zeros_like(input_ids)
            x = core.word_embeddings(input_ids) + core.position_embeddings(positions) + core.token_type_embeddings(token_type_ids)
            keep = torch.ones(B, S, device=x.device) if attention_mask is None else attention_mask.to(x.device)
            return core.drop(core.emb_norm(x)), keep.to(x.dtype)

        def encoder(layer):
            def run(x, keep):
                return layer(x, keep[:, None, None, :] > 0.5), keep

            return run

        def head(x, keep):
            pooled = torch.tanh(core.pooler(x[:, 0]))
            return (self.classifier(self.dropout(pooled)),)

        blocks = [Block("bert.word_embeddings", embed, [core.word_embeddings, core.position_embeddings, core.token_type_embeddings, core.emb_norm])]
        blocks += [Block(f"bert.layers.{i}", encoder(layer), layer) for i, layer in enumerate(core.layers)]
        blocks.append(Block("bert.pooler", head, [core.pooler, self.classifier]))
        return blocks


def synthetic_mrpc(n: int, vocab_size: int = 28996, max_len: int = 128, seed: int = 0):
    """MRPC-shaped synthetic sentence pairs: variable lengths, ``[CLS] a [SEP] b [SEP]``, binary labels.

    Returns a list of dicts with ``input_ids`` / ``token_type_ids`` / ``attention_mask`` / ``labels``
    (un-padded; use :func:`pad_collate`).  3668 train / 408 validation pairs in the real task.
    """
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        la = int(torch.randint(8, max_len // 2 - 2, (1,), generator=g))
        lb = int(torch.randint(8, max_len // 2 - 2, (1,), generator=g))
        ids = torch.randint(min(1000, vocab_size // 4), vocab_size, (la + lb + 3,), generator=g)
        ids[0], ids[la + 1], ids[-1] = 101, 102, 102
        tt = torch.cat([torch.zeros(la + 2, dtype=torch.long), torch.ones(lb + 1, dtype=torch.long)])
        label = int(torch.randint(0, 2, (1,), generator=g))
        out.append({"input_ids": ids, "token_type_ids": tt, "attention_mask": torch.ones_like(ids), "labels": torch.tensor(label)})
    return out


def pad_collate(examples):
    """Pad to the longest sequence of the batch (the reference example's ``pad_to_longest`` collate)."""
    m = max(e["input_ids"].numel() for e in examples)
    batch = {}
    for k in ("input_ids", "token_type_ids", "attention_mask"):
        batch[k] = torch.stack([torch.nn.functional.pad(e[k], (0, m - e[k].numel())) for e in examples])
    batch["labels"] = torch.stack([e["labels"] for e in examples])
    return batch
