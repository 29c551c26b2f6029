"""The tiny workloads every distributed parity test trains: a noisy line ``y = a x + b`` and the model that fits it.

Two scalar parameters make "did the gradients synchronise / did the optimizer step" checks exact and cheap; the
MRPC-shaped loaders feed the example scripts when no dataset can be downloaded.  (Counterpart of the reference's
``accelerate.test_utils.training``.)
"""

from __future__ import annotations

import numpy as np
import torch
from torch.utils.data import DataLoader


class RegressionDataset:
    """``length`` samples ``{"x": x_i, "y": a * x_i + b + noise}`` (float32 scalars), reproducible through ``seed``."""

    def __init__(self, a: float = 2, b: float = 3, length: int = 64, seed=None):
        draw = np.random.default_rng(seed)
        self.length = int(length)
        self.x = draw.normal(size=self.length).astype(np.float32)
        noise = draw.normal(scale=0.1, size=self.length).astype(np.float32)
        self.y = (a * self.x + b + noise).astype(np.float32)

    def __len__(self) -> int:
        return self.length

    def __getitem__(self, index):
        return {"x": self.x[index], "y": self.y[index]}


class RegressionModel(torch.nn.Module):
    """``forward(x) = a * x + b`` with scalar ``a`` and ``b`` (both start at the given values)."""

    def __init__(self, a: float = 0, b: float = 0, double_output: bool = False):
        super().__init__()
        self.a = torch.nn.Parameter(torch.tensor(float(a)))
        self.b = torch.nn.Parameter(torch.tensor(float(b)))
        self.double_output = double_output
        self.first_batch = True

    def forward(self, x=None):
        self.first_batch = False
        y = self.a * x + self.b
        return (y, y) if self.double_output else y


def mocked_dataloaders(accelerator, batch_size: int = 16):
    """MRPC-shaped synthetic train / eval dataloaders (no hub access in this environment)."""
    from ..models.bert import pad_collate, synthetic_mrpc

    def loader(n, seed, shuffle):
        data = synthetic_mrpc(n, vocab_size=1000, max_len=32, seed=seed)
        return DataLoader(data, shuffle=shuffle, collate_fn=pad_collate, batch_size=batch_size)

    return loader(64, 0, True), loader(32, 1, False)


def mocked_dataloaders_for_autoregressive_models(accelerator, batch_size: int = 16):
    """Train / eval dataloaders for a causal language model: fixed-length token sequences whose labels are the inputs themselves
    (the model shifts them), padded positions masked out of the loss.  Synthetic and reproducible — the reference tokenises the MRPC
    sentence pairs with a hub tokenizer, which cannot be fetched here."""

    def loader(count, seed, shuffle, length=32, vocab=1000):
        g = torch.Generator().manual_seed(seed)
        rows = []
        for _ in range(count):
            used = int(torch.randint(length // 2, length + 1, (1,), generator=g))
            ids = torch.zeros(length, dtype=torch.long)
            ids[:used] = torch.randint(1, vocab, (used,), generator=g)
            mask = (torch.arange(length) < used).long()
            labels = ids.clone()
            labels[used:] = -100
            rows.append({"input_ids": ids, "attention_mask": mask, "labels": labels})
        collate = lambda batch: {key: torch.stack([row[key] for row in batch]) for key in batch[0]}  # noqa: E731
        return DataLoader(rows, shuffle=shuffle, collate_fn=collate, batch_size=batch_size)

    return loader(64, 0, True), loader(32, 1, False)
