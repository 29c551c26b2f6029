"""Tiny regression workload used by the distributed parity tests (reference src/accelerate/test_utils/training.py:22-47)."""

import numpy as np
import torch
from torch.utils.data import DataLoader


class RegressionDataset:
    def __init__(self, a=2, b=3, length=64, seed=None):
        rng = np.random.default_rng(seed)
        self.length = length
        self.x = rng.normal(size=(length,)).astype(np.float32)
        self.y = a * self.x + b + rng.normal(scale=0.1, size=(length,)).astype(np.float32)

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        return {"x": self.x[i], "y": self.y[i]}


class RegressionModel(torch.nn.Module):
    """``y = a * x + b`` with scalar parameters."""

    def __init__(self, a=0, b=0, double_output=False):
        super().__init__()
        self.a = torch.nn.Parameter(torch.tensor(a).float())
        self.b = torch.nn.Parameter(torch.tensor(b).float())
        self.first_batch = True

    def forward(self, x=None):
        if self.first_batch:
            self.first_batch = False
        return x * self.a + self.b


class RegressionModel4XPU(RegressionModel):
    pass


def mocked_dataloaders(accelerator, batch_size: int = 16):
    """MRPC-shaped synthetic train / eval dataloaders (no hub access in this environment)."""
    from ..models.bert import pad_collate, synthetic_mrpc

    train = synthetic_mrpc(64, vocab_size=1000, max_len=32, seed=0)
    val = synthetic_mrpc(32, vocab_size=1000, max_len=32, seed=1)
    return (DataLoader(train, shuffle=True, collate_fn=pad_collate, batch_size=batch_size),
            DataLoader(val, shuffle=False, collate_fn=pad_collate, batch_size=batch_size))
