"""``LocalSGD``: skip gradient synchronisation and average the *parameters* every ``local_sgd_steps`` steps
(reference src/accelerate/local_sgd.py:41-106)."""

import torch

from .accelerator import Accelerator
from .utils.dataclasses import DistributedType


class LocalSGD:
    """Context manager; call ``step()`` after each optimizer step.

        with LocalSGD(accelerator=accelerator, model=model, local_sgd_steps=8) as local_sgd:
            for batch in loader:
                ...; optimizer.step(); local_sgd.step()
    """

    def __enter__(self):
        if self.enabled:
            self.model_sync_obj = self.model.no_sync()
            self.model_sync_obj.__enter__()
        return self

    def __exit__(self, type, value, tb):
        if self.enabled:
            self._sync_and_avg_model_params()
            self.model_sync_obj.__exit__(type, value, tb)

    def __init__(self, accelerator: Accelerator, model: torch.nn.Module, local_sgd_steps: int, enabled: bool = True):
        if accelerator.distributed_type not in [DistributedType.NO, DistributedType.MULTI_CPU, DistributedType.MULTI_GPU]:
            raise NotImplementedError("LocalSGD is supported only for CPUs and GPUs (no DeepSpeed or MegatronLM)")
        self.enabled = enabled and accelerator.distributed_type != DistributedType.NO
        self.num_steps = 0
        if self.enabled:
            self.accelerator = accelerator
            self.model = model
            self.local_sgd_steps = local_sgd_steps

    def step(self):
        self.num_steps += 1
        if not self.enabled:
            return
        if self.num_steps % self.local_sgd_steps == 0:
            self._sync_and_avg_model_params()

    def _sync_and_avg_model_params(self):
        """Average every parameter over the ranks: one flat all-reduce per dtype instead of one per tensor."""
        self.accelerator.wait_for_everyone()
        with self.accelerator.autocast():
            by_dtype = {}
            for param in self.model.parameters():
                by_dtype.setdefault(param.dtype, []).append(param)
            for params in by_dtype.values():
                flat = torch.cat([p.data.reshape(-1) for p in params])
                flat = self.accelerator.reduce(flat, reduction="mean")
                off = 0
                for p in params:
                    n = p.numel()
                    p.data.copy_(flat[off : off + n].view_as(p.data))
                    off += n
