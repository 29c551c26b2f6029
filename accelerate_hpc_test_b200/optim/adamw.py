"""FusedAdamW: AdamW with fp32 master weights and moments updated by one sm_100a kernel per tensor.

The bf16 parameters the model computes with are "shadows" rewritten by the same kernel, the
gradient clip coefficient and the GradScaler ``found_inf`` flag are read from device memory inside
the kernel, and no step of the update synchronises with the host.

Reference parity: the reference uses ``torch.optim.AdamW`` wrapped by ``AcceleratedOptimizer``
(reference src/accelerate/optimizer.py:38-213) and clips with ``torch.nn.utils.clip_grad_norm_``
(reference src/accelerate/accelerator.py:2951-2979).  The ``state_dict`` layout keeps torch's
``exp_avg`` / ``exp_avg_sq`` / ``step`` keys (plus ``master``) so ``optimizer.bin`` stays loadable.
"""

from __future__ import annotations

from typing import Iterable, Optional

import torch

from ..ops import functional as F


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, master_weights: bool = True):
        if lr < 0 or eps < 0 or not 0 <= betas[0] < 1 or not 0 <= betas[1] < 1:
            raise ValueError("invalid AdamW hyper-parameters")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.master_weights = master_weights
        self._clip_coef: Optional[torch.Tensor] = None
        self._found_inf: Optional[torch.Tensor] = None
        self.grad_scale: float = 1.0

    # -- hooks used by Accelerator.clip_grad_norm_ / GradScaler integration --------------------
    def set_clip_coef(self, coef: Optional[torch.Tensor]):
        """Device scalar multiplied into every gradient inside the update kernel (consumed by the next step)."""
        self._clip_coef = coef

    def set_found_inf(self, flag: Optional[torch.Tensor]):
        self._found_inf = flag

    def _init_state(self, p: torch.Tensor):
        st = self.state[p]
        st["step"] = 0
        st["exp_avg"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
        st["exp_avg_sq"] = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
        if p.dtype != torch.float32 and self.master_weights:
            st["master"] = p.detach().to(torch.float32).clone()
        return st

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                g = p.grad
                if g is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st = self._init_state(p)
                st["step"] += 1
                if not g.is_contiguous():
                    g = g.contiguous()
                if p.dtype == torch.float32:
                    master, shadow = p.data, None
                elif "master" in st:
                    master, shadow = st["master"], p.data
                else:  # no master copy requested: update an fp32 staging copy then write back
                    master, shadow = p.data.float(), p.data
                if g.dtype not in (torch.float32, torch.bfloat16):
                    g = g.float()
                F.adamw_flat_(master.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1), g.view(-1),
                              shadow.view(-1) if shadow is not None else None,
                              lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"], weight_decay=group["weight_decay"],
                              step=st["step"], grad_scale=self.grad_scale, clip_coef=self._clip_coef,
                              found_inf=self._found_inf)
        self._clip_coef = None
        return loss


def grad_global_norm(params: Iterable[torch.Tensor], max_norm: Optional[float] = None):
    """Global L2 norm of the gradients on the device (no host sync).

    Returns ``(norm, clip_coef, found_inf)`` as 1-element device tensors; ``clip_coef`` is
    ``min(1, max_norm / (norm + 1e-6))`` when ``max_norm`` is given.
    """
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        z = torch.zeros(1)
        return z, torch.ones(1), z.clone()
    dev = grads[0].device
    acc = torch.zeros(2, dtype=torch.float32, device=dev)
    for g in grads:
        F.sumsq_(g, acc)
    norm = torch.empty(1, dtype=torch.float32, device=dev)
    coef = torch.ones(1, dtype=torch.float32, device=dev)
    F.clip_coef_(acc, float(max_norm) if max_norm is not None else float("inf"), norm, coef)
    return norm, coef, acc[1:2]
