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

import os
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
        # ---- "optimizer in backward" (see enable_update_in_backward)
        self._bwd_mode = "off"          # off | probing | active
        self._bwd_stream = None
        self._bwd_hooks = []
        self._bwd_fired = {}            # param -> hook firings since the last step (probing)
        self._bwd_updated = set()       # params already updated during this step's backward
        self._saw_grad_consumer = False  # clip / found_inf / grad_scale used: gradients are needed after backward
        self._sync_fn = None
        self._probe_clean_steps = 0

    # -- hooks used by Accelerator.clip_grad_norm_ / GradScaler integration --------------------
    def set_clip_coef(self, coef: Optional[torch.Tensor]):
        """Device scalar multiplied into every gradient inside the update kernel (consumed by the next step)."""
        self._note_grad_consumer("gradient clipping")
        self._clip_coef = coef

    def set_found_inf(self, flag: Optional[torch.Tensor]):
        self._note_grad_consumer("a GradScaler")
        self._found_inf = flag

    # -- optimizer step fused into the backward pass ----------------------------------------------
    def enable_update_in_backward(self, sync_gradients_fn=None, mode: Optional[str] = None):
        """Update every parameter as soon as its gradient is final, on a side stream, while the rest of the backward
        pass still runs (the update is pure HBM traffic, the backward GEMMs are tensor-core bound, so the two overlap
        almost for free; the gradient is released right away).  The single-device counterpart of the sharded engine's
        fused reduce-scatter + AdamW.

        ``mode``: ``"1"`` on from the first step, ``"0"`` never, ``"auto"`` (default, or
        ``ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD``): the first step runs normally and the fusion switches on only if
        nothing needed the gradients after backward (no clipping, no GradScaler, one backward per step).  Once on,
        ``p.grad`` is ``None`` after backward; asking for clipping then raises with a pointer to the switch.
        ``sync_gradients_fn`` returns False on gradient-accumulation micro-steps (no update then)."""
        mode = mode or os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto")
        if mode == "0" or self._bwd_hooks:
            return
        params = [p for g in self.param_groups for p in g["params"] if p.requires_grad]
        if not params or not all(p.is_cuda for p in params):
            return
        self._sync_fn = sync_gradients_fn
        self._bwd_stream = torch.cuda.Stream(device=params[0].device)
        self._group_of = {p: g for g in self.param_groups for p in g["params"]}
        for p in params:
            self._bwd_hooks.append(p.register_post_accumulate_grad_hook(self._on_grad_ready))
        self._bwd_mode = "active" if mode == "1" else "probing"

    def enable_update_after_reduce(self, stream, sync_gradients_fn=None, mode: Optional[str] = None) -> bool:
        """The same fusion under a data-parallel engine (``parallel/ddp.py``): a gradient is final once its bucket has been
        all-reduced, so the engine calls ``update_reduced`` for the bucket's parameters right behind the all-reduce, on its
        communication ``stream`` — the update of bucket *i* then overlaps the backward GEMMs that produce bucket *i+1*.
        Same ``mode`` / probing rules as ``enable_update_in_backward``.  Returns whether the fusion was armed."""
        mode = mode or os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto")
        if mode == "0" or self._bwd_hooks or self._bwd_mode != "off":
            return False
        self._sync_fn = sync_gradients_fn
        self._bwd_stream = stream
        self._group_of = {p: g for g in self.param_groups for p in g["params"]}
        self._bwd_mode = "active" if mode == "1" else "probing"
        return True

    @torch.no_grad()
    def update_reduced(self, p: torch.Tensor, reduced_grad: torch.Tensor) -> bool:
        """Engine hook: ``reduced_grad`` (a view of the reduced bucket) is final.  Must be called with the engine's stream current,
        after the all-reduce was enqueued on it.  True when the parameter was updated (its gradient is then spent)."""
        if self._bwd_mode == "probing":
            self._bwd_fired[p] = self._bwd_fired.get(p, 0) + 1
            return False
        if self._bwd_mode != "active" or p not in self._group_of:
            return False
        if self._sync_fn is not None and not self._sync_fn():
            return False
        if p in self._bwd_updated:
            raise RuntimeError("a parameter's bucket was reduced twice in one step (several synchronised backward passes per step): "
                               "set ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD=0.")
        self._update_one(p, reduced_grad, self._group_of[p])
        self._bwd_updated.add(p)
        return True

    def _note_grad_consumer(self, what: str):
        self._saw_grad_consumer = True
        if self._bwd_mode == "active" and self._bwd_updated:
            raise RuntimeError(
                f"{what} was requested after the optimizer update already ran inside backward. "
                "Set ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD=0 when gradients are needed after backward on some steps only."
            )
        if self._bwd_mode == "probing":
            self._bwd_mode = "off"

    @torch.no_grad()
    def _on_grad_ready(self, p: torch.Tensor):
        if self._bwd_mode == "probing":
            self._bwd_fired[p] = self._bwd_fired.get(p, 0) + 1
            return
        if self._bwd_mode != "active" or p.grad is None:
            return
        if self._sync_fn is not None and not self._sync_fn():
            return  # accumulation micro-step: keep summing into p.grad
        if p in self._bwd_updated:
            raise RuntimeError("a parameter received a second gradient after it was already updated inside backward "
                               "(several backward passes per step): set ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD=0.")
        g = p.grad
        cur = torch.cuda.current_stream(p.device)
        self._bwd_stream.wait_stream(cur)
        with torch.cuda.stream(self._bwd_stream):
            self._update_one(p, g, self._group_of[p])
        g.record_stream(self._bwd_stream)
        p.grad = None
        self._bwd_updated.add(p)

    def _update_one(self, p, g, group):
        st = self.state[p]
        if len(st) == 0:
            st = self._init_state(p)
        st["step"] += 1
        b1, b2 = group["betas"]
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
            for p in group["params"]:
                if p.grad is None or p in self._bwd_updated:
                    continue
                self._update_one(p, p.grad, group)
        if self._bwd_updated:
            # the next forward reads the parameters the side stream has just written
            if self._bwd_stream is not None:
                torch.cuda.current_stream().wait_stream(self._bwd_stream)
            self._bwd_updated.clear()
        if self._bwd_mode == "probing":
            # switch on after two steps in a row that each followed exactly one backward and needed no gradient afterwards; a
            # loop that accumulates by hand (several backward passes per step) keeps the ordinary update
            clean = (not self._saw_grad_consumer and self.grad_scale == 1.0 and bool(self._bwd_fired)
                     and all(n == 1 for n in self._bwd_fired.values()))
            self._probe_clean_steps = self._probe_clean_steps + 1 if clean else 0
            if self._saw_grad_consumer or self.grad_scale != 1.0:
                self._bwd_mode = "off"
            elif self._probe_clean_steps >= 2:
                self._bwd_mode = "active"
            self._bwd_fired.clear()
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
