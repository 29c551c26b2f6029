"""Tensor parallelism with sequence-parallel activations (Megatron style), on fused kernels.

The reference expects the model to arrive already TP-sharded by transformers (``tp_plan="auto"``) and
only wraps leftovers as replicated DTensors (reference src/accelerate/accelerator.py:1579-1639); the
collectives are DTensor/NCCL.  Here TP is native:

* activations between blocks are sharded along tokens: ``[T/tp, H]`` per rank;
* ``ColumnParallelLinear`` (QKV, gate/up): **all-gather -> GEMM in one kernel** — warp 3 of every
  GEMM CTA pulls the peers' activation shards with ``cp.async.bulk`` while the tensor cores already
  work on the local rows; per-shard counters gate the TMA producer (``b200_gemm_bf16_allgather_a``);
* ``RowParallelLinear`` (attention out, MLP down): **GEMM -> reduce-scatter** — the epilogue stores
  each partial tile straight into the owning rank's slot buffer over NVLink
  (``b200_gemm_bf16_scatter``), a barrier and a slot reduction finish it;
* the backward passes are the same two primitives with the roles swapped.

``TorchTensorParallelConfig.enable_async_tp`` selects the fused kernels (default on CUDA); the
un-fused path (separate all-gather / reduce-scatter kernels, or c10d on CPU) is the oracle.
"""

from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch
import torch.nn as nn

from ..ops import _lib
from ..ops import functional as F
from ..ops._lib import check, ptr

BF16 = torch.bfloat16


# SM-pair (cta_group::2) kernels by default; ACCELERATE_B200_GEMM_2CTA=0 selects the single-CTA versions
_TWO_CTA = os.environ.get("ACCELERATE_B200_GEMM_2CTA", "1") != "0"
_AG_GEMM = "b200_gemm2_bf16_allgather_a" if _TWO_CTA else "b200_gemm_bf16_allgather_a"
_GEMM_RS = "b200_gemm2_bf16_scatter" if _TWO_CTA else "b200_gemm_bf16_scatter"


class TPComm:
    """Communication context of one tensor-parallel group (symmetric buffers, channels, epochs)."""

    def __init__(self, group, device, hidden: int, max_tokens: int, fused: bool = True):
        self.group = group
        self.world = torch.distributed.get_world_size(group)
        self.rank = torch.distributed.get_rank(group)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.fused = fused and self.cuda
        self.hidden = hidden
        self.max_tokens = max_tokens
        # ACCELERATE_B200_TP_KEEP_GATHERED=0 trades the per-layer copy of the gathered input (tokens x hidden x 2 bytes per
        # column-parallel layer) for a second gather in backward
        self.keep_gathered = os.environ.get("ACCELERATE_B200_TP_KEEP_GATHERED", "1") != "0"
        if self.cuda:
            from .symm import SymmetricWorkspace

            per = max_tokens * hidden * 2
            self.ws = SymmetricWorkspace(group, 4 * per + 16 * 2**20, self.device)
            self.ag_off = [self.ws.alloc(per) for _ in range(2)]  # gathered activations, double-buffered
            self.rs_off = [self.ws.alloc(per) for _ in range(2)]  # [world][T/world][N] slot buffers
            self.ready = [torch.zeros(self.world, dtype=torch.int64, device=self.device) for _ in range(2)]
            self.ready_epoch = [0, 0]
            self.ch_ag = self.ws.new_channel()
            self.ch_rs = self.ws.new_channel()
            self.ch_a2a = (self.ws.new_channel(), self.ws.new_channel())
            self.flip_ag = 0
            self.flip_rs = 0

    def private_copy(self, gathered: torch.Tensor) -> torch.Tensor:
        """A copy of gathered rows that outlives the rotating gather buffer (a no-op when nothing was gathered into one)."""
        if self.world == 1 or not self.cuda:
            return gathered
        return gathered.clone()

    # ------------------------------------------------------------------ un-fused primitives
    def all_gather_rows(self, x: torch.Tensor) -> torch.Tensor:
        """[t, H] per rank -> [t * world, H] (rank order)."""
        if self.world == 1:
            return x
        if not self.cuda:
            parts = [torch.empty_like(x) for _ in range(self.world)]
            torch.distributed.all_gather(parts, x.contiguous(), group=self.group)
            return torch.cat(parts, dim=0)
        full, _ = self._stage_and_gather(x)
        return full

    def _stage_and_gather(self, x: torch.Tensor):
        """Copy my shard into my slot of the symmetric gather buffer, barrier, pull the peers' slots."""
        t, h = x.shape
        buf = self.flip_ag
        self.flip_ag ^= 1
        off = self.ag_off[buf]
        full = self.ws.tensor(off, (t * self.world, h), BF16)
        full[self.rank * t : (self.rank + 1) * t].copy_(x)
        self.ws.barrier(self.ch_ag)
        shard_bytes = t * h * 2
        from .symm import _stream

        for s in range(1, self.world):
            p = (self.rank + s) % self.world
            # pull peer p's own slot (it sits at the same offset in its buffer)
            check(self.ws.lib.b200_memcpy_async(ctypes.c_void_p(self.ws.base + off + p * shard_bytes),
                                                ctypes.c_void_p(self.ws.peer_bases[p] + off + p * shard_bytes), shard_bytes, _stream(None)))
        return full, buf

    def reduce_scatter_rows(self, y: torch.Tensor) -> torch.Tensor:
        """[t * world, N] partial sums per rank -> [t, N] reduced rows owned by this rank."""
        if self.world == 1:
            return y
        T, n = y.shape
        t = T // self.world
        if not self.cuda:
            full = y.float().contiguous()
            torch.distributed.all_reduce(full, group=self.group)
            return full[self.rank * t : (self.rank + 1) * t].to(y.dtype)
        buf = self.flip_rs
        self.flip_rs ^= 1
        off = self.rs_off[buf]
        from .symm import _stream

        y = y.contiguous()
        for s in range(self.world):
            p = (self.rank + s) % self.world
            dst = self.ws.peer_bases[p] + off + self.rank * t * n * 2
            check(self.ws.lib.b200_memcpy_async(ctypes.c_void_p(dst), ctypes.c_void_p(y.data_ptr() + p * t * n * 2), t * n * 2, _stream(None)))
        self.ws.barrier(self.ch_rs)
        out = torch.empty((t, n), dtype=y.dtype, device=y.device)
        check(self.ws.lib.b200_reduce_slots(ctypes.c_void_p(self.ws.base + off), ptr(out), self.world, t * n, _lib.cur_stream()))
        return out

    # ------------------------------------------------------------------ fused primitives
    def ag_gemm(self, x: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False):
        """``all_gather_rows(x) @ W^T`` (or ``@ W`` when ``b_mn_major``). Returns ``(y, x_full)``."""
        if not self.fused or self.world == 1:
            full = self.all_gather_rows(x)
            return F.gemm(full, w, b_mn_major=b_mn_major), full
        t, k = x.shape
        n = w.shape[1] if b_mn_major else w.shape[0]
        buf = self.flip_ag
        self.flip_ag ^= 1
        off = self.ag_off[buf]
        full = self.ws.tensor(off, (t * self.world, k), BF16)
        full[self.rank * t : (self.rank + 1) * t].copy_(x)
        self.ws.barrier(self.ch_ag)  # every rank's own slot is written (and nobody still reads this buffer)
        self.ready_epoch[buf] += 1
        y = torch.empty((t * self.world, n), dtype=BF16, device=x.device)
        check(getattr(self.ws.lib, _AG_GEMM)(
            self.ws.peer_array, self.rank, self.world, off, ctypes.c_void_p(self.ws.base + off), t, ptr(w), ptr(y), n, k,
            w.stride(0), y.stride(0), int(b_mn_major), ptr(self.ready[buf]), self.ready_epoch[buf], _lib.cur_stream()), "gemm_allgather_a")
        return y, full

    def gemm_rs(self, x: torch.Tensor, w: torch.Tensor, b_mn_major: bool = False) -> torch.Tensor:
        """``reduce_scatter_rows(x @ W^T)`` with the scatter done by the GEMM epilogue."""
        if not self.fused or self.world == 1:
            return self.reduce_scatter_rows(F.gemm(x, w, b_mn_major=b_mn_major))
        T, k = x.shape
        n = w.shape[1] if b_mn_major else w.shape[0]
        t = T // self.world
        buf = self.flip_rs
        self.flip_rs ^= 1
        off = self.rs_off[buf]
        x = F._rowmajor2d(x)
        check(getattr(self.ws.lib, _GEMM_RS)(ptr(x), ptr(w), self.ws.peer_array, self.rank, self.world, off, t, n, k, x.stride(0),
                                                 w.stride(0), 0, int(b_mn_major), _lib.cur_stream()), "gemm_scatter")
        self.ws.barrier(self.ch_rs)  # all partial tiles have landed in my slots
        out = torch.empty((t, n), dtype=BF16, device=x.device)
        check(self.ws.lib.b200_reduce_slots(ctypes.c_void_p(self.ws.base + off), ptr(out), self.world, t * n, _lib.cur_stream()))
        return out

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            torch.distributed.all_reduce(t, group=self.group)
        return t


class _ColumnParallelFn(torch.autograd.Function):
    """y[T, N/tp] = all_gather(x)[T, K] @ W_local^T ; backward: dx = reduce_scatter(dy @ W_local), dW = dy^T x_full."""

    @staticmethod
    def forward(ctx, x, weight, comm: TPComm):
        y, x_full = comm.ag_gemm(x, weight)
        # the gathered activations sit in a rotating symmetric buffer: a private copy (one D2D pass) is cheaper than the
        # stage + barrier + peer pulls of gathering them again for the weight gradient
        ctx.kept_full = comm.keep_gathered and weight.requires_grad
        ctx.save_for_backward(comm.private_copy(x_full) if ctx.kept_full else x, weight)
        ctx.comm = comm
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        comm = ctx.comm
        dy = dy.contiguous()
        dx = comm.gemm_rs(dy, weight, b_mn_major=True) if ctx.needs_input_grad[0] else None
        dw = None
        if ctx.needs_input_grad[1]:
            x_full = x if ctx.kept_full else comm.all_gather_rows(x)
            dw = F.deliver_weight_grad(weight, lambda out, acc: F.gemm(dy, x_full, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc))
        return dx, dw, None


class _RowParallelFn(torch.autograd.Function):
    """y[T/tp, N] = reduce_scatter(x[T, K/tp] @ W_local^T) ; backward: dx = all_gather(dy) @ W_local, dW = dy_full^T x."""

    @staticmethod
    def forward(ctx, x, weight, comm: TPComm):
        y = comm.gemm_rs(x, weight)
        ctx.save_for_backward(x, weight)
        ctx.comm = comm
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        comm = ctx.comm
        dx, dy_full = comm.ag_gemm(dy.contiguous(), weight, b_mn_major=True)
        dw = None
        if ctx.needs_input_grad[1]:
            dw = F.deliver_weight_grad(weight, lambda out, acc: F.gemm(dy_full, x, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc))
        return (dx if ctx.needs_input_grad[0] else None), dw, None


class _ScatterRowsFn(torch.autograd.Function):
    """Replicated [T, H] -> my [T/tp, H] slice; backward all-gathers the slices."""

    @staticmethod
    def forward(ctx, x, comm: TPComm):
        t = x.shape[0] // comm.world
        ctx.comm = comm
        return x[comm.rank * t : (comm.rank + 1) * t].contiguous()

    @staticmethod
    def backward(ctx, dy):
        return ctx.comm.all_gather_rows(dy.contiguous()), None


class _GatherRowsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, comm: TPComm):
        ctx.comm = comm
        return comm.all_gather_rows(x.contiguous())

    @staticmethod
    def backward(ctx, dy):
        comm = ctx.comm
        t = dy.shape[0] // comm.world
        return dy[comm.rank * t : (comm.rank + 1) * t].contiguous(), None


class ColumnParallelLinear(nn.Module):
    """Output features sharded over the group; consumes token-sharded input, produces full-token output."""

    def __init__(self, weight_local: torch.Tensor, comm: TPComm):
        super().__init__()
        self.weight = nn.Parameter(weight_local)
        self.comm = comm

    def forward(self, x):
        return _ColumnParallelFn.apply(x, self.weight, self.comm)


class RowParallelLinear(nn.Module):
    """Input features sharded over the group; consumes full-token input, produces token-sharded output."""

    def __init__(self, weight_local: torch.Tensor, comm: TPComm):
        super().__init__()
        self.weight = nn.Parameter(weight_local)
        self.comm = comm

    def forward(self, x):
        return _RowParallelFn.apply(x, self.weight, self.comm)


def _shard_rows(w: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    n = w.shape[0] // world
    return w[rank * n : (rank + 1) * n].detach().clone()


def _shard_cols(w: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    n = w.shape[1] // world
    return w[:, rank * n : (rank + 1) * n].detach().clone().contiguous()


def _mark_replicated(p: nn.Parameter, comm: TPComm):
    """Parameters every TP rank keeps whole see only 1/tp of the tokens: their gradients are summed over the group."""
    p._b200_tp_group = comm.group
    if getattr(p, "_b200_tp_hooked", False):
        return
    p._b200_tp_hooked = True

    def hook(param):
        if getattr(param, "_b200_fsdp_unit", None) is not None:
            return  # the sharded engine folds this reduction into its own
        if param.grad is not None and comm.world > 1:
            torch.distributed.all_reduce(param.grad, group=comm.group)

    p.register_post_accumulate_grad_hook(hook)


def apply_tensor_parallel(model: nn.Module, tp_mesh, tp_config=None, max_tokens: Optional[int] = None) -> nn.Module:
    """Shard a native ``LlamaForCausalLM`` over the ``tp`` mesh axis (in place) and return it.

    QKV / gate-up become column-parallel (split by heads / by intermediate features), attention-out /
    down become row-parallel, norms / embedding / LM head stay replicated with gradient all-reduce,
    activations between blocks are token-sharded.
    """
    from ..models.llama import LlamaForCausalLM

    if not isinstance(model, LlamaForCausalLM):
        raise NotImplementedError("native tensor parallelism is implemented for accelerate_hpc_test_b200.models.LlamaForCausalLM")
    group = tp_mesh.get_group() if hasattr(tp_mesh, "get_group") else tp_mesh
    cfg = model.config
    device = next(model.parameters()).device
    world = torch.distributed.get_world_size(group)
    rank = torch.distributed.get_rank(group)
    if cfg.num_key_value_heads % world or cfg.num_attention_heads % world or cfg.intermediate_size % world:
        raise ValueError(f"tp_size={world} must divide the head counts ({cfg.num_attention_heads}/{cfg.num_key_value_heads}) and the FFN size")
    fused = True if tp_config is None else bool(getattr(tp_config, "enable_async_tp", False) or os.environ.get("ACCELERATE_B200_FUSED_TP", "1") == "1")
    max_tokens = max_tokens or int(os.environ.get("ACCELERATE_B200_TP_MAX_TOKENS", "16384"))
    comm = TPComm(group, device, cfg.hidden_size, max_tokens, fused=fused)
    D = cfg.head_dim
    nq, nk = cfg.num_attention_heads * D, cfg.num_key_value_heads * D
    I = cfg.intermediate_size
    before = {name: param for name, param in model.named_parameters()}
    for layer in model.model.layers:
        attn, mlp = layer.self_attn, layer.mlp
        w = attn.qkv_proj.weight.data
        wq, wk, wv = w[:nq], w[nq : nq + nk], w[nq + nk :]
        local = torch.cat([_shard_rows(wq, rank, world), _shard_rows(wk, rank, world), _shard_rows(wv, rank, world)], dim=0)
        attn.qkv_proj = ColumnParallelLinear(local, comm)
        attn.o_proj = RowParallelLinear(_shard_cols(attn.o_proj.weight.data, rank, world), comm)
        gu = mlp.gate_up_proj.weight.data
        mlp.gate_up_proj = ColumnParallelLinear(torch.cat([_shard_rows(gu[:I], rank, world), _shard_rows(gu[I:], rank, world)], dim=0), comm)
        mlp.down_proj = RowParallelLinear(_shard_cols(mlp.down_proj.weight.data, rank, world), comm)
        attn.tp_heads = (cfg.num_attention_heads // world, cfg.num_key_value_heads // world)
        for p in (layer.input_layernorm.weight, layer.post_attention_layernorm.weight):
            _mark_replicated(p, comm)
    # the embedding sees the all-gathered gradient of every token (backward of the token scatter), so its
    # gradient is already complete on each rank; the final norm and the LM head only see their token shard
    for p in (model.model.norm.weight, model.lm_head.weight):
        _mark_replicated(p, comm)
    model.model.tp_comm = comm
    model.tp_comm = comm
    # the sharded linears hold new Parameter objects: tell `Accelerator.prepare` how to re-point optimizers created before
    # (chained behind an earlier replacement, e.g. a transformers model adopted as the native one)
    after = dict(model.named_parameters())
    replaced = {id(old): after[name] for name, old in before.items() if name in after and after[name] is not old}
    earlier = getattr(model, "_b200_param_remap", None)
    if earlier:
        model._b200_param_remap = {key: (replaced.get(id(new), new) if new is not None else None) for key, new in earlier.items()}
        model._b200_param_remap.update({key: new for key, new in replaced.items() if key not in model._b200_param_remap})
    else:
        model._b200_param_remap = replaced
    return model


def scatter_tokens(x, comm: TPComm):
    return _ScatterRowsFn.apply(x, comm)


def gather_tokens(x, comm: TPComm):
    return _GatherRowsFn.apply(x, comm)
