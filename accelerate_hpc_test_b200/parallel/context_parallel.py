"""Context parallelism (``cp_size``): the sequence is sharded across the ``cp`` group for the whole model and
attention sees the full context through the group.

The reference wraps the step in ``torch.distributed.tensor.experimental.context_parallel`` (reference
src/accelerate/accelerator.py:1618-1680 ``_prepare_cp`` / ``maybe_context_parallel``: buffers sharded in place along
their sequence dims, SDPA replaced by ring attention with an ``allgather`` or ``alltoall`` rotate method,
utils/dataclasses.py:2186 ``TorchContextParallelConfig``).  The B200 design:

* **layout** — causal load balancing: the sequence is cut into ``2*cp`` chunks and rank ``r`` owns chunks ``r`` and
  ``2cp-1-r`` (one cheap "head" chunk, one expensive "tail" chunk), so every rank does the same attention work;
* **comm** — K/V of one layer is small next to the attention math (GQA: 2*Hkv*D per token), so they are all-gathered
  once per layer over NVSwitch (``cp_comm_strategy="allgather"``) and dK/dV return with one reduce-scatter;
* **compute** — every (local Q chunk, KV chunk) pair runs the tcgen05 flash kernel (full or causal), partial
  results are merged with their log-sum-exp; the backward reuses the *global* LSE / delta so each pair yields exact
  dQ (accumulated in the kernel's fp32 buffer), dK and dV contributions.
"""

from __future__ import annotations

import math
from contextlib import contextmanager

import torch
import torch.distributed as dist
from torch import nn

from ..ops import _lib
from ..ops import functional as F


def _group_of(mesh_or_group):
    return mesh_or_group.get_group() if hasattr(mesh_or_group, "get_group") else mesh_or_group


def chunk_owner(j: int, world: int):
    """(rank, half) holding global chunk ``j`` of the ``2*world`` chunks."""
    return (j, 0) if j < world else (2 * world - 1 - j, 1)


def shard_sequence(t: torch.Tensor, dim: int, rank: int, world: int) -> torch.Tensor:
    """This rank's head+tail chunks of ``t`` along ``dim`` (length divisible by ``2*world``)."""
    n = t.shape[dim]
    if n % (2 * world):
        raise ValueError(f"context parallelism needs the sequence length ({n}) to be divisible by 2*cp_size ({2 * world}).")
    c = n // (2 * world)
    head = t.narrow(dim, rank * c, c)
    tail = t.narrow(dim, (2 * world - 1 - rank) * c, c)
    return torch.cat([head, tail], dim=dim)


def unshard_sequence(parts, dim: int) -> torch.Tensor:
    """Inverse of :func:`shard_sequence` given every rank's shard (list indexed by rank)."""
    world = len(parts)
    chunks = [None] * (2 * world)
    for r, p in enumerate(parts):
        c = p.shape[dim] // 2
        chunks[r] = p.narrow(dim, 0, c)
        chunks[2 * world - 1 - r] = p.narrow(dim, c, c)
    return torch.cat(chunks, dim=dim)


class CPComm:
    def __init__(self, group, device, strategy: str = "allgather"):
        self.group = _group_of(group)
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.device = torch.device(device)
        self.strategy = strategy
        # K/V all-gather and dK/dV reduce-scatter run on the peer-memory kernels (TMA pull / NVLS reduction); c10d is only
        # the CPU oracle path and the fallback for shapes the kernels do not take
        from .peer_collectives import PeerCollectives

        self.peer = PeerCollectives(self.group, self.device)

    def local_positions(self, S_loc: int, device) -> torch.Tensor:
        c = S_loc // 2
        a = torch.arange(c, device=device)
        return torch.cat([a + self.rank * c, a + (2 * self.world - 1 - self.rank) * c])

    def all_reduce_(self, t):
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        return t

    def all_gather_rows(self, x: torch.Tensor) -> torch.Tensor:
        """``[rows, w] -> [world, rows, w]``"""
        return self.peer.all_gather(x)

    def reduce_scatter_rows(self, x: torch.Tensor) -> torch.Tensor:
        """``[world, rows, w] -> [rows, w]`` (sum over ranks of slice ``rank``)"""
        return self.peer.reduce_scatter(x)

    def shift_labels_across_ranks(self, labels: torch.Tensor) -> torch.Tensor:
        parts = [torch.empty_like(labels) for _ in range(self.world)]
        dist.all_gather(parts, labels.contiguous(), group=self.group)
        full = unshard_sequence(parts, 1)
        sh = torch.full_like(full, -100)
        sh[:, :-1] = full[:, 1:]
        return shard_sequence(sh, 1, self.rank, self.world)


# ------------------------------------------------------------------------------------------- chunk-pair attention
def _pair_fwd(q, k, v, S, Hq, Hkv, D, causal, scale):
    """One (Q chunk, KV chunk) block: ``q [S, Hq*D]``, ``k, v [S, Hkv*D]`` (strided views) -> ``o`` fp32/bf16, ``lse [Hq, S]``."""
    if q.is_cuda:
        lib = _lib.require_gpu_lib()
        o = torch.empty((S, Hq * D), dtype=torch.bfloat16, device=q.device)
        lse = torch.empty((1, Hq, S), dtype=torch.float32, device=q.device)
        _lib.check(lib.b200_flash_attn_fwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(o), _lib.ptr(lse), 1, S, Hq, Hkv, D,
                                           q.stride(0), k.stride(0), v.stride(0), o.stride(0), float(scale), int(causal),
                                           _lib.cur_stream()), "flash_attn_fwd")
        return o, lse[0]
    rep = Hq // Hkv
    qf = q.float().view(S, Hq, D).transpose(0, 1)
    kf = k.float().view(S, Hkv, D).transpose(0, 1).repeat_interleave(rep, dim=0)
    vf = v.float().view(S, Hkv, D).transpose(0, 1).repeat_interleave(rep, dim=0)
    s = (qf @ kf.transpose(-1, -2)) * scale
    if causal:
        s = s.masked_fill(~torch.ones(S, S, dtype=torch.bool).tril(), float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    o = (torch.exp(s - lse[..., None]) @ vf).transpose(0, 1).reshape(S, Hq * D)
    return o, lse


def _pair_bwd(q, k, v, do, lse, delta, dq_acc, S, Hq, Hkv, D, causal, scale):
    """Backward of one block with the GLOBAL ``lse`` / ``delta`` of the Q rows.  Adds the unscaled dQ into ``dq_acc``
    (fp32 ``[Hq, S_pad, D]``, the kernel's accumulator layout) and returns this block's ``(dk, dv)`` ``[S, Hkv*D]``."""
    if q.is_cuda:
        lib = _lib.require_gpu_lib()
        dkv = torch.empty((S, 2 * Hkv * D), dtype=torch.bfloat16, device=q.device)
        dk, dv = dkv[:, : Hkv * D], dkv[:, Hkv * D :]
        _lib.check(lib.b200_flash_attn_bwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(do), _lib.ptr(lse), _lib.ptr(delta),
                                           _lib.ptr(dq_acc), _lib.ptr(dk), _lib.ptr(dv), 1, S, Hq, Hkv, D, q.stride(0), k.stride(0),
                                           v.stride(0), do.stride(0), dkv.stride(0), float(scale), int(causal), _lib.cur_stream()),
                   "flash_attn_bwd")
        return dk, dv
    rep = Hq // Hkv
    qf = q.float().view(S, Hq, D).transpose(0, 1)
    kf = k.float().view(S, Hkv, D).transpose(0, 1).repeat_interleave(rep, dim=0)
    vf = v.float().view(S, Hkv, D).transpose(0, 1).repeat_interleave(rep, dim=0)
    dof = do.float().view(S, Hq, D).transpose(0, 1)
    s = (qf @ kf.transpose(-1, -2)) * scale
    if causal:
        s = s.masked_fill(~torch.ones(S, S, dtype=torch.bool).tril(), float("-inf"))
    p = torch.exp(s - lse[..., None])
    dv = p.transpose(-1, -2) @ dof
    ds = p * (dof @ vf.transpose(-1, -2) - delta[..., None])
    dq_acc[:, :S].add_(ds @ kf)  # unscaled, like the kernel's accumulator
    dk = (ds.transpose(-1, -2) @ qf) * scale
    dk = dk.view(Hkv, rep, S, D).sum(1).transpose(0, 1).reshape(S, Hkv * D)
    dv = dv.view(Hkv, rep, S, D).sum(1).transpose(0, 1).reshape(S, Hkv * D)
    return dk, dv


def _merge(o_acc, lse_acc, o_j, lse_j):
    """Fold one block's normalised output into the running (output, lse) pair.  ``o*: [S, H, D]`` fp32, ``lse*: [H, S]``."""
    new = torch.logaddexp(lse_acc, lse_j)
    wa = torch.exp(lse_acc - new).transpose(0, 1).unsqueeze(-1)
    wj = torch.exp(lse_j - new).transpose(0, 1).unsqueeze(-1)
    return o_acc * wa + o_j * wj, new


class _CPAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, cos, sin, pos, B, S_loc, Hq, Hkv, D, scale, comm: CPComm):
        w, r = comm.world, comm.rank
        c = S_loc // 2
        T = B * S_loc
        qkv2 = qkv.reshape(T, (Hq + 2 * Hkv) * D)
        if not qkv2.is_cuda:
            qkv2 = qkv2.clone()
        F.rope_inplace(qkv2, cos, sin, pos, Hq + Hkv, D)
        kv_all = comm.all_gather_rows(qkv2[:, Hq * D :])  # [w, T, 2*Hkv*D]
        o = torch.empty((T, Hq * D), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((B, 2, Hq, c), dtype=torch.float32, device=qkv.device)
        for b in range(B):
            for half, qi in enumerate((r, 2 * w - 1 - r)):
                rows = slice(b * S_loc + half * c, b * S_loc + half * c + c)
                q = qkv2[rows, : Hq * D]
                o_acc = l_acc = None
                for j in range(qi + 1):
                    owner, h2 = chunk_owner(j, w)
                    kv = kv_all[owner, b * S_loc + h2 * c : b * S_loc + h2 * c + c]
                    o_j, l_j = _pair_fwd(q, kv[:, : Hkv * D], kv[:, Hkv * D :], c, Hq, Hkv, D, j == qi, scale)
                    o_j = o_j.float().view(c, Hq, D)
                    if o_acc is None:
                        o_acc, l_acc = o_j, l_j
                    else:
                        o_acc, l_acc = _merge(o_acc, l_acc, o_j, l_j)
                o[rows] = o_acc.reshape(c, Hq * D).to(o.dtype)
                lse[b, half] = l_acc
        ctx.save_for_backward(qkv2, kv_all, o, lse, cos, sin, pos)
        ctx.cfg = (B, S_loc, Hq, Hkv, D, scale)
        ctx.comm = comm
        ctx.in_shape = qkv.shape
        return o

    @staticmethod
    def backward(ctx, do):
        qkv2, kv_all, o, lse, cos, sin, pos = ctx.saved_tensors
        B, S_loc, Hq, Hkv, D, scale = ctx.cfg
        comm = ctx.comm
        w, r = comm.world, comm.rank
        c = S_loc // 2
        T = B * S_loc
        do = do.reshape(T, Hq * D).contiguous()
        # delta[b, half, h, s] = sum_d do * o for every local q row
        delta = (do.float() * o.float()).view(B, 2, c, Hq, D).sum(-1).permute(0, 1, 3, 2).contiguous()
        c_pad = (c + 63) // 64 * 64
        dq_acc = torch.zeros((B, 2, Hq, c_pad, D), dtype=torch.float32, device=do.device)
        dkv_all = torch.zeros(kv_all.shape, dtype=torch.float32, device=do.device)
        for b in range(B):
            for half, qi in enumerate((r, 2 * w - 1 - r)):
                rows = slice(b * S_loc + half * c, b * S_loc + half * c + c)
                q = qkv2[rows, : Hq * D]
                for j in range(qi + 1):
                    owner, h2 = chunk_owner(j, w)
                    krows = slice(b * S_loc + h2 * c, b * S_loc + h2 * c + c)
                    kv = kv_all[owner, krows]
                    dk, dv = _pair_bwd(q, kv[:, : Hkv * D], kv[:, Hkv * D :], do[rows], lse[b, half], delta[b, half],
                                       dq_acc[b, half], c, Hq, Hkv, D, j == qi, scale)
                    dkv_all[owner, krows, : Hkv * D] += dk
                    dkv_all[owner, krows, Hkv * D :] += dv
        dkv = comm.reduce_scatter_rows(dkv_all)
        dqkv = torch.empty_like(qkv2)
        # [B, 2, Hq, c, D] -> token-major [B*S_loc, Hq*D]
        dq = dq_acc[:, :, :, :c].permute(0, 1, 3, 2, 4).reshape(T, Hq * D)
        dqkv[:, : Hq * D] = (dq * scale).to(dqkv.dtype)
        dqkv[:, Hq * D :] = dkv.to(dqkv.dtype)
        F.rope_inplace(dqkv, cos, sin, pos, Hq + Hkv, D, backward=True)
        return (dqkv.view(ctx.in_shape),) + (None,) * 10


def cp_attention(qkv, cos, sin, pos, B, S_loc, Hq, Hkv, D, comm: CPComm, scale=None):
    """Causal attention for a head+tail sequence shard; ``qkv`` is the local fused projection ``[B*S_loc, (Hq+2Hkv)*D]``."""
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if S_loc % 2:
        raise ValueError("context parallelism needs an even local sequence length (head + tail chunk).")
    return _CPAttentionFn.apply(qkv, cos, sin, pos, B, S_loc, Hq, Hkv, D, scale, comm)


def apply_context_parallel(model: nn.Module, cp_mesh, cp_config=None) -> nn.Module:
    group = _group_of(cp_mesh)
    if dist.get_world_size(group) == 1:
        return model
    strategy = getattr(cp_config, "cp_comm_strategy", None) or "allgather"
    attn_blocks = [m for m in model.modules() if hasattr(m, "qkv_proj") and hasattr(m, "o_proj")]
    if not attn_blocks:
        raise ValueError("apply_context_parallel found no attention blocks (native models expose `qkv_proj` / `o_proj`).")
    comm = CPComm(group, next(model.parameters()).device, strategy)
    for m in attn_blocks:
        m.cp_comm = comm
    model.cp_comm = comm
    inner = getattr(model, "model", None)
    if inner is not None:
        inner.cp_comm = comm
    return model


@contextmanager
def context_parallel_buffers(cp_mesh, buffers, buffer_seq_dims, no_restore_buffers=()):
    """Shard ``buffers`` in place along ``buffer_seq_dims`` (head+tail layout) for the body of the ``with`` block and put
    the originals back afterwards, except those in ``no_restore_buffers`` (same contract as torch's
    ``context_parallel`` used at reference accelerator.py:1660-1680)."""
    group = _group_of(cp_mesh)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if len(buffers) != len(buffer_seq_dims):
        raise ValueError("`buffers` and `buffer_seq_dims` must have the same length.")
    keep = {id(b) for b in no_restore_buffers}
    saved = []
    for buf, dim in zip(buffers, buffer_seq_dims):
        saved.append(None if id(buf) in keep else buf.clone())
        shard = shard_sequence(buf, dim, rank, world).contiguous()
        buf.resize_(shard.shape)
        buf.copy_(shard)
    try:
        yield
    finally:
        for buf, orig in zip(buffers, saved):
            if orig is not None:
                buf.resize_(orig.shape)
                buf.copy_(orig)
