"""Ulysses sequence parallelism (``sp_size``): activations are sharded along the sequence everywhere
except inside attention, where an all-to-all re-shards them along the heads so that every rank runs
ordinary flash attention over the *whole* sequence for ``heads / sp`` heads.

The reference reaches this through DeepSpeed (reference src/accelerate/accelerator.py:2358-2410:
``UlyssesSPAttentionHF.register_with_transformers`` + ``UlyssesSPDataLoaderAdapter``; config
utils/dataclasses.py:2214 ``DeepSpeedSequenceParallelConfig``).  Here it is part of the engine:

* the QKV projection and the all-to-all are ONE kernel: the GEMM epilogue TMA-stores every output block straight into the
  head-sharded buffer of the rank that owns those heads (``SPComm.gemm_scatter``); the same kernel serves the backward of the
  output projection (dgrad GEMM -> all-to-all).  Its two destination buffers alternate, and every call ends with one device
  barrier, which also proves that all ranks are done reading the buffer the next call will overwrite;
* in the other direction (head- to sequence-sharded) the two layout changes are folded into the copy itself —
  ``b200_seq_head_exchange`` (csrc/comm.cu) pulls the needed column segments row by row from every
  peer's symmetric buffer with TMA bulk copies over NVLink, so there is no pack / unpack pass;
* on CPU (gloo) the same exchange is ``all_to_all_single`` plus the two permutes (the oracle path);
* the loss is the global token mean: every rank contributes ``sum_local / count_global`` and the
  data-parallel gradient averaging (which spans the ``sp`` axis) is compensated by ``sp``.
"""

from __future__ import annotations

import ctypes
import math
import os

import torch
import torch.distributed as dist
from torch import nn

from ..ops import _lib
from ..ops import functional as F


def _group_of(mesh_or_group):
    return mesh_or_group.get_group() if hasattr(mesh_or_group, "get_group") else mesh_or_group


class SPComm:
    """Sequence <-> head exchange over one ``sp`` process group."""

    def __init__(self, group, device, max_tokens: int = 0, row_width: int = 0, fused: bool = True):
        self.group = _group_of(group)
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.device = torch.device(device)
        self.ws = None
        self._bufs = {}
        if fused and self.device.type == "cuda" and self.world > 1 and _lib.available():
            from .symm import SymmetricWorkspace

            self._max_bytes = 0
            self._ws_cls = SymmetricWorkspace
            if max_tokens and row_width:
                self._ensure(max_tokens * row_width * 2)

    # ------------------------------------------------------------------ symmetric staging buffer
    def _ensure(self, nbytes: int):
        """One symmetric staging buffer, grown collectively (all ranks see the same shapes)."""
        if self.ws is not None and nbytes <= self._max_bytes:
            return
        if self.ws is not None:
            self.ws.close()
        self.ws = self._ws_cls(self.group, int(nbytes * 1.25) + (1 << 20), self.device)
        self._src_off = self.ws.alloc(int(nbytes * 1.25))
        self._max_bytes = int(nbytes * 1.25)
        self._ch_arrive = self.ws.new_channel()
        self._ch_done = self.ws.new_channel()
        # CTAs of the head<->sequence exchange kernel (0: the library's default, three per SM)
        self._exchange_ctas = int(os.environ.get("ACCELERATE_B200_EXCHANGE_CTAS", "0"))

    def _native(self, x2d: torch.Tensor) -> bool:
        return x2d.is_cuda and self.world > 1 and hasattr(self, "_ws_cls")

    # ------------------------------------------------------------------ the exchange
    def exchange(self, x: torch.Tensor, mode: int, B: int, S_loc: int, widths) -> torch.Tensor:
        """``mode 0``: ``x [B*S_loc, sum(widths)]`` (every segment holds all heads) -> ``[B*S, sum(widths)/sp]``;
        ``mode 1``: ``x [B*S, sum(widths)/sp]`` -> ``[B*S_loc, sum(widths)]``.  ``widths`` are the full
        (un-sharded) segment widths in elements, e.g. ``(Hq*D, Hkv*D, Hkv*D)`` for a fused QKV buffer."""
        w = self.world
        S = S_loc * w
        full = sum(widths)
        part = full // w
        if w == 1:
            return x
        rows_in, cols_in = (B * S_loc, full) if mode == 0 else (B * S, part)
        rows_out, cols_out = (B * S, part) if mode == 0 else (B * S_loc, full)
        assert x.shape == (rows_in, cols_in), (tuple(x.shape), (rows_in, cols_in))
        if not self._native(x):
            return self._exchange_reference(x, mode, B, S_loc, widths)
        es = x.element_size()
        self._ensure(rows_in * cols_in * es)
        stage = self.ws.tensor(self._src_off, (rows_in, cols_in), x.dtype)
        stage.copy_(x)
        out = torch.empty((rows_out, cols_out), dtype=x.dtype, device=x.device)
        segs = []
        fo = po = 0
        for wd in widths:
            pw = wd // w
            if mode == 0:
                segs += [fo * es, po * es, pw * es]
            else:
                segs += [po * es, fo * es, pw * es]
            fo += wd
            po += pw
        arr = (ctypes.c_longlong * len(segs))(*segs)
        e = self.ws.next_epoch(self._ch_arrive)
        self.ws._epochs[self._ch_done] = e
        _lib.check(
            self.ws.lib.b200_seq_head_exchange(self.ws.peer_array, self.rank, w, self._src_off, _lib.ptr(out), mode, B, S_loc,
                                               cols_in * es, cols_out * es, arr, len(widths), self._ch_arrive, self._ch_done, e,
                                               self.ws.counter_ptr(self._ch_arrive), self._exchange_ctas, _lib.cur_stream()),
            "seq_head_exchange",
        )
        return out

    # ------------------------------------------------------------------ GEMM fused with the sequence->head exchange
    def gemm_scatter(self, x2: torch.Tensor, w: torch.Tensor, B: int, S_loc: int, widths, b_mn_major: bool = False) -> torch.Tensor:
        """``exchange(x2 @ W^T, mode 0)`` in one kernel: the GEMM epilogue TMA-stores every 32 x 64 block straight into the
        buffer of the rank that owns those heads (csrc/gemm2_tcgen05.cu, ``b200_gemm2_bf16_ulysses_scatter``)."""
        w_ = self.world
        full = sum(widths)
        part = full // w_
        fused = (self._native(x2) and x2.dtype == torch.bfloat16 and S_loc % 32 == 0 and all(wd % (64 * w_) == 0 for wd in widths)
                 and os.environ.get("ACCELERATE_B200_ULYSSES_FUSED", "1") != "0")
        if not fused:
            y = F.gemm(x2, w, b_mn_major=b_mn_major)
            return self.exchange(y, 0, B, S_loc, widths)
        rows = B * S_loc * w_
        nbytes = rows * part * 2
        self._ensure_dst(nbytes)
        off = self._dst_off[self._dst_flip]
        self._dst_flip ^= 1
        x2 = F._rowmajor2d(x2)
        arr = (ctypes.c_int * len(widths))(*widths)
        _lib.check(
            self.ws_dst.lib.b200_gemm2_bf16_ulysses_scatter(_lib.ptr(x2), _lib.ptr(w), self.ws_dst.peer_array, self.rank, w_, off, B, S_loc, arr,
                                                            len(widths), x2.shape[1], x2.stride(0), w.stride(0), int(b_mn_major),
                                                            _lib.cur_stream()),
            "gemm_ulysses_scatter",
        )
        # one barrier per call: everybody's blocks have landed here; the two destination buffers alternate, so a rank
        # that passed this barrier has also finished reading the buffer the NEXT call writes (see module docstring)
        self.ws_dst.barrier(self._dst_ch)
        return self.ws_dst.tensor(off, (rows, part), torch.bfloat16).clone()

    def _ensure_dst(self, nbytes: int):
        if getattr(self, "ws_dst", None) is not None and nbytes <= self._dst_bytes:
            return
        if getattr(self, "ws_dst", None) is not None:
            self.ws_dst.close()
        cap = int(nbytes * 1.1) + 4096
        self.ws_dst = self._ws_cls(self.group, 2 * cap + (1 << 20), self.device)
        self._dst_off = [self.ws_dst.alloc(cap, align=1024), self.ws_dst.alloc(cap, align=1024)]
        self._dst_bytes = cap
        self._dst_flip = 0
        self._dst_ch = self.ws_dst.new_channel()

    def gather_positions(self, pos: torch.Tensor, B: int, S_loc: int) -> torch.Tensor:
        """Positions of the whole sequence in head-sharded row order ``(b, rank, s)`` from every rank's local ones."""
        parts = [torch.empty_like(pos) for _ in range(self.world)]
        dist.all_gather(parts, pos.contiguous(), group=self.group)
        return torch.stack([p_.view(B, S_loc) for p_ in parts], dim=1).reshape(-1).contiguous()

    def _exchange_reference(self, x, mode, B, S_loc, widths):
        w = self.world
        outs = []
        off = 0
        for wd in widths:
            pw = wd // w
            if mode == 0:
                seg = x[:, off : off + wd].reshape(B, S_loc, w, pw).permute(2, 0, 1, 3).contiguous()  # [dst, B, S_loc, pw]
                recv = torch.empty_like(seg)
                dist.all_to_all_single(recv, seg, group=self.group)  # [src, B, S_loc, pw]
                outs.append(recv.permute(1, 0, 2, 3).reshape(B * S_loc * w, pw))
                off += wd
            else:
                seg = x[:, off : off + pw].reshape(B, w, S_loc, pw).permute(1, 0, 2, 3).contiguous()  # [dst, B, S_loc, pw]
                recv = torch.empty_like(seg)
                dist.all_to_all_single(recv, seg, group=self.group)  # [src(head block), B, S_loc, pw]
                outs.append(recv.permute(1, 2, 0, 3).reshape(B * S_loc, wd))
                off += pw
        return torch.cat(outs, dim=1)

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        return t

    def shift_labels_across_ranks(self, labels: torch.Tensor) -> torch.Tensor:
        """Next-token labels for a contiguous sequence slice: the successor of the last local token is the
        first label of the next rank's slice (``-100`` on the last rank)."""
        firsts = [torch.empty_like(labels[:, :1]) for _ in range(self.world)]
        dist.all_gather(firsts, labels[:, :1].contiguous(), group=self.group)
        sh = torch.full_like(labels, -100)
        sh[:, :-1] = labels[:, 1:]
        if self.rank + 1 < self.world:
            sh[:, -1:] = firsts[self.rank + 1]
        return sh


class _UlyssesAttentionFn(torch.autograd.Function):
    """RoPE on the local tokens, sequence->head exchange, flash attention over the full sequence for the
    local heads, head->sequence exchange.  Backward runs the same four steps mirrored."""

    @staticmethod
    def forward(ctx, qkv, cos, sin, pos, B, S_loc, Hq, Hkv, D, causal, scale, comm: SPComm):
        w = comm.world
        S = S_loc * w
        hq, hkv = Hq // w, Hkv // w
        qkv2 = qkv.reshape(B * S_loc, (Hq + 2 * Hkv) * D)
        gpu = qkv2.is_cuda
        if not gpu:
            qkv2 = qkv2.clone()
        F.rope_inplace(qkv2, cos, sin, pos, Hq + Hkv, D)
        qkv_h = comm.exchange(qkv2, 0, B, S_loc, (Hq * D, Hkv * D, Hkv * D))
        if gpu:
            o_h, lse = F.flash_attn_fwd(qkv_h, B, S, hq, hkv, D, causal, scale)
        else:
            o_h, lse = F._attn_reference(qkv_h, B, S, hq, hkv, D, causal, scale).to(qkv.dtype), torch.empty(0)
        o = comm.exchange(o_h, 1, B, S_loc, (Hq * D,))
        ctx.save_for_backward(qkv_h, o_h, lse, cos, sin, pos)
        ctx.cfg = (B, S_loc, Hq, Hkv, D, causal, scale)
        ctx.comm = comm
        ctx.in_shape = qkv.shape
        return o

    @staticmethod
    def backward(ctx, do):
        qkv_h, o_h, lse, cos, sin, pos = ctx.saved_tensors
        B, S_loc, Hq, Hkv, D, causal, scale = ctx.cfg
        comm = ctx.comm
        w = comm.world
        S = S_loc * w
        hq, hkv = Hq // w, Hkv // w
        do_h = comm.exchange(do.reshape(B * S_loc, Hq * D).contiguous(), 0, B, S_loc, (Hq * D,))
        if qkv_h.is_cuda:
            dqkv_h = F.flash_attn_bwd(do_h, qkv_h, o_h, lse, B, S, hq, hkv, D, causal, scale)
        else:
            with torch.enable_grad():
                x = qkv_h.detach().float().requires_grad_(True)
                ref = F._attn_reference(x, B, S, hq, hkv, D, causal, scale)
                (dqkv_h,) = torch.autograd.grad(ref, x, do_h.float())
            dqkv_h = dqkv_h.to(qkv_h.dtype)
        dqkv = comm.exchange(dqkv_h, 1, B, S_loc, (Hq * D, Hkv * D, Hkv * D))
        F.rope_inplace(dqkv, cos, sin, pos, Hq + Hkv, D, backward=True)
        return (dqkv.view(ctx.in_shape),) + (None,) * 11


class _UlyssesQKVFn(torch.autograd.Function):
    """QKV projection of the local tokens, delivered head-sharded: ``[B*S, (Hq+2Hkv)*D / sp]`` for the whole sequence.
    Forward is one kernel (GEMM + all-to-all); backward brings dQKV back (exchange kernel) and runs dgrad / wgrad."""

    @staticmethod
    def forward(ctx, x, weight, comm: SPComm, B, S_loc, widths):
        x2 = x.reshape(B * S_loc, x.shape[-1])
        out = comm.gemm_scatter(x2, weight, B, S_loc, widths)
        ctx.save_for_backward(x2, weight)
        ctx.comm, ctx.cfg, ctx.in_shape = comm, (B, S_loc, widths), x.shape
        return out

    @staticmethod
    def backward(ctx, dqkv_h):
        x2, weight = ctx.saved_tensors
        B, S_loc, widths = ctx.cfg
        dy = ctx.comm.exchange(dqkv_h.contiguous(), 1, B, S_loc, widths)  # [B*S_loc, N]
        if dy.dtype != weight.dtype:
            dy = dy.to(weight.dtype)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = F.gemm(dy, weight, b_mn_major=True).view(ctx.in_shape)
        if ctx.needs_input_grad[1]:
            dw = F.deliver_weight_grad(weight, lambda out, acc: F.gemm(dy, x2, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc))
        return dx, dw, None, None, None, None


class _UlyssesOutProjFn(torch.autograd.Function):
    """Attention output projection fed by head-sharded attention output: forward brings the local tokens' heads together
    (exchange kernel) and multiplies; backward is one kernel again — dgrad GEMM whose epilogue scatters d(attn out) back to the
    ranks that own the heads."""

    @staticmethod
    def forward(ctx, o_h, weight, comm: SPComm, B, S_loc):
        width = weight.shape[1]
        o = comm.exchange(o_h.contiguous(), 1, B, S_loc, (width,))  # [B*S_loc, Hq*D]
        y = F.gemm(o, weight)
        ctx.save_for_backward(o, weight)
        ctx.comm, ctx.cfg = comm, (B, S_loc, width)
        return y

    @staticmethod
    def backward(ctx, dy):
        o, weight = ctx.saved_tensors
        B, S_loc, width = ctx.cfg
        dy2 = dy.reshape(B * S_loc, dy.shape[-1])
        if dy2.dtype != weight.dtype:
            dy2 = dy2.to(weight.dtype)
        do_h = dw = None
        if ctx.needs_input_grad[0]:
            do_h = ctx.comm.gemm_scatter(dy2, weight, B, S_loc, (width,), b_mn_major=True)  # dO = dY W, scattered by heads
        if ctx.needs_input_grad[1]:
            dw = F.deliver_weight_grad(weight, lambda out, acc: F.gemm(dy2, o, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc))
        return do_h, dw, None, None, None


def ulysses_attention_block(x, qkv_weight, o_weight, cos, sin, pos_full, B, S_loc, Hq, Hkv, D, comm: SPComm, causal=True):
    """The whole attention block under Ulysses: fused QKV-projection + all-to-all, flash attention over the full sequence for
    this rank's heads (RoPE with global positions), all-to-all + output projection."""
    w = comm.world
    qkv_h = _UlyssesQKVFn.apply(x, qkv_weight, comm, B, S_loc, (Hq * D, Hkv * D, Hkv * D))
    o_h = F.rope_attention(qkv_h, cos, sin, pos_full, B, S_loc * w, Hq // w, Hkv // w, D, causal=causal)
    return _UlyssesOutProjFn.apply(o_h, o_weight, comm, B, S_loc)


def ulysses_attention(qkv, cos, sin, pos, B, S_loc, Hq, Hkv, D, comm: SPComm, causal=True, scale=None):
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    return _UlyssesAttentionFn.apply(qkv, cos, sin, pos, B, S_loc, Hq, Hkv, D, causal, scale, comm)


def apply_ulysses(model: nn.Module, sp_mesh, sp_config=None) -> nn.Module:
    """Switch every attention block of a native model to Ulysses attention over ``sp_mesh``.

    Needs ``num_attention_heads`` and ``num_key_value_heads`` divisible by ``sp`` (DeepSpeed's
    implementation has the same requirement unless KV heads are replicated)."""
    group = _group_of(sp_mesh)
    world = dist.get_world_size(group)
    if world == 1:
        return model
    cfg = getattr(model, "config", None)
    if cfg is None or not hasattr(cfg, "num_attention_heads"):
        raise ValueError("Ulysses sequence parallelism expects a model with a `config` describing its attention heads.")
    if cfg.num_attention_heads % world or cfg.num_key_value_heads % world:
        raise ValueError(f"sp_size={world} must divide num_attention_heads={cfg.num_attention_heads} and "
                         f"num_key_value_heads={cfg.num_key_value_heads}.")
    attn_blocks = [m for m in model.modules() if hasattr(m, "qkv_proj") and hasattr(m, "o_proj")]
    if not attn_blocks:
        raise ValueError("apply_ulysses found no attention blocks (native models expose `qkv_proj` / `o_proj`).")
    device = next(model.parameters()).device
    seq = getattr(sp_config, "sp_seq_length", None) if sp_config is not None else None
    width = (cfg.num_attention_heads + 2 * cfg.num_key_value_heads) * cfg.head_dim
    comm = SPComm(group, device, max_tokens=(int(seq) // world) if seq else 0, row_width=width)
    for m in attn_blocks:
        m.sp_comm = comm
    model.sp_comm = comm
    inner = getattr(model, "model", None)
    if inner is not None:
        inner.sp_comm = comm
    return model


def sp_loss(local_sum_mean: torch.Tensor, valid: torch.Tensor, comm: SPComm) -> torch.Tensor:
    """Global token-mean loss from this rank's mean over its ``valid`` tokens.

    Value: ``sum_r mean_r * n_r / sum_r n_r``.  Gradient: this rank's share times ``sp`` — the sharded /
    data-parallel engine averages gradients over a group that contains the ``sp`` ranks."""
    cnt = valid.to(torch.float32)
    tot = comm.all_reduce_(cnt.clone()).clamp(min=1.0)
    part = local_sum_mean * (cnt / tot)
    whole = comm.all_reduce_(part.detach().clone())
    scaled = part * comm.world
    return scaled + (whole - scaled.detach())


class UlyssesDataLoaderAdapter:
    """Shard every batch along the sequence for the ``sp`` group.

    All ranks of an ``sp`` group receive the same batch from the prepared dataloader (it treats the
    model-parallel axes as one data rank); the adapter keeps this rank's ``S / sp`` slice of every
    ``[B, S]`` tensor and adds ``shift_labels`` (labels shifted *before* slicing, so the last local
    token still sees its successor) and global ``position_ids``.  With ``shared_batches=False`` the
    ranks hold different batches (a loader prepared without the mesh): they are then processed one
    after the other by the whole group, like DeepSpeed's ``UlyssesSPDataLoaderAdapter``
    (reference accelerator.py:2398-2408).
    """

    def __init__(self, dataloader, sp_mesh, device=None, shared_batches: bool = True, seq_keys=None):
        self.dl = dataloader
        self.group = _group_of(sp_mesh)
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.device = device
        self.shared_batches = shared_batches
        self.seq_keys = seq_keys
        self._is_accelerate_prepared = True

    def __len__(self):
        n = len(self.dl)
        return n if self.shared_batches else n * self.world

    def __getattr__(self, name):
        if name in ("dl", "__setstate__"):
            raise AttributeError(name)
        return getattr(self.dl, name)

    def _shard(self, batch: dict) -> dict:
        ids = batch["input_ids"]
        B, S = ids.shape
        w, r = self.world, self.rank
        if S % w:
            raise ValueError(f"sequence length {S} is not divisible by sp_size={w}; pad the batch (pad_to_multiple_of).")
        c = S // w
        out = dict(batch)
        if "labels" in batch and "shift_labels" not in batch:
            lab = batch["labels"]
            sh = torch.full_like(lab, -100)
            sh[:, :-1] = lab[:, 1:]
            out["shift_labels"] = sh
            out.pop("labels")
        if "position_ids" not in out:
            out["position_ids"] = torch.arange(S, device=ids.device).unsqueeze(0).expand(B, S)
        for k, v in list(out.items()):
            if torch.is_tensor(v) and v.dim() >= 2 and v.shape[1] == S and (self.seq_keys is None or k in self.seq_keys):
                out[k] = v[:, r * c : (r + 1) * c].contiguous()
        return out

    def __iter__(self):
        for batch in self.dl:
            if not isinstance(batch, dict):
                raise TypeError("UlyssesDataLoaderAdapter expects dict batches with an `input_ids` entry.")
            if self.shared_batches or self.world == 1:
                yield self._shard(batch)
                continue
            for src in range(self.world):
                objs = [batch if self.rank == src else None]
                # batches are small token tensors: an object broadcast keeps arbitrary batch layouts working
                dist.broadcast_object_list(objs, src=dist.get_global_rank(self.group, src), group=self.group)
                b = objs[0]
                if self.device is not None:
                    b = {k: (v.to(self.device) if torch.is_tensor(v) else v) for k, v in b.items()}
                yield self._shard(b)
