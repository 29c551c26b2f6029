"""Tensor-level wrappers of the sm_100a kernels plus the autograd Functions built on them.

Every public op has two bodies:

* CUDA tensors -> the hand-written kernel in ``lib/libb200.so`` (raises if the library is missing:
  a GPU run must never silently use eager PyTorch);
* CPU tensors  -> a plain-PyTorch fp32 restatement of the same math.  That restatement is the
  oracle of the GPU numerics tests and lets the model/engine layers be unit-tested without a GPU.

Reference parity: these are the K1-K12 rows of SURVEY.md §2.6.  The reference
(hang-zou/accelerate-hpc-test) has no kernels; it reaches the same math through ``nn.Linear`` /
transformers' Llama modules under ``torch.autocast`` (reference src/accelerate/accelerator.py:1791-1802).
"""

from __future__ import annotations

import math
from typing import Optional

import os

import torch

from . import _lib
from ._lib import check, cur_stream, ptr

BF16 = torch.bfloat16


def _on_gpu(*ts) -> bool:
    return any(t is not None and t.is_cuda for t in ts)


def _rowmajor2d(t: torch.Tensor) -> torch.Tensor:
    """Return a 2-D view with unit inner stride and a 16-byte aligned row stride (copy only if needed)."""
    assert t.dim() == 2
    if t.stride(1) != 1 or t.stride(0) % 8 != 0 or t.data_ptr() % 16 != 0:
        rows, cols = t.shape
        if cols % 8 == 0:
            return t.contiguous()
        # TMA wants 16-byte row pitches: copy into rows padded to 8 elements and hand out the [rows, cols] window
        padded = torch.empty((rows, (cols + 7) // 8 * 8), dtype=t.dtype, device=t.device)
        padded[:, :cols].copy_(t)
        return padded[:, :cols]
    return t


# ----------------------------------------------------------------------------------------- GEMM
def gemm(
    a: torch.Tensor,
    b: torch.Tensor,
    *,
    a_mn_major: bool = False,
    b_mn_major: bool = False,
    bias: Optional[torch.Tensor] = None,
    out: Optional[torch.Tensor] = None,
    out_dtype: Optional[torch.dtype] = None,
    accumulate: bool = False,
    max_ctas: int = 0,
) -> torch.Tensor:
    """``D[M,N] (+)= sum_k A(m,k) B(n,k) (+ bias[n])`` on the tcgen05 GEMM.

    ``a`` is ``[M,K]`` (or ``[K,M]`` when ``a_mn_major``), ``b`` is ``[N,K]`` (or ``[K,N]`` when
    ``b_mn_major``); both bf16.  ``out`` may be a strided row-major view (e.g. a column slice).
    """
    if a_mn_major:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn_major:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    if K != Kb:
        raise ValueError(f"gemm reduction mismatch: {tuple(a.shape)} vs {tuple(b.shape)}")
    if out is None:
        if accumulate:
            raise ValueError("accumulate=True needs an existing `out`")
        if out_dtype is None:
            out_dtype = a.dtype
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    else:
        out_dtype = out.dtype
        if out.shape != (M, N) or out.stride(1) != 1:
            raise ValueError("bad `out` for gemm")
    if not _on_gpu(a, b):
        af = a.float().t() if a_mn_major else a.float()
        bf = b.float() if b_mn_major else b.float().t()
        r = af @ bf
        if bias is not None:
            r = r + bias.float()
        if accumulate:
            r = r + out.float()
        out.copy_(r.to(out_dtype))
        return out
    lib = _lib.require_gpu_lib()
    if a.dtype != BF16 or b.dtype != BF16:
        raise TypeError("b200 gemm expects bf16 operands")
    if out_dtype not in (BF16, torch.float32):
        raise TypeError("b200 gemm writes bf16 or fp32")
    # a dim-0-sharded FSDP weight whose peers' rows have not been fetched yet (parallel/fsdp.py): this GEMM fetches them
    job = _pending_gather(b)
    if isinstance(b, torch.nn.Parameter) and getattr(b, "_b200_fsdp_unit", None) is not None:
        b._b200_gemm_b_seen = True  # from the next step on the engine leaves this weight's gather to the GEMMs
    a_job = _pending_gather(a)
    if a_job is not None:
        a_job.materialize()
    a = _rowmajor2d(a)
    b = _rowmajor2d(b)
    if bias is not None:
        bias = bias.to(BF16).contiguous()
    two_cta = _GEMM_2CTA and max_ctas == 0 and M > 128
    if job is not None and not (two_cta and not a_mn_major):
        job.materialize()
        job = None
    mode = 0
    if job is not None:
        mode = 2 if b_mn_major else 1  # B rows are the reduction dimension (dgrad) / the output columns (forward)
    elif two_cta and not (a_mn_major and not b_mn_major) and _CARRY_SOURCES and 2.0 * M * N * K >= 4e9:
        # nothing of its own to fetch: carry the next weight the stream will need (backward: the weight-gradient GEMMs)
        for src in _CARRY_SOURCES:
            job = src.pop_carry_job()
            if job is not None:
                mode = 3
                break
    if mode:
        peers, rank, world, src_off, dst, piece_bytes, rows, ready, expect = job.launch_args()
        check(
            lib.b200_gemm2_bf16_gather(
                ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, a.stride(0), b.stride(0), out.stride(0),
                int(a_mn_major), int(b_mn_major), int(out_dtype == torch.float32), int(accumulate),
                peers, rank, world, src_off, dst, piece_bytes, rows, mode, ready, expect, cur_stream(),
            ),
            "gemm2_bf16_gather",
        )
        return out
    if two_cta:
        # SM-pair kernel (cta_group::2, 256x256 tiles, 6 stages): the default for plain GEMMs
        check(
            lib.b200_gemm2_bf16(
                ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, a.stride(0), b.stride(0), out.stride(0),
                int(a_mn_major), int(b_mn_major), int(out_dtype == torch.float32), int(accumulate), cur_stream(),
            ),
            "gemm2_bf16",
        )
        return out
    check(
        lib.b200_gemm_bf16(
            ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, a.stride(0), b.stride(0), out.stride(0),
            int(a_mn_major), int(b_mn_major), int(out_dtype == torch.float32), int(accumulate),
            int(max_ctas), cur_stream(),
        ),
        "gemm_bf16",
    )
    return out


_GEMM_2CTA = os.environ.get("ACCELERATE_B200_GEMM_2CTA", "1") != "0"
_CARRY_SOURCES: list = []  # sharded engines that may hand a weight gather to any big GEMM (see B200ShardedModule.pop_carry_job)


def _pending_gather(t):
    job = getattr(t, "_b200_ag", None)
    return job if (job is not None and job.pending) else None


def register_carry_source(engine) -> None:
    if engine not in _CARRY_SOURCES:
        _CARRY_SOURCES.append(engine)


def unregister_carry_source(engine) -> None:
    if engine in _CARRY_SOURCES:
        _CARRY_SOURCES.remove(engine)


def materialize(t: torch.Tensor) -> torch.Tensor:
    """Make sure a sharded-engine weight is completely gathered before an op other than the GEMM reads it."""
    job = _pending_gather(t)
    if job is not None:
        job.materialize()
    return t


def deliver_weight_grad(weight: torch.Tensor, compute):
    """Weight-gradient hand-off between the native ops and the data-parallel engines.

    ``compute(out, accumulate)`` produces dW (into ``out`` when given).  If an engine has published a
    destination for this parameter (``weight._b200_grad_buf``: a view of its flat gradient bucket /
    shard buffer), dW is written — or accumulated — there directly, the engine is notified through
    ``weight._b200_grad_hook`` and autograd receives ``None`` (no second copy, no AccumulateGrad pass).
    Otherwise dW is returned to autograd as usual.
    """
    buf = getattr(weight, "_b200_grad_buf", None)
    if buf is None:
        return compute(None, False)
    g = weight.grad
    # accumulate when an earlier micro-batch (no_sync) or an earlier use of a shared parameter (tied embeddings) already
    # wrote this step's gradient into the buffer
    acc = (bool(getattr(weight, "_b200_has_partial", False)) or bool(getattr(weight, "_b200_grad_seen", False))
           or (g is not None and g.data_ptr() == buf.data_ptr()))
    compute(buf, acc)
    weight._b200_grad_seen = True
    weight._b200_grad_hook(weight)
    return None


class _LinearFn(torch.autograd.Function):
    """y = x W^T (+ b).  fwd / dgrad / wgrad all run on the tcgen05 GEMM with no transposes in memory."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        shp = x.shape
        x2 = x.reshape(-1, shp[-1])
        y = gemm(x2, weight, bias=bias)
        ctx.save_for_backward(x2, weight)
        ctx.weight_ref = weight  # the Parameter object itself: the engines hang their hand-off attributes on it
        ctx.has_bias = bias is not None
        ctx.in_shape = shp
        return y.view(*shp[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        weight = ctx.weight_ref
        dy2 = dy.reshape(-1, dy.shape[-1])
        if dy2.dtype != weight.dtype:
            dy2 = dy2.to(weight.dtype)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # dX[T,K] = dY[T,N] W[N,K] : B is stored [reduction, out] -> MN-major
            dx = gemm(dy2, weight, b_mn_major=True).view(ctx.in_shape)
        if ctx.needs_input_grad[1]:
            # dW[N,K] = dY^T X : both operands stored [reduction(T), out] -> MN-major
            dw = deliver_weight_grad(
                weight, lambda out, acc: gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc)
            )
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.float().sum(0).to(dy2.dtype)
        return dx, dw, db


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    return _LinearFn.apply(x, weight, bias)


# ----------------------------------------------------------------------------------------- RMSNorm
def rmsnorm_fwd(x, weight, eps: float, residual=None):
    """Returns ``(y, s, rstd)`` with ``s = x + residual`` (``s is x`` when residual is None)."""
    H = x.shape[-1]
    x2 = x.reshape(-1, H)
    T = x2.shape[0]
    if not _on_gpu(x):
        s = x2.float() if residual is None else (x2.float() + residual.reshape(-1, H).float()).to(x.dtype).float()
        rstd = torch.rsqrt(s.pow(2).mean(-1) + eps)
        y = (s * rstd[:, None] * weight.float()).to(x.dtype)
        s_out = x2 if residual is None else s.to(x.dtype)
        return y.view_as(x), s_out.view_as(x), rstd
    lib = _lib.require_gpu_lib()
    x2 = x2.contiguous()
    y = torch.empty_like(x2)
    rstd = torch.empty(T, dtype=torch.float32, device=x.device)
    if residual is not None:
        r2 = residual.reshape(-1, H).contiguous()
        s = torch.empty_like(x2)
    else:
        r2, s = None, None
    check(lib.b200_rmsnorm_fwd(ptr(x2), ptr(r2), ptr(weight), ptr(y), ptr(s), ptr(rstd), T, H, float(eps), cur_stream()), "rmsnorm_fwd")
    return y.view_as(x), (s if s is not None else x2).view_as(x), rstd


def rmsnorm_bwd(dy, s, weight, rstd, dres=None):
    """Returns ``(dx, dw)``; ``dres`` (gradient arriving on the residual stream) is added to dx."""
    H = s.shape[-1]
    s2 = s.reshape(-1, H)
    dy2 = dy.reshape(-1, H)
    T = s2.shape[0]
    if not _on_gpu(s):
        xh = s2.float() * rstd[:, None]
        g = dy2.float() * weight.float()
        dx = rstd[:, None] * (g - xh * (g * xh).mean(-1, keepdim=True))
        if dres is not None:
            dx = dx + dres.reshape(-1, H).float()
        dw = (dy2.float() * xh).sum(0)
        return dx.to(s.dtype).view_as(s), dw.to(weight.dtype)
    lib = _lib.require_gpu_lib()
    dy2 = dy2.contiguous()
    s2 = s2.contiguous()
    d2 = dres.reshape(-1, H).contiguous() if dres is not None else None
    dx = torch.empty_like(s2)
    dw = torch.empty_like(weight)
    parts = lib.b200_rmsnorm_bwd_parts(T)
    scratch = torch.empty((parts, H), dtype=torch.float32, device=s.device)
    check(lib.b200_rmsnorm_bwd(ptr(dy2), ptr(s2), ptr(weight), ptr(rstd), ptr(d2), ptr(dx), ptr(dw), ptr(scratch), T, H, 0, cur_stream()), "rmsnorm_bwd")
    return dx.view_as(s), dw


class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, eps):
        y, s, rstd = rmsnorm_fwd(x, weight, eps)
        ctx.save_for_backward(s, weight, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        s, weight, rstd = ctx.saved_tensors
        dx, dw = rmsnorm_bwd(dy, s, weight, rstd)
        return dx, dw, None


class _AddRMSNormFn(torch.autograd.Function):
    """(x, residual) -> (rmsnorm(x + residual), x + residual) in one pass each way."""

    @staticmethod
    def forward(ctx, x, residual, weight, eps):
        y, s, rstd = rmsnorm_fwd(x, weight, eps, residual=residual)
        ctx.save_for_backward(s, weight, rstd)
        return y, s

    @staticmethod
    def backward(ctx, dy, ds):
        s, weight, rstd = ctx.saved_tensors
        dx, dw = rmsnorm_bwd(dy, s, weight, rstd, dres=ds)
        return dx, dx, dw, None


def rmsnorm(x, weight, eps: float = 1e-5):
    return _RMSNormFn.apply(x, weight, eps)


def add_rmsnorm(x, residual, weight, eps: float = 1e-5):
    return _AddRMSNormFn.apply(x, residual, weight, eps)


# ----------------------------------------------------------------------------------------- RoPE
def rope_tables(max_pos: int, head_dim: int, theta: float, device) -> tuple[torch.Tensor, torch.Tensor]:
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32, device=device) / head_dim))
    ang = torch.arange(max_pos, dtype=torch.float32, device=device)[:, None] * inv[None, :]
    return ang.cos().contiguous(), ang.sin().contiguous()


def rope_inplace(buf2d: torch.Tensor, cos, sin, pos_i32, n_heads: int, head_dim: int, backward: bool = False):
    """Rotate heads ``0..n_heads-1`` (width ``head_dim``, back to back) of every row of ``buf2d`` in place."""
    T = buf2d.shape[0]
    if not _on_gpu(buf2d):
        half = head_dim // 2
        v = buf2d[:, : n_heads * head_dim].float().view(T, n_heads, head_dim)
        c = cos[pos_i32.long()][:, None, :]
        s = sin[pos_i32.long()][:, None, :] * (-1.0 if backward else 1.0)
        lo, hi = v[..., :half], v[..., half:]
        out = torch.cat([lo * c - hi * s, hi * c + lo * s], dim=-1)
        buf2d[:, : n_heads * head_dim] = out.view(T, -1).to(buf2d.dtype)
        return buf2d
    lib = _lib.require_gpu_lib()
    assert buf2d.stride(1) == 1
    check(lib.b200_rope_inplace(ptr(buf2d), ptr(cos), ptr(sin), ptr(pos_i32), T, buf2d.stride(0), n_heads, head_dim, int(backward), cur_stream()), "rope")
    return buf2d


# ----------------------------------------------------------------------------------------- SwiGLU
def swiglu_fwd(gu: torch.Tensor) -> torch.Tensor:
    I = gu.shape[-1] // 2
    g2 = gu.reshape(-1, 2 * I)
    if not _on_gpu(gu):
        g, u = g2[:, :I].float(), g2[:, I:].float()
        return (torch.nn.functional.silu(g) * u).to(gu.dtype).view(*gu.shape[:-1], I)
    lib = _lib.require_gpu_lib()
    g2 = g2.contiguous()
    act = torch.empty((g2.shape[0], I), dtype=gu.dtype, device=gu.device)
    check(lib.b200_swiglu_fwd(ptr(g2), ptr(act), g2.shape[0], I, cur_stream()), "swiglu_fwd")
    return act.view(*gu.shape[:-1], I)


def swiglu_bwd(gu: torch.Tensor, dact: torch.Tensor) -> torch.Tensor:
    I = gu.shape[-1] // 2
    g2 = gu.reshape(-1, 2 * I)
    d2 = dact.reshape(-1, I)
    if not _on_gpu(gu):
        g, u, d = g2[:, :I].float(), g2[:, I:].float(), d2.float()
        sg = torch.sigmoid(g)
        dg = d * u * sg * (1 + g * (1 - sg))
        du = d * g * sg
        return torch.cat([dg, du], dim=-1).to(gu.dtype).view_as(gu)
    lib = _lib.require_gpu_lib()
    g2 = g2.contiguous()
    d2 = d2.contiguous()
    dgu = torch.empty_like(g2)
    check(lib.b200_swiglu_bwd(ptr(g2), ptr(d2), ptr(dgu), g2.shape[0], I, cur_stream()), "swiglu_bwd")
    return dgu.view_as(gu)


class _SwiGLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        ctx.save_for_backward(gu)
        return swiglu_fwd(gu)

    @staticmethod
    def backward(ctx, dact):
        (gu,) = ctx.saved_tensors
        return swiglu_bwd(gu, dact)


def swiglu(gate_up: torch.Tensor) -> torch.Tensor:
    """``silu(gate) * up`` for the fused ``[gate | up]`` projection output."""
    return _SwiGLUFn.apply(gate_up)


# ----------------------------------------------------------------------------------------- embedding
class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, table):
        flat = ids.reshape(-1).contiguous()
        materialize(table)
        V, H = table.shape
        if not _on_gpu(table):
            out = table[flat]
        else:
            lib = _lib.require_gpu_lib()
            out = torch.empty((flat.numel(), H), dtype=table.dtype, device=table.device)
            check(lib.b200_embedding_fwd(ptr(flat), ptr(table), ptr(out), flat.numel(), H, V, cur_stream()), "embedding_fwd")
        ctx.save_for_backward(flat)
        ctx.table = table  # the Parameter object (for the engines' gradient hand-off), not its data
        ctx.tshape = (V, H)
        ctx.tdtype = table.dtype
        return out.view(*ids.shape, H)

    @staticmethod
    def backward(ctx, dout):
        (flat,) = ctx.saved_tensors
        V, H = ctx.tshape
        d2 = dout.reshape(-1, H).contiguous()

        def compute(out, acc):
            if out is None:
                out = torch.zeros((V, H), dtype=ctx.tdtype, device=dout.device)
            elif not acc:
                out.zero_()
            if not _on_gpu(dout):
                out.index_add_(0, flat, d2.to(ctx.tdtype))
            else:
                lib = _lib.require_gpu_lib()
                check(lib.b200_embedding_bwd(ptr(flat), ptr(d2), ptr(out), flat.numel(), H, V, cur_stream()), "embedding_bwd")
            return out

        return None, deliver_weight_grad(ctx.table, compute)


def embedding(ids: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    return _EmbeddingFn.apply(ids, table)


# ----------------------------------------------------------------------------------------- casts
_CAST_KIND = {(torch.float32, BF16): 0, (BF16, torch.float32): 1, (torch.float32, torch.float32): 2, (BF16, BF16): 3}


def scale_cast(src: torch.Tensor, dst: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """``dst = (dst.dtype)(src * scale)`` in one pass (bf16<->fp32 casts, grad/GA division)."""
    if not _on_gpu(src):
        dst.copy_((src.float() * scale).to(dst.dtype))
        return dst
    lib = _lib.require_gpu_lib()
    kind = _CAST_KIND[(src.dtype, dst.dtype)]
    assert src.is_contiguous() and dst.is_contiguous() and src.numel() == dst.numel()
    check(lib.b200_scale_cast(ptr(src), ptr(dst), src.numel(), float(scale), kind, cur_stream()), "scale_cast")
    return dst


# ----------------------------------------------------------------------------------------- cross entropy
def cross_entropy_chunk_(logits: torch.Tensor, labels: torch.Tensor, inv_count: Optional[torch.Tensor], scale: float = 1.0,
                         ignore_index: int = -100, write_grad: bool = True):
    """Per-row CE loss of a bf16 ``[T, V]`` logits chunk; overwrites it with d(loss)/d(logits) in place."""
    T, V = logits.shape
    if not _on_gpu(logits):
        z = logits.float()
        lse = torch.logsumexp(z, dim=-1)
        valid = labels != ignore_index
        safe = labels.clamp(min=0)
        loss_rows = torch.where(valid, lse - z.gather(1, safe[:, None]).squeeze(1), torch.zeros_like(lse))
        if write_grad:
            g = scale * (float(inv_count) if inv_count is not None else 1.0)
            p = torch.exp(z - lse[:, None])
            p[torch.arange(T), safe] -= 1.0
            p = p * g * valid[:, None].float()
            logits.copy_(p.to(logits.dtype))
        return loss_rows
    lib = _lib.require_gpu_lib()
    assert logits.stride(1) == 1 and labels.dtype == torch.int64
    loss_rows = torch.empty(T, dtype=torch.float32, device=logits.device)
    check(lib.b200_ce_fwd_bwd(ptr(logits), ptr(labels), ptr(loss_rows), None, T, logits.stride(0), V, ptr(inv_count), float(scale), ignore_index, int(write_grad), cur_stream()), "ce_fwd_bwd")
    return loss_rows


def _scale_by_device_scalar(src: torch.Tensor, dst: torch.Tensor, scale: torch.Tensor, accumulate: bool) -> torch.Tensor:
    """``dst (+)= src * scale`` with ``scale`` a 0-dim fp32 tensor on the same device (never read by the host)."""
    fast = (_on_gpu(src) and src.dtype == BF16 and dst.dtype == BF16 and src.is_contiguous() and dst.is_contiguous()
            and src.numel() % 8 == 0 and src.data_ptr() % 16 == 0 and dst.data_ptr() % 16 == 0 and scale.dtype == torch.float32)
    if not fast:
        s = scale.to(src.dtype)
        if accumulate:
            return dst.add_(src * s)
        return src.mul_(s) if dst is src else torch.mul(src, s, out=dst)
    lib = _lib.require_gpu_lib()
    check(lib.b200_scale_by_scalar_bf16(ptr(src), ptr(dst), src.numel(), ptr(scale.reshape(1)), int(accumulate), cur_stream()), "scale_by_scalar")
    return dst


class _FusedLinearCEFn(torch.autograd.Function):
    """mean CE(h W^T, labels) without ever holding full logits: per token-chunk GEMM -> CE(+grad in
    place) -> dgrad / wgrad GEMMs.  Gradients are produced in forward and scaled in backward."""

    @staticmethod
    def forward(ctx, h, weight, labels, ignore_index, chunk_tokens):
        H = h.shape[-1]
        h2 = h.reshape(-1, H)
        lab = labels.reshape(-1).contiguous()
        T = h2.shape[0]
        V = weight.shape[0]
        need_grad = h.requires_grad or weight.requires_grad
        on_gpu = _on_gpu(h2)
        if on_gpu:
            lib = _lib.require_gpu_lib()
            inv = torch.empty(1, dtype=torch.float32, device=h.device)
            check(lib.b200_count_valid(ptr(lab), T, ignore_index, ptr(inv), None, cur_stream()), "count_valid")
        else:
            inv = 1.0 / (lab != ignore_index).sum().clamp(min=1).float()
        dh = torch.empty_like(h2) if need_grad else None
        dw = torch.empty_like(weight) if need_grad else None
        loss_rows = []
        first = True
        for t0 in range(0, T, chunk_tokens):
            t1 = min(T, t0 + chunk_tokens)
            if V % 8:
                # odd vocabularies (GPT-2's 50257): rows on a 16-byte pitch so that TMA and the vector loads stay aligned
                logits = gemm(h2[t0:t1], weight, out=torch.empty((t1 - t0, (V + 7) // 8 * 8), dtype=h2.dtype, device=h2.device)[:, :V])
            else:
                logits = gemm(h2[t0:t1], weight)  # [tc, V] bf16
            rows = cross_entropy_chunk_(logits, lab[t0:t1], inv, 1.0, ignore_index, write_grad=need_grad)
            loss_rows.append(rows)
            if need_grad:
                gemm(logits, weight, b_mn_major=True, out=dh[t0:t1])
                gemm(logits, h2[t0:t1], a_mn_major=True, b_mn_major=True, out=dw, accumulate=not first)
            first = False
            del logits
        rows = torch.cat(loss_rows) if len(loss_rows) > 1 else loss_rows[0]
        if on_gpu:
            loss = torch.empty(1, dtype=torch.float32, device=h.device)
            check(lib.b200_loss_reduce(ptr(rows), T, ptr(inv), 1.0, ptr(loss), 0, cur_stream()), "loss_reduce")
            loss = loss[0]
        else:
            loss = rows.sum() * inv
        if need_grad:
            ctx.save_for_backward(dh, dw)
        ctx.weight = weight
        ctx.h_shape = h.shape
        return loss

    @staticmethod
    def backward(ctx, dloss):
        dh, dw = ctx.saved_tensors
        g = dloss.to(torch.float32)
        # upstream scale (1/GA from Accelerator.backward, loss scaling): one pass, the scale is read on the device
        dh = _scale_by_device_scalar(dh, dh, g, False) if ctx.needs_input_grad[0] else None
        if ctx.needs_input_grad[1]:

            def compute(out, acc):
                return _scale_by_device_scalar(dw, dw if out is None else out, g, acc and out is not None)

            dw = deliver_weight_grad(ctx.weight, compute)
        else:
            dw = None
        return (dh.view(ctx.h_shape) if dh is not None else None), dw, None, None, None


def fused_linear_cross_entropy(h, weight, labels, ignore_index: int = -100, chunk_tokens: int = 4096):
    """Mean cross-entropy of ``h @ weight.T`` against ``labels`` (already shifted by the caller)."""
    return _FusedLinearCEFn.apply(h, weight, labels, ignore_index, chunk_tokens)


# ----------------------------------------------------------------------------------------- attention
def _attn_reference(qkv, B, S, Hq, Hkv, D, causal, scale):
    T = B * S
    q = qkv[:, : Hq * D].float().view(B, S, Hq, D).transpose(1, 2)
    k = qkv[:, Hq * D : (Hq + Hkv) * D].float().view(B, S, Hkv, D).transpose(1, 2)
    v = qkv[:, (Hq + Hkv) * D :].float().view(B, S, Hkv, D).transpose(1, 2)
    rep = Hq // Hkv
    k = k.repeat_interleave(rep, dim=1)
    v = v.repeat_interleave(rep, dim=1)
    s = (q @ k.transpose(-1, -2)) * scale
    if causal:
        mask = torch.ones(S, S, dtype=torch.bool, device=qkv.device).tril()
        s = s.masked_fill(~mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = (p @ v).transpose(1, 2).reshape(T, Hq * D)
    return o


def flash_attn_fwd(qkv: torch.Tensor, B: int, S: int, Hq: int, Hkv: int, D: int, causal: bool, scale: float):
    """Attention over the fused (post-RoPE) QKV buffer ``[B*S, (Hq+2Hkv)*D]`` -> ``(o [B*S, Hq*D], lse)``."""
    T = B * S
    lib = _lib.require_gpu_lib()
    assert qkv.stride(1) == 1 and qkv.dtype == BF16
    o = torch.empty((T, Hq * D), dtype=BF16, device=qkv.device)
    lse = torch.empty((B, Hq, S), dtype=torch.float32, device=qkv.device)
    q = qkv
    k = qkv[:, Hq * D :]
    v = qkv[:, (Hq + Hkv) * D :]
    check(
        lib.b200_flash_attn_fwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(lse), B, S, Hq, Hkv, D,
                                qkv.stride(0), qkv.stride(0), qkv.stride(0), o.stride(0), float(scale), int(causal), cur_stream()),
        "flash_attn_fwd",
    )
    return o, lse


def flash_attn_bwd(do, qkv, o, lse, B, S, Hq, Hkv, D, causal, scale):
    """Returns ``dqkv`` laid out like ``qkv`` (dq | dk | dv)."""
    T = B * S
    lib = _lib.require_gpu_lib()
    do = do.contiguous()
    dqkv = torch.empty_like(qkv)
    delta = torch.empty((B, Hq, S), dtype=torch.float32, device=qkv.device)
    s_pad = (S + 63) // 64 * 64
    dq_acc = torch.zeros((B, Hq, s_pad, D), dtype=torch.float32, device=qkv.device)  # tiles of 64 q rows are contiguous
    check(lib.b200_flash_attn_bwd_prep(ptr(do), ptr(o), ptr(delta), B, S, Hq, do.stride(0), o.stride(0), cur_stream()), "flash_attn_bwd_prep")
    q = qkv
    k = qkv[:, Hq * D :]
    v = qkv[:, (Hq + Hkv) * D :]
    dk = dqkv[:, Hq * D :]
    dv = dqkv[:, (Hq + Hkv) * D :]
    check(
        lib.b200_flash_attn_bwd(ptr(q), ptr(k), ptr(v), ptr(do), ptr(lse), ptr(delta), ptr(dq_acc), ptr(dk), ptr(dv),
                                B, S, Hq, Hkv, D, qkv.stride(0), qkv.stride(0), qkv.stride(0), do.stride(0), dqkv.stride(0),
                                float(scale), int(causal), cur_stream()),
        "flash_attn_bwd",
    )
    check(lib.b200_dq_finalize(ptr(dq_acc), ptr(dqkv), B, S, Hq, dqkv.stride(0), float(scale), cur_stream()), "dq_finalize")
    return dqkv


class _RopeAttentionFn(torch.autograd.Function):
    """RoPE (in place on the fused QKV projection output) + causal GQA flash attention."""

    @staticmethod
    def forward(ctx, qkv, cos, sin, pos, B, S, Hq, Hkv, D, causal, scale):
        T = B * S
        qkv2 = qkv.reshape(T, (Hq + 2 * Hkv) * D)
        on_gpu = _on_gpu(qkv2)
        if on_gpu:
            rope_inplace(qkv2, cos, sin, pos, Hq + Hkv, D)  # the producing GEMM does not need its output
            o, lse = flash_attn_fwd(qkv2, B, S, Hq, Hkv, D, causal, scale)
            ctx.save_for_backward(qkv2, o, lse, cos, sin, pos)
        else:
            qkv2 = qkv2.clone()
            rope_inplace(qkv2, cos, sin, pos, Hq + Hkv, D)
            o = _attn_reference(qkv2, B, S, Hq, Hkv, D, causal, scale).to(qkv.dtype)
            ctx.save_for_backward(qkv2, o, torch.empty(0), cos, sin, pos)
        ctx.cfg = (B, S, Hq, Hkv, D, causal, scale)
        ctx.in_shape = qkv.shape
        return o

    @staticmethod
    def backward(ctx, do):
        qkv2, o, lse, cos, sin, pos = ctx.saved_tensors
        B, S, Hq, Hkv, D, causal, scale = ctx.cfg
        if _on_gpu(qkv2):
            dqkv = flash_attn_bwd(do, qkv2, o, lse, B, S, Hq, Hkv, D, causal, scale)
        else:
            with torch.enable_grad():
                x = qkv2.detach().float().requires_grad_(True)
                ref = _attn_reference(x, B, S, Hq, Hkv, D, causal, scale)
                (dqkv,) = torch.autograd.grad(ref, x, do.float())
            dqkv = dqkv.to(qkv2.dtype)
        rope_inplace(dqkv, cos, sin, pos, Hq + Hkv, D, backward=True)
        return (dqkv.view(ctx.in_shape),) + (None,) * 10


def rope_attention(qkv, cos, sin, pos, B, S, Hq, Hkv, D, causal=True, scale=None):
    """``qkv``: output of the fused QKV projection, ``[B*S, (Hq+2Hkv)*D]`` (modified in place on GPU)."""
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    return _RopeAttentionFn.apply(qkv, cos, sin, pos, B, S, Hq, Hkv, D, causal, scale)


# ----------------------------------------------------------------------------------------- optimizer kernels
def adamw_flat_(master, m, v, grad, shadow, *, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0,
                clip_coef=None, found_inf=None):
    """Fused AdamW on flat fp32 state; writes the bf16 shadow parameter when given."""
    n = master.numel()
    if not _on_gpu(master):
        if found_inf is not None and float(found_inf) != 0.0:
            return
        g = grad.float() * grad_scale * (float(clip_coef) if clip_coef is not None else 1.0)
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        master.mul_(1 - lr * weight_decay)
        denom = (v / (1 - beta2**step)).sqrt_().add_(eps)
        master.addcdiv_(m, denom, value=-lr / (1 - beta1**step))
        if shadow is not None:
            shadow.copy_(master.to(shadow.dtype))
        return
    lib = _lib.require_gpu_lib()
    check(
        lib.b200_adamw_flat(ptr(master), ptr(m), ptr(v), ptr(grad), ptr(shadow), n, int(grad.dtype == BF16), float(lr), float(beta1),
                            float(beta2), float(eps), float(weight_decay), int(step), float(grad_scale), ptr(clip_coef), ptr(found_inf), cur_stream()),
        "adamw_flat",
    )


def sumsq_(grad: torch.Tensor, out2: torch.Tensor):
    """``out2[0] += sum(grad^2)``, ``out2[1] = 1`` if any element is non-finite."""
    if not _on_gpu(grad):
        g = grad.float()
        out2[0] += (g * g).sum()
        if not torch.isfinite(g).all():
            out2[1] = 1.0
        return
    lib = _lib.require_gpu_lib()
    g = grad if grad.is_contiguous() else grad.contiguous()
    check(lib.b200_sumsq(ptr(g), g.numel(), int(g.dtype == BF16), ptr(out2), cur_stream()), "sumsq")


def clip_coef_(sumsq: torch.Tensor, max_norm: float, norm_out: torch.Tensor, coef_out: torch.Tensor):
    if not _on_gpu(sumsq):
        norm = sumsq[0].sqrt()
        norm_out[0] = norm
        coef_out[0] = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
        return
    lib = _lib.require_gpu_lib()
    check(lib.b200_clip_coef(ptr(sumsq), float(max_norm), ptr(norm_out), ptr(coef_out), cur_stream()), "clip_coef")
