"""Native FP8 training path (``mixed_precision="fp8"`` with ``FP8BackendType.NATIVE``).

The reference delegates fp8 to libraries that swap ``nn.Linear`` for their own modules (reference
src/accelerate/utils/transformer_engine.py:26-110 ``convert_model``, utils/ao.py:104-140
``convert_model_to_fp8_ao``, accelerator.py:1910-1940 back-end choice, first/last layer kept in high
precision by ``filter_first_and_last_linear_layers`` utils/ao.py:78-101).  Here the linear layers keep their
bf16 master weights and run

* forward  ``y = x W^T``     e4m3 x e4m3 on ``tcgen05.mma.kind::f8f6f4`` (csrc/gemm2_tcgen05.cu, FP8 variant),
* dgrad    ``dx = dy W``     e4m3 x e4m3 (the weight is quantised once per forward into ``W_q`` and ``W_q^T``),
* wgrad    ``dW = dy^T x``   bf16 (kept in high precision: it feeds the optimizer directly),

in one of two recipes:

* ``mxfp8`` (default): OCP microscaling — one UE8M0 power-of-two scale per 32 elements along the reduction dimension,
  applied inside the tensor core by ``tcgen05.mma.kind::mxf8f6f4.block_scale`` (csrc/gemm_mxfp8.cu; quantisers produce the
  scale bytes directly in the layout ``tcgen05.cp`` loads into TMEM);
* ``tensorwise``: one just-in-time scale per tensor (``amax`` pass + fused scale/cast/transpose, csrc/fp8.cu) on the
  2-CTA ``kind::f8f6f4`` GEMM.  Select with ``ACCELERATE_B200_FP8_RECIPE=tensorwise`` or
  ``TERecipeKwargs(use_mxfp8_block_scaling=False)``.

On CPU the same arithmetic is emulated with ``torch.float8_e4m3fn`` casts (the test oracle).
"""

from __future__ import annotations

from typing import Iterable, Optional

import torch
from torch import nn

from ..ops import _lib
from ..ops import functional as F

E4M3_MAX = 448.0
FP8 = torch.float8_e4m3fn


def quantize_e4m3(x: torch.Tensor, transpose: bool = False):
    """``x`` bf16 ``[R, C]`` -> ``(q [R,C] e4m3, q_t [C,R] or None, deq fp32 scalar)`` with ``x ~= q * deq``."""
    R, C = x.shape
    if not x.is_cuda:
        amax = x.float().abs().max().clamp(min=1e-12)
        scale = E4M3_MAX / amax
        q = (x.float() * scale).clamp(-E4M3_MAX, E4M3_MAX).to(FP8)
        return q, (q.t().contiguous() if transpose else None), (1.0 / scale).reshape(1)
    lib = _lib.require_gpu_lib()
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    if not x.is_contiguous():
        x = x.contiguous()
    amax = torch.zeros(1, dtype=torch.float32, device=x.device)
    _lib.check(lib.b200_amax_bf16(_lib.ptr(x), x.numel(), _lib.ptr(amax), _lib.cur_stream()), "amax")
    q = torch.empty((R, C), dtype=FP8, device=x.device)
    qt = torch.empty((C, R), dtype=FP8, device=x.device) if transpose else None
    deq = torch.empty(1, dtype=torch.float32, device=x.device)
    _lib.check(lib.b200_quantize_e4m3(_lib.ptr(x), x.stride(0), R, C, _lib.ptr(amax), _lib.ptr(q), q.stride(0), _lib.ptr(qt),
                                      qt.stride(0) if qt is not None else 0, _lib.ptr(deq), _lib.cur_stream()), "quantize_e4m3")
    return q, qt, deq


def gemm_e4m3(a_q, b_q, deq_a, deq_b, bias=None, out_dtype=torch.bfloat16, out=None, accumulate=False):
    """``D[M,N] = deq_a*deq_b * A_q[M,K] B_q[N,K]^T (+ bias)`` on the fp8 tensor-core path."""
    M, K = a_q.shape
    N = b_q.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a_q.device)
    if not a_q.is_cuda:
        r = (a_q.float() @ b_q.float().t()) * (deq_a.float() * deq_b.float())
        if bias is not None:
            r = r + bias.float()
        if accumulate:
            r = r + out.float()
        out.copy_(r.to(out.dtype))
        return out
    lib = _lib.require_gpu_lib()
    if K % 16:
        raise ValueError("fp8 GEMM needs the reduction dimension to be a multiple of 16")
    if bias is not None:
        bias = bias.to(torch.bfloat16).contiguous()
    _lib.check(lib.b200_gemm2_e4m3(_lib.ptr(a_q), _lib.ptr(b_q), _lib.ptr(out), _lib.ptr(bias), _lib.ptr(deq_a), _lib.ptr(deq_b),
                                   M, N, K, a_q.stride(0), b_q.stride(0), out.stride(0), int(out.dtype == torch.float32),
                                   int(accumulate), _lib.cur_stream()), "gemm2_e4m3")
    return out


# ------------------------------------------------------------------------------------------- MXFP8 (block scaled)
MX_BLOCK = 32


def _sf_bytes(rows: int, k: int) -> int:
    # whole 256-row tiles: the SM-pair kernel's tensor map covers both 128-row blocks of its last tile
    return ((rows + 255) // 256) * 2 * (k // 128) * 512


def _mx_emulate(x: torch.Tensor):
    """fp32 emulation of the OCP MX quantiser along the last dim: returns (q e4m3 [R, C], scale fp32 [R, C/32])."""
    R, C = x.shape
    g = x.float().view(R, C // MX_BLOCK, MX_BLOCK)
    amax = g.abs().amax(-1)
    e = torch.floor(torch.log2(amax.clamp(min=2.0**-126))) - 8.0
    e = torch.where(amax == 0, torch.full_like(e, -127.0), e).clamp(-127.0, 127.0)
    scale = torch.exp2(e)
    q = (g / scale[..., None]).clamp(-E4M3_MAX, E4M3_MAX).to(FP8).view(R, C)
    return q, scale


class MXTensor:
    """An e4m3 matrix ``[rows, k]`` with one power-of-two scale per 32 elements along ``k``.

    On the GPU ``sf`` is the UE8M0 byte buffer in the 512-byte block layout the tensor core reads (csrc/gemm_mxfp8.cu);
    on the CPU it is the fp32 scale matrix ``[rows, k/32]`` of the emulation."""

    __slots__ = ("q", "sf")

    def __init__(self, q, sf):
        self.q, self.sf = q, sf

    def dequantize(self) -> torch.Tensor:
        if self.q.is_cuda:
            raise NotImplementedError("dequantize() is the CPU oracle")
        R, C = self.q.shape
        return (self.q.float().view(R, C // MX_BLOCK, MX_BLOCK) * self.sf[..., None]).view(R, C)


def quantize_mx(x: torch.Tensor, transpose: bool = False) -> MXTensor:
    """MX-quantise ``x [R, C]`` along ``C``; with ``transpose`` the result is ``x^T [C, R]`` quantised along ``R``."""
    R, C = x.shape
    if not x.is_cuda:
        return MXTensor(*_mx_emulate(x.t().contiguous() if transpose else x))
    lib = _lib.require_gpu_lib()
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    if x.stride(1) != 1:
        x = x.contiguous()
    rows, k = (C, R) if transpose else (R, C)
    if k % 128:
        raise ValueError(f"MXFP8 needs the reduction dimension ({k}) to be a multiple of 128")
    q = torch.empty((rows, k), dtype=FP8, device=x.device)
    # rows past the end of the tensor keep whatever the allocator left there: they only ever scale zero-filled data
    # whose products are never stored
    sf = torch.empty(_sf_bytes(rows, k), dtype=torch.uint8, device=x.device)
    fn = lib.b200_quantize_mx_colwise if transpose else lib.b200_quantize_mx_rowwise
    _lib.check(fn(_lib.ptr(x), x.stride(0), R, C, _lib.ptr(q), q.stride(0), _lib.ptr(sf), _lib.cur_stream()), "quantize_mx")
    return MXTensor(q, sf)


def gemm_mxfp8(a: MXTensor, b: MXTensor, bias=None, out_dtype=torch.bfloat16, out=None, accumulate=False):
    """``D[M,N] = dequant(A)[M,K] dequant(B)[N,K]^T (+ bias)`` with the scales applied inside the tensor core."""
    M, K = a.q.shape
    N = b.q.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.q.device)
    if not a.q.is_cuda:
        r = a.dequantize() @ b.dequantize().t()
        if bias is not None:
            r = r + bias.float()
        if accumulate:
            r = r + out.float()
        out.copy_(r.to(out.dtype))
        return out
    lib = _lib.require_gpu_lib()
    if bias is not None:
        bias = bias.to(torch.bfloat16).contiguous()
    import os

    fn = lib.b200_gemm_mxfp8 if os.environ.get("ACCELERATE_B200_GEMM_2CTA", "1") == "0" or M <= 128 else lib.b200_gemm2_mxfp8
    _lib.check(fn(_lib.ptr(a.q), _lib.ptr(b.q), _lib.ptr(a.sf), _lib.ptr(b.sf), _lib.ptr(out), _lib.ptr(bias), M, N, K,
                                   a.q.stride(0), b.q.stride(0), out.stride(0), int(out.dtype == torch.float32), int(accumulate),
                                   _lib.cur_stream()), "gemm_mxfp8")
    return out


class _MxFp8LinearFn(torch.autograd.Function):
    """fwd and dgrad on the block-scaled fp8 tensor cores; wgrad in bf16 (see module docstring)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        shp = x.shape
        x2 = x.reshape(-1, shp[-1])
        cdt = torch.bfloat16 if x2.is_cuda else x2.dtype
        y = gemm_mxfp8(quantize_mx(x2), quantize_mx(weight.detach()), bias=bias, out_dtype=cdt)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.in_shape = shp
        return y.view(*shp[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if dy2.dtype != weight.dtype:
            dy2 = dy2.to(weight.dtype)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # dX[M,K] = dY[M,N] W[N,K]: reduction over N -> dY scaled along N, W^T [K,N] scaled along N
            dx = gemm_mxfp8(quantize_mx(dy2), quantize_mx(weight.detach(), transpose=True), out_dtype=dy2.dtype).view(ctx.in_shape)
        if ctx.needs_input_grad[1]:
            dw = F.deliver_weight_grad(
                weight, lambda out, acc: F.gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc)
            )
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.float().sum(0).to(dy2.dtype)
        return dx, dw, db


def mxfp8_linear(x, weight, bias=None):
    return _MxFp8LinearFn.apply(x, weight, bias)


class _Fp8LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        shp = x.shape
        x2 = x.reshape(-1, shp[-1])
        cdt = torch.bfloat16 if x2.is_cuda else x2.dtype
        xq, _, x_deq = quantize_e4m3(x2)
        need_dx = ctx.needs_input_grad[0]
        wq, wqt, w_deq = quantize_e4m3(weight.detach(), transpose=need_dx)
        y = gemm_e4m3(xq, wq, x_deq, w_deq, bias=bias, out_dtype=cdt)
        ctx.save_for_backward(x2, weight, wqt if wqt is not None else torch.empty(0), w_deq)
        ctx.has_bias = bias is not None
        ctx.in_shape = shp
        return y.view(*shp[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, wqt, w_deq = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if dy2.dtype != weight.dtype:
            dy2 = dy2.to(weight.dtype)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dyq, _, dy_deq = quantize_e4m3(dy2)
            dx = gemm_e4m3(dyq, wqt, dy_deq, w_deq, out_dtype=dy2.dtype).view(ctx.in_shape)
        if ctx.needs_input_grad[1]:
            dw = F.deliver_weight_grad(
                weight, lambda out, acc: F.gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc)
            )
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.float().sum(0).to(dy2.dtype)
        return dx, dw, db


def fp8_linear(x, weight, bias=None):
    return _Fp8LinearFn.apply(x, weight, bias)


def _eligible(mod: nn.Module, multiple: int = 16) -> bool:
    w = getattr(mod, "weight", None)
    return w is not None and w.dim() == 2 and w.shape[0] % multiple == 0 and w.shape[1] % multiple == 0


def _recipe_name(recipe) -> str:
    """``mxfp8`` (1x32 block scales, the default) or ``tensorwise`` (one just-in-time scale per tensor)."""
    import os

    env = os.environ.get("ACCELERATE_B200_FP8_RECIPE")
    if env:
        return env.lower()
    if recipe is not None and getattr(recipe, "use_mxfp8_block_scaling", None) is False:
        return "tensorwise"
    return "mxfp8"


def convert_model_to_mxfp8(model: nn.Module, recipe=None, skip: Optional[Iterable[str]] = None) -> nn.Module:
    """Route the linear layers of ``model`` through the fp8 GEMM path (in place; parameters are untouched).

    Like the reference's torchao conversion the first and last linear layers stay in bf16 unless the recipe says
    otherwise (they are the most sensitive ones), as does anything whose dimensions are not multiples of 16
    (reference utils/ao.py:78-101, utils/transformer_engine.py:44-60)."""
    from ..models.llama import B200Linear

    skip = set(skip or ())
    linears = [(n, m) for n, m in model.named_modules() if isinstance(m, (B200Linear, nn.Linear))]
    if not linears:
        return model
    keep_ends = True
    if recipe is not None and getattr(recipe, "use_during_eval", None) is not None:
        model._fp8_use_during_eval = bool(recipe.use_during_eval)
    if keep_ends and len(linears) > 2:
        skip.update({linears[0][0], linears[-1][0]})
    kind = _recipe_name(recipe)
    converted = 0
    for name, mod in linears:
        if name in skip or name.endswith("lm_head") or not _eligible(mod, 128 if kind == "mxfp8" else 16):
            continue
        _patch_forward(mod, kind)
        converted += 1
    model._b200_fp8_layers = converted
    model._b200_fp8_recipe = kind
    return model


def _patch_forward(mod: nn.Module, kind: str = "mxfp8"):
    if getattr(mod, "_b200_fp8", False):
        return
    mod._b200_fp8 = True
    fn = mxfp8_linear if kind == "mxfp8" else fp8_linear

    def forward(x, _m=mod, _fn=fn):
        return _fn(x, _m.weight, _m.bias)

    mod.forward = forward


def has_fp8_layers(model: nn.Module) -> bool:
    return any(getattr(m, "_b200_fp8", False) for m in model.modules())
