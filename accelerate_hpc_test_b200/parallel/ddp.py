"""``B200DataParallel``: replicated data parallelism with gradient buckets in peer-mapped memory.

What the reference does with ``torch.nn.parallel.DistributedDataParallel`` + the c10d Reducer +
``ncclAllReduce`` (reference src/accelerate/accelerator.py:1865-1869; knobs in
``DistributedDataParallelKwargs``, utils/dataclasses.py:155-237) is done here with:

* flat gradient buckets allocated in the :class:`~.symm.SymmetricWorkspace`; ``param.grad`` *is* a
  view of its bucket (``gradient_as_bucket_view`` semantics, always), and the native weight-gradient
  GEMMs write straight into that view (no copy, see ``ops.functional._LinearFn``);
* one fused kernel per bucket — sum over the 8 peers' buckets through NVLink loads, scale by
  ``1/world``, cast, store the result into every peer's bucket — launched on a communication stream
  as soon as the bucket's last gradient has been produced, i.e. overlapped with the rest of backward;
* ``no_sync()`` / ``require_backward_grad_sync`` with DDP's meaning.

Buckets are filled in reverse parameter order (≈ the order backward produces gradients).
"""

from __future__ import annotations

import contextlib
from typing import Optional

import os

import torch
import torch.nn as nn

from ..ops import functional as F
from ..utils.dataclasses import DDPCommunicationHookType, DistributedDataParallelKwargs
from .symm import SymmetricWorkspace


class _Bucket:
    def __init__(self, index: int):
        self.index = index
        self.params: list[nn.Parameter] = []
        self.offsets: list[int] = []
        self.numel = 0
        self.padded = 0
        self.offset_bytes = 0
        self.buffer: Optional[torch.Tensor] = None
        self.pending = 0
        self.ch_arrive = self.ch_done = -1
        self.launched = False
        self.updated = False
        self.done_event: Optional[torch.cuda.Event] = None


class B200DataParallel(nn.Module):
    """Wraps ``module``; forward is a passthrough, backward reduces gradients bucket by bucket."""

    def __init__(self, module: nn.Module, process_group=None, handler: Optional[DistributedDataParallelKwargs] = None,
                 gradient_state=None, workspace: Optional[SymmetricWorkspace] = None):
        super().__init__()
        self.module = module
        self.process_group = process_group if process_group is not None else torch.distributed.group.WORLD
        self.world = torch.distributed.get_world_size(self.process_group)
        self.rank = torch.distributed.get_rank(self.process_group)
        self.handler = handler or DistributedDataParallelKwargs()
        self.require_backward_grad_sync = True
        self.gradient_state = gradient_state
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise ValueError("B200DataParallel needs at least one parameter that requires grad")
        self.device = params[0].device
        if self.device.type != "cuda":
            raise ValueError("B200DataParallel runs on CUDA; CPU jobs use torch DistributedDataParallel over gloo")
        # replicas start identical: one broadcast per tensor at wrap time (reference: DDP does the same)
        with torch.no_grad():
            for t in list(module.parameters()) + (list(module.buffers()) if self.handler.broadcast_buffers else []):
                torch.distributed.broadcast(t.data, src=torch.distributed.get_global_rank(self.process_group, 0), group=self.process_group)
        wire = self.handler.comm_hook
        if wire in (DDPCommunicationHookType.POWER_SGD, DDPCommunicationHookType.BATCHED_POWER_SGD):
            raise NotImplementedError(
                f"comm_hook={wire} (low-rank gradient compression) is not implemented by the native data-parallel engine: its "
                "all-reduce already runs at NVLink speed inside one node. Use comm_hook NO / BF16, or set ACCELERATE_B200_DDP=0 to "
                "run torch DistributedDataParallel with the PowerSGD hook."
            )
        if wire == DDPCommunicationHookType.FP16:
            import warnings

            warnings.warn("comm_hook=FP16: the native engine reduces 16-bit gradients as bfloat16 (same bytes on the wire, no "
                          "overflow handling needed); pass BF16 to silence this.")
        self._wire_dtype = {DDPCommunicationHookType.BF16: torch.bfloat16, DDPCommunicationHookType.FP16: torch.bfloat16}.get(wire)
        # a parameter registered in several modules (tied embeddings) is handed one gradient per use
        self._uses: dict[int, int] = {}
        for _, q in module.named_parameters(remove_duplicate=False):
            self._uses[id(q)] = self._uses.get(id(q), 0) + 1
        total_bytes = sum(p.numel() * (2 if (self._wire_dtype or p.dtype) == torch.bfloat16 else 4) for p in params)
        if workspace is None:
            workspace = SymmetricWorkspace(self.process_group, total_bytes + 64 * 2**20, self.device)
            self._own_workspace = True
        else:
            self._own_workspace = False
        self.ws = workspace
        self.comm_stream = torch.cuda.Stream(device=self.device)
        # auto: in-switch reduction (NVLS) from 4 ranks up, where it beats NCCL by 1.16-1.37x; between two GPUs the switch has
        # nothing to save and the TMA two-phase kernel is the faster of ours (422 vs 390 GB/s)
        algo = os.environ.get("ACCELERATE_B200_ALLREDUCE", "auto")
        if algo == "auto":
            algo = "nvls" if self.world >= 4 else "tma"
        # CTAs of the in-switch all-reduce: it runs beside the backward GEMMs, so it should hold as few SM slots as still fill the links
        self._ar_ctas = int(os.environ.get("ACCELERATE_B200_DDP_AR_CTAS", "16"))
        self._tma_ctas = int(os.environ.get("ACCELERATE_B200_DDP_TMA_CTAS", "148"))
        self._tma_allreduce = algo in ("tma", "nvls")
        self._nvls_allreduce = algo == "nvls"
        self._buckets: list[_Bucket] = []
        self._param_slot: dict[int, tuple[_Bucket, int]] = {}
        self._fused_opt = None  # FusedAdamW that updates buckets right behind their all-reduce (bind_optimizer)
        self._build_buckets(params)
        self._callback_queued = False
        self._hooks = []
        self._install_hooks()

    # ------------------------------------------------------------------ buckets
    def _bucket_dtype(self, p):
        return self._wire_dtype or (torch.bfloat16 if p.dtype == torch.bfloat16 else torch.float32)

    def _build_buckets(self, params):
        cap = max(1, int(self.handler.bucket_cap_mb)) * 2**20
        cur: Optional[_Bucket] = None
        cur_dtype = None
        cur_bytes = 0
        for p in reversed(params):
            dt = self._bucket_dtype(p)
            esz = 2 if dt == torch.bfloat16 else 4
            if cur is None or dt != cur_dtype or (cur_bytes + p.numel() * esz > cap and cur.params):
                cur = _Bucket(len(self._buckets))
                cur.dtype = dt
                self._buckets.append(cur)
                cur_dtype, cur_bytes = dt, 0
            cur.params.append(p)
            cur.offsets.append(cur.numel)
            # keep every view 16-byte aligned so the GEMM epilogue can store vectors into it
            cur.numel += (p.numel() + 7) // 8 * 8
            cur_bytes += p.numel() * esz
        for b in self._buckets:
            epv = 8 if b.dtype == torch.bfloat16 else 4
            quantum = self.world * epv
            b.padded = (b.numel + quantum - 1) // quantum * quantum
            b.offset_bytes, b.buffer = self.ws.alloc_tensor((b.padded,), b.dtype)
            b.buffer.zero_()
            b.ch_arrive, b.ch_done = self.ws.new_channel(), self.ws.new_channel()
            b.ch_mid = self.ws.new_channel()  # between the two phases of the TMA all-reduce
            for p, off in zip(b.params, b.offsets):
                self._param_slot[id(p)] = (b, off)
                view = b.buffer[off : off + p.numel()].view(p.shape)
                if b.dtype == p.dtype:
                    p._b200_grad_buf = view  # native ops write their weight gradient here
                    p._b200_grad_hook = self._native_grad_ready
                p._b200_bucket_view = view

    def _install_hooks(self):
        for b in self._buckets:
            for p in b.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._autograd_grad_ready))

    # ------------------------------------------------------------------ gradient arrival
    def _begin_backward(self):
        if self._callback_queued:
            return
        self._callback_queued = True
        for b in self._buckets:
            b.pending = len(b.params)
            b.launched = False
            b._ready = set()
            b._deliveries = {}
        torch.autograd.Variable._execution_engine.queue_callback(self._finish_backward)

    def _native_grad_ready(self, p):
        """Called by native ops after they wrote (or accumulated) dW into ``p._b200_grad_buf``."""
        self._begin_backward()
        p.grad = p._b200_grad_buf
        b, _ = self._param_slot[id(p)]
        got = b._deliveries.get(id(p), 0) + 1
        b._deliveries[id(p)] = got
        if got >= self._uses.get(id(p), 1):  # the bucket may only be reduced once every use has added its share
            self._mark_ready(p)

    def _autograd_grad_ready(self, p):
        """Post-accumulate hook: gradients produced by foreign ops are moved into the bucket view."""
        self._begin_backward()
        view = p._b200_bucket_view
        g = p.grad
        if g is None:
            return  # a native op delivered this gradient straight into the bucket (and reported it)
        if g.data_ptr() != view.data_ptr():
            if view.dtype != p.dtype:
                # compressed wire (bf16 bucket, fp32 parameter): p.grad keeps accumulating in full precision across
                # no_sync micro-steps, the bucket is always a fresh cast of that running total
                F.scale_cast(g.contiguous().view(-1), view.view(-1), 1.0) if g.dtype in (torch.float32, torch.bfloat16) else view.copy_(g)
            elif getattr(p, "_b200_has_partial", False):
                view.add_(g.to(view.dtype))
                p.grad = view
            else:
                F.scale_cast(g.contiguous().view(-1), view.view(-1), 1.0) if g.dtype in (torch.float32, torch.bfloat16) else view.copy_(g)
                p.grad = view
        self._mark_ready(p)

    def _mark_ready(self, p):
        b, _ = self._param_slot[id(p)]
        if id(p) in b._ready:
            return
        b._ready.add(id(p))
        b.pending -= 1
        if b.pending == 0 and self._sync_now():
            self._launch(b)

    def _sync_now(self) -> bool:
        return self.require_backward_grad_sync

    def _launch(self, b: _Bucket):
        b.launched = True
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.comm_stream.wait_event(ev)
        with torch.cuda.stream(self.comm_stream):
            if b.dtype == torch.bfloat16 and self.ws.has_multicast and self._nvls_allreduce:
                # the NVSwitch sums the 8 copies and multicasts the result: one pass over the bucket, no shared-memory staging
                self.ws.allreduce_nvls_(b.offset_bytes, b.padded, 1.0 / self.world, b.ch_arrive, b.ch_done, ctas=self._ar_ctas, stream=self.comm_stream)
            elif b.dtype == torch.bfloat16 and self._tma_allreduce:
                # TMA engine end to end (reduce my slice from all peers, then pull the other slices): ~1.5x the LSU kernel
                self.ws.allreduce_tma_(b.offset_bytes, b.padded, 1.0 / self.world, (b.ch_arrive, b.ch_mid, b.ch_done), ctas=self._tma_ctas, stream=self.comm_stream)
            else:
                self.ws.allreduce_(b.offset_bytes, b.padded, b.dtype, 1.0 / self.world, b.ch_arrive, b.ch_done, stream=self.comm_stream)
            b.updated = False
            if self._fused_opt is not None and all(q.dtype == b.dtype for q in b.params):
                # the reduced gradients of this bucket are final: update its parameters right here, behind the all-reduce
                done = [self._fused_opt.update_reduced(q, q._b200_bucket_view.reshape(-1)) for q in b.params]
                b.updated = all(done)
            b.done_event = torch.cuda.Event()
            b.done_event.record(self.comm_stream)
        if b.updated:
            for q in b.params:
                q.grad = None  # spent

    def _finish_backward(self):
        self._callback_queued = False
        sync = self._sync_now()
        for b in self._buckets:
            for p in b.params:
                p._b200_grad_seen = False  # per-backward flag of ops.functional.deliver_weight_grad
        for b in self._buckets:
            if b.pending > 0:
                # parameters that took no part in this backward contribute zeros
                for p, off in zip(b.params, b.offsets):
                    if id(p) not in b._ready and not getattr(p, "_b200_has_partial", False):
                        b.buffer[off : off + p.numel()].zero_()
            if sync and not b.launched:
                self._launch(b)
        cur = torch.cuda.current_stream(self.device)
        for b in self._buckets:
            if sync:
                if b.done_event is not None:
                    cur.wait_event(b.done_event)
                for p in b.params:
                    p._b200_has_partial = False
                    if b.dtype != p.dtype and p.grad is not None:
                        F.scale_cast(p._b200_bucket_view.reshape(-1), p.grad.view(-1), 1.0)
            else:
                for p in b.params:
                    if id(p) in b._ready:
                        p._b200_has_partial = True

    # ------------------------------------------------------------------ public surface (DDP-compatible)
    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    @contextlib.contextmanager
    def trigger_sync_in_backward(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = True
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def bind_optimizer(self, optimizer):
        """The replicated update stays with the user's optimizer.  A ``FusedAdamW`` is additionally offered the chance to update
        every bucket's parameters right behind that bucket's all-reduce, on the communication stream (it decides by its own
        probing rules whether that is safe: no clipping, no loss scaler, one synchronised backward per step)."""
        inner = getattr(optimizer, "optimizer", optimizer)
        if self._fused_opt is None and hasattr(inner, "enable_update_after_reduce") and getattr(optimizer, "scaler", None) is None:
            gs = self.gradient_state
            sync = (lambda: gs.sync_gradients) if gs is not None else None
            if inner.enable_update_after_reduce(self.comm_stream, sync_gradients_fn=sync):
                self._fused_opt = inner
        return None

    def state_dict(self, *args, **kwargs):
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, *args, **kwargs):
        return self.module.load_state_dict(*args, **kwargs)

    def release(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []
        if self._own_workspace:
            self.ws.close()

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)
