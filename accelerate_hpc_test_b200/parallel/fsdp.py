"""``B200ShardedModule``: fully sharded data parallelism (FSDP2 semantics) on peer-mapped memory.

The reference hands the model to torch's ``fully_shard`` (reference src/accelerate/utils/fsdp_utils.py:
621-737) and lets FSDP2 issue ``ncclAllGather`` / ``ncclReduceScatter`` per wrapped module.  Here:

* every wrapped unit (the modules picked by the plugin's auto-wrap policy, plus one unit per large
  leftover tensor group) owns a flat bf16 parameter buffer; each rank keeps ``1/world`` of it — its
  *shard* — in the symmetric workspace, together with local fp32 master / Adam moments;
* **all-gather** is a *pull* kernel: a rank reads its peers' shards over NVLink into a pooled full
  buffer, on the communication stream, one unit ahead of the compute stream (forward and backward
  prefetch).  No rank-to-rank handshake is needed per gather: shards only change in the optimizer
  step, which ends with one cross-rank barrier;
* **reduce-scatter + AdamW** is one kernel: when a unit's backward is complete, each rank sums its
  shard of the 8 flat gradient buffers through NVLink loads, applies AdamW to its fp32 shard and
  writes the new bf16 shard — overlapped with the backward of the next unit.  If gradients must
  survive until ``optimizer.step()`` (gradient clipping, foreign optimizers) the same kernel only
  reduces (``do_adam=0``) and the sharded update runs at ``step()``;
* native weight-gradient GEMMs write straight into the unit's flat gradient buffer.

On CPU (gloo) the same engine runs with c10d collectives — that path is the test oracle.
Checkpoint layout: ``pytorch_model_fsdp_{i}/`` and ``optimizer_{i}/`` directories (SHARDED) or
``pytorch_model_fsdp.bin`` / ``optimizer.bin`` (FULL), as in reference utils/fsdp_utils.py:103-335.
"""

from __future__ import annotations

import contextlib
import json
import os
from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as F
from ..utils.constants import FSDP_MODEL_NAME, OPTIMIZER_NAME
from ..utils.other import get_module_children_bottom_up

BF16 = torch.bfloat16


def apply_activation_checkpointing(model: nn.Module, plugin) -> None:
    """Recompute the wrapped units in backward (reference utils/fsdp_utils.py:588-618)."""
    from torch.utils.checkpoint import checkpoint

    for module in model.modules():
        if module is not model and plugin.should_wrap(module) and not getattr(module, "_b200_checkpointed", False):
            orig = module.forward

            def fwd(*args, _orig=orig, **kwargs):
                if torch.is_grad_enabled():
                    return checkpoint(_orig, *args, use_reentrant=False, **kwargs)
                return _orig(*args, **kwargs)

            module.forward = fwd
            module._b200_checkpointed = True


class _Pool:
    """Equal-sized reusable buffers with the CUDA event that must pass before each can be reused."""

    def __init__(self, make):
        self._make = make  # () -> (symmetric byte offset, tensor)
        self.bufs = []
        self.events = []
        self.free = []

    def add(self, n: int = 1):
        for _ in range(n):
            self.bufs.append(self._make())
            self.events.append(None)
            self.free.append(len(self.bufs) - 1)

    def take(self) -> int:
        if not self.free:
            self.add(1)  # grow rather than dead-lock (deep prefetch / many live units)
        return self.free.pop(0)

    def give(self, idx: int, event=None):
        self.events[idx] = event
        self.free.append(idx)


from ..utils.diagnostics import shadow_check


class _Unit:
    """One sharded group of parameters (a wrapped module, or a leftover group of the root)."""

    def __init__(self, name: str, module: Optional[nn.Module], params: list, names: list):
        self.name = name
        self.module = module
        self.params = params
        self.param_names = names
        self.offsets = []
        n = 0
        for p in params:
            self.offsets.append(n)
            n += (p.numel() + 7) // 8 * 8
        self.numel = n
        self.padded = 0
        self.shard_numel = 0
        self.full: Optional[torch.Tensor] = None  # pooled full bf16 buffer while unsharded
        self.grad: Optional[torch.Tensor] = None  # pooled flat gradient buffer during backward
        self.grad_off = 0
        self.ag_event = None
        self.free_event = None
        self.unsharded = False
        self.in_backward = False
        self.ready = set()
        self.has_partial = False
        self.rs_launched = False


class B200ShardedModule(nn.Module):
    def __init__(self, module: nn.Module, plugin, device, shard_group=None, replicate_group=None, gradient_state=None,
                 shard_group_is_world: bool = True):
        super().__init__()
        self.module = module
        self.plugin = plugin
        self.device = torch.device(device)
        self.gradient_state = gradient_state
        self.group = shard_group if shard_group is not None else torch.distributed.group.WORLD
        from ..utils.diagnostics import CommTrace, comm_poison_enabled, comm_shadow_enabled

        self.trace = CommTrace()  # exposed-communication accounting (ACCELERATE_B200_TRACE_COMM=1 or .enabled = True)
        self._shadow, self._poison = comm_shadow_enabled(), comm_poison_enabled()
        self.replicate_group = replicate_group
        self.world = torch.distributed.get_world_size(self.group)
        self.rank = torch.distributed.get_rank(self.group)
        self.use_symm = self.device.type == "cuda"
        self.reshard_after_forward = plugin.effective_reshard_after_forward
        self.requires_gradient_sync = True
        self.param_dtype = BF16 if self.use_symm else next(module.parameters()).dtype
        pol = plugin.mixed_precision_policy
        if pol is not None and pol.param_dtype is not None and self.use_symm and pol.param_dtype != BF16:
            raise ValueError("the native sharded engine computes in bf16; use mixed_precision='bf16'")
        self.step_count = 0
        self._optimizer = None
        self._hyper = None
        self._fused = False
        self._hyper_ok = False
        self._fused_allowed = False
        self._clip_requested_this_step = False
        self._saw_clip = False
        self._clip_coef = None
        self._callback_queued = False
        self._units: list[_Unit] = []
        self._order: list[_Unit] = []  # forward execution order, learnt on the first iteration
        self._order_locked = False
        self._build_units()
        self._allocate()
        self._install_hooks()

    # ------------------------------------------------------------------ construction
    def _build_units(self):
        plugin = self.plugin
        claimed = set()
        ignored = set()
        if plugin.ignored_modules is not None and not isinstance(plugin.ignored_modules, str):
            for m in plugin.ignored_modules:
                ignored.update(id(p) for p in m.parameters())
        names = {id(p): n for n, p in self.module.named_parameters()}
        if plugin.activation_checkpointing:
            apply_activation_checkpointing(self.module, plugin)
        for fqn, sub in get_module_children_bottom_up(self.module, return_fqns=True):
            if sub is self.module or not plugin.should_wrap(sub):
                continue
            ps = [p for p in sub.parameters() if id(p) not in claimed and id(p) not in ignored and p.requires_grad]
            if not ps:
                continue
            claimed.update(id(p) for p in ps)
            self._units.append(_Unit(fqn, sub, ps, [names[id(p)] for p in ps]))
        # leftovers of the root: big tensors (embedding, LM head) get their own unit so the pooled
        # buffers are sized by the largest *layer*, not by the whole root
        rest = [p for p in self.module.parameters() if id(p) not in claimed and id(p) not in ignored and p.requires_grad]
        seen = set()
        rest = [p for p in rest if not (id(p) in seen or seen.add(id(p)))]
        big = [p for p in rest if p.numel() >= 2**24]
        small = [p for p in rest if p.numel() < 2**24]
        root_units = []
        for p in big:
            root_units.append(_Unit(f"<root:{names[id(p)]}>", None, [p], [names[id(p)]]))
        if small:
            root_units.append(_Unit("<root>", None, small, [names[id(p)] for p in small]))
        self._root_units = root_units
        self._units = self._units + root_units
        self._unit_of_param = {}
        for u in self._units:
            for p in u.params:
                self._unit_of_param[id(p)] = u
                p._b200_fsdp_unit = u
                u.engine = self
        self._ignored_params = [p for p in self.module.parameters() if id(p) in ignored]

    def _allocate(self):
        world = self.world
        for u in self._units:
            q = world * 8
            u.padded = (u.numel + q - 1) // q * q
            u.shard_numel = u.padded // world
        layer_units = [u for u in self._units if u.module is not None]
        root_units = [u for u in self._units if u.module is None]
        max_layer = max([u.padded for u in layer_units], default=0)
        total_shard = sum(u.shard_numel for u in self._units)
        esz = 2 if self.param_dtype == BF16 else 4
        n_grad_bufs = int(os.environ.get("ACCELERATE_B200_FSDP_GRAD_BUFFERS", "3"))
        n_full_bufs = int(os.environ.get("ACCELERATE_B200_FSDP_PARAM_BUFFERS", "3"))
        if self.use_symm:
            from .symm import SymmetricWorkspace

            # shards + dedicated root gradient buffers + pooled layer gradient buffers (+ slack for pool growth)
            need = total_shard * esz + sum(u.padded for u in root_units) * esz + (n_grad_bufs + 2) * max_layer * esz + 64 * 2**20
            self.ws = SymmetricWorkspace(self.group, need, self.device)
            # gathers and reductions get their own streams: a reduce-scatter waits for its peers, and a
            # prefetching all-gather must not queue behind that wait
            self.comm_stream = torch.cuda.Stream(device=self.device)
            self.rs_stream = torch.cuda.Stream(device=self.device)
            self.ag_ctas = int(os.environ.get("ACCELERATE_B200_AG_CTAS", "148"))
            self.rs_ctas = int(os.environ.get("ACCELERATE_B200_RS_CTAS", "148"))
            self.ch_step = self.ws.new_channel()
        else:
            self.ws = None
            self.comm_stream = None
            self.rs_stream = None
        # shards: bf16 parameter shard (symmetric), fp32 master + moments (local)
        for u in self._units:
            flat = torch.zeros(u.padded, dtype=self.param_dtype, device=self.device)
            for p, off in zip(u.params, u.offsets):
                flat[off : off + p.numel()].copy_(p.detach().reshape(-1).to(self.param_dtype))
            # replicas must start identical: rank 0's values win (reference: fsdp2_load_full_state_dict)
            torch.distributed.broadcast(flat, src=torch.distributed.get_global_rank(self.group, 0), group=self.group)
            lo = self.rank * u.shard_numel
            if self.use_symm:
                u.shard_off, u.shard = self.ws.alloc_tensor((u.shard_numel,), self.param_dtype)
                u.shard.copy_(flat[lo : lo + u.shard_numel])
                u.ch_arrive, u.ch_done = self.ws.new_channel(), self.ws.new_channel()
            else:
                u.shard = flat[lo : lo + u.shard_numel].clone()
            u.master = u.shard.to(torch.float32)
            u.exp_avg = torch.zeros_like(u.master)
            u.exp_avg_sq = torch.zeros_like(u.master)
            u.grad_shard = None
            del flat
            for p in u.params:
                p._b200_full_shape = tuple(p.shape)
                p.data = torch.empty(0, dtype=self.param_dtype, device=self.device)
        # layer units share pools sized by the largest layer; root units (live for the whole
        # forward+backward) get dedicated buffers of their own size
        def full_maker(n):
            return lambda: (0, torch.empty(n, dtype=self.param_dtype, device=self.device))

        def grad_maker(n):
            if self.use_symm:
                return lambda: self.ws.alloc_tensor((n,), self.param_dtype)
            return lambda: (0, torch.empty(n, dtype=self.param_dtype, device=self.device))

        layer_full, layer_grad = _Pool(full_maker(max_layer)), _Pool(grad_maker(max_layer))
        if layer_units:
            layer_full.add(n_full_bufs)
            layer_grad.add(n_grad_bufs)
        for u in self._units:
            if u.module is not None:
                u.fpool, u.gpool = layer_full, layer_grad
            else:
                u.fpool, u.gpool = _Pool(full_maker(u.padded)), _Pool(grad_maker(u.padded))
                u.fpool.add(1)
                u.gpool.add(1)
        if self.use_symm:
            torch.cuda.synchronize(self.device)
            self.ws.barrier(self.ch_step)
            torch.cuda.synchronize(self.device)

    def _install_hooks(self):
        self._handles = []
        for u in self._units:
            if u.module is not None:
                self._handles.append(u.module.register_forward_pre_hook(lambda m, a, kw, _u=u: self._pre_forward(_u, a, kw), with_kwargs=True))
                self._handles.append(u.module.register_forward_hook(lambda m, a, out, _u=u: self._post_forward(_u, out)))
            for p in u.params:
                p._b200_grad_hook = self._native_grad_ready
                self._handles.append(p.register_post_accumulate_grad_hook(self._autograd_grad_ready))

    # ------------------------------------------------------------------ all-gather / reshard
    def _launch_gather(self, u: _Unit):
        """Start (on the comm stream) the gather of ``u`` into a pooled buffer; idempotent."""
        if u.full is not None:
            return
        idx = u.fpool.take()
        u.full_idx = idx
        u.full = u.fpool.bufs[idx][1][: u.padded]
        if self.use_symm:
            ev = u.fpool.events[idx]
            if ev is not None:
                self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                self.ws.allgather_pull(u.shard_off, u.shard_numel * 2, u.full, ctas=self.ag_ctas, stream=self.comm_stream)
                u.ag_event = torch.cuda.Event()
                u.ag_event.record(self.comm_stream)
        else:
            parts = [torch.empty_like(u.shard) for _ in range(self.world)]
            torch.distributed.all_gather(parts, u.shard, group=self.group)
            u.full.copy_(torch.cat(parts))

    def unshard(self, u: Optional[_Unit] = None):
        """Make the parameters of ``u`` (or of every unit) whole: ``param.data`` becomes a view of the gathered buffer."""
        if u is None:
            for x in self._units:
                self.unshard(x)
            return
        if u.unsharded:
            return
        self._launch_gather(u)
        if self.use_symm and u.ag_event is not None:
            self.trace.wait_event(torch.cuda.current_stream(self.device), u.ag_event, "all_gather")
            if self._shadow:
                # debug: the same gather through torch.distributed must give the same bytes
                ref = torch.empty_like(u.full[: self.world * u.shard_numel])
                torch.distributed.all_gather_into_tensor(ref, u.shard, group=self.group)
                shadow_check(f"all_gather[{u.name}]", u.full[: self.world * u.shard_numel], ref)
        for p, off in zip(u.params, u.offsets):
            p.data = u.full[off : off + _numel(p._b200_full_shape)].view(p._b200_full_shape)
        u.unsharded = True

    def reshard(self, u: Optional[_Unit] = None, write_back: bool = False):
        if u is None:
            for x in self._units:
                self.reshard(x, write_back)
            return
        if not u.unsharded:
            return
        if write_back:
            lo = self.rank * u.shard_numel
            u.shard.copy_(u.full[lo : lo + u.shard_numel])
            u.master.copy_(u.shard)
        for p in u.params:
            p.data = torch.empty(0, dtype=self.param_dtype, device=self.device)
        ev = None
        if self._poison:
            u.full.fill_(float("nan"))  # debug: anything that still reads the released buffer turns into NaN
        if self.use_symm:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
        u.fpool.give(u.full_idx, ev)
        u.full = None
        u.ag_event = None
        u.unsharded = False

    # ------------------------------------------------------------------ forward / backward hooks
    def _next_in_order(self, u: _Unit, step: int) -> Optional[_Unit]:
        if not self._order_locked or u not in self._order:
            return None
        i = self._order.index(u) + step
        return self._order[i] if 0 <= i < len(self._order) else None

    def _pre_forward(self, u: _Unit, args, kwargs):
        if not self._order_locked and u not in self._order:
            self._order.append(u)
        self.unshard(u)
        nxt = self._next_in_order(u, +1)
        if nxt is not None and not u.in_backward:
            self._launch_gather(nxt)
        if torch.is_grad_enabled() and not u.in_backward:
            self._register_pre_backward(u, args, kwargs)
        return None

    def _register_pre_backward(self, u: _Unit, args, kwargs):
        """Arrange for ``_pre_backward(u)`` to run when gradients reach the unit's outputs."""
        u._want_pre_backward = True

    def _post_forward(self, u: _Unit, out):
        if torch.is_grad_enabled() and getattr(u, "_want_pre_backward", False) and not u.in_backward:
            u._want_pre_backward = False
            tensors = [t for t in _flatten(out) if isinstance(t, torch.Tensor) and t.requires_grad]
            if tensors:
                torch.autograd.graph.register_multi_grad_hook(tensors, lambda *_g, _u=u: self._pre_backward(_u), mode="any")
        if u.in_backward:
            return None  # recomputation inside backward: the parameters are needed right after
        if self.reshard_after_forward and u.module is not None:
            self.reshard(u)
        return None

    def _queue_final_callback(self):
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finish_backward)

    def _pre_backward(self, u: _Unit):
        if u.in_backward:
            return
        self._queue_final_callback()
        u.in_backward = True
        self.unshard(u)
        prv = self._next_in_order(u, -1)
        if prv is not None and not prv.in_backward:
            self._launch_gather(prv)
        self._attach_grad_buffer(u)

    def _attach_grad_buffer(self, u: _Unit):
        if u.grad is None:
            idx = u.gpool.take()
            u.grad_idx = idx
            u.grad_off, buf = u.gpool.bufs[idx]
            u.grad = buf[: u.padded]
            if self.use_symm and u.gpool.events[idx] is not None:
                self.trace.wait_event(torch.cuda.current_stream(self.device), u.gpool.events[idx], "grad_buffer_reuse")
            u.has_partial = False
        u.ready = set()
        u.rs_launched = False
        for p, off in zip(u.params, u.offsets):
            p._b200_grad_buf = u.grad[off : off + _numel(p._b200_full_shape)].view(p._b200_full_shape)
            p._b200_has_partial = u.has_partial

    def _native_grad_ready(self, p):
        u = self._unit_of_param[id(p)]
        self._mark_ready(u, p)

    def _autograd_grad_ready(self, p):
        """Gradient produced by a foreign op landed in ``p.grad``: move it into the flat buffer."""
        u = self._unit_of_param[id(p)]
        if u.grad is None:
            # root-level parameters have no pre-backward hook of their own
            self._queue_final_callback()
            u.in_backward = True
            self._attach_grad_buffer(u)
        g = p.grad
        if g is not None:
            view = p._b200_grad_buf
            if u.has_partial or id(p) in u.ready:
                view.add_(g.to(view.dtype))
            else:
                view.copy_(g)
            p.grad = None
        self._mark_ready(u, p)

    def _mark_ready(self, u: _Unit, p):
        u.ready.add(id(p))
        if len(u.ready) == len(u.params) and u.module is not None:
            self._post_backward(u)

    def _post_backward(self, u: _Unit):
        """All gradients of ``u`` are in its flat buffer: reduce (+ update) and let go of the parameters."""
        if u.rs_launched:
            return
        u.rs_launched = True
        if u.grad is not None:
            missing = [i for i, p in enumerate(u.params) if id(p) not in u.ready]
            if missing and not u.has_partial:
                for i in missing:
                    off = u.offsets[i]
                    u.grad[off : off + _numel(u.params[i]._b200_full_shape)].zero_()
            if self.requires_gradient_sync:
                self._reduce_model_parallel_replicas(u)
                self._reduce_unit(u)
                self._release_grad(u)
            else:
                u.has_partial = True
        if u.unsharded and (self.reshard_after_forward or u.module is None or True):
            self.reshard(u)
        u.in_backward = False

    def _reduce_model_parallel_replicas(self, u: _Unit):
        """Parameters that a tensor-parallel group keeps whole (norms, LM head) only saw that rank's token shard: their
        gradients are summed over the tp group before the data-parallel reduce-scatter (tensor_parallel._mark_replicated
        leaves this to the engine when the parameter belongs to a sharded unit)."""
        for p, off in zip(u.params, u.offsets):
            g = getattr(p, "_b200_tp_group", None)
            if g is None or torch.distributed.get_world_size(g) == 1:
                continue
            view = u.grad[off : off + _numel(p._b200_full_shape)]
            torch.distributed.all_reduce(view, group=g)

    def _release_grad(self, u: _Unit):
        for p in u.params:
            p._b200_grad_buf = None
            p._b200_has_partial = False
        u.gpool.give(u.grad_idx, u.rs_event if self.use_symm else None)
        u.grad = None
        u.has_partial = False

    def _finish_backward(self):
        self._callback_queued = False
        if self._order and not self._order_locked:
            self._order_locked = True
        for u in self._units:
            if u.grad is not None and not u.rs_launched:
                self._post_backward(u)
            elif u.in_backward:
                if u.unsharded:
                    self.reshard(u)
                u.in_backward = False
        if self.use_symm:
            cur = torch.cuda.current_stream(self.device)
            for u in self._units:
                ev = getattr(u, "rs_event", None)
                if ev is not None:
                    self.trace.wait_event(cur, ev, "reduce_scatter_tail")

    # ------------------------------------------------------------------ reduce-scatter (+ AdamW)
    def _adam_hyper(self):
        opt = self._optimizer
        if opt is None:
            return None
        g = opt.param_groups[0]
        b1, b2 = g.get("betas", (0.9, 0.999))
        return dict(lr=float(g["lr"]), beta1=float(b1), beta2=float(b2), eps=float(g.get("eps", 1e-8)), weight_decay=float(g.get("weight_decay", 0.0)))

    def _reduce_unit(self, u: _Unit):
        fused = self._fused and self._hyper_ok
        scale = 1.0 / self.world
        if self.use_symm:
            from ..ops._lib import check, ptr
            import ctypes

            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.rs_stream.wait_event(ev)
            e = self.ws.next_epoch(u.ch_arrive)
            self.ws._epochs[u.ch_done] = e
            if not fused and u.grad_shard is None:
                u.grad_shard = torch.empty(u.shard_numel, dtype=torch.float32, device=self.device)
            h = self._adam_hyper() if fused else dict(lr=0.0, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0)
            with torch.cuda.stream(self.rs_stream):
                check(self.ws.lib.b200_reduce_scatter_adamw(
                    self.ws.peer_array, self.rank, self.world, u.grad_off, u.shard_numel, ptr(u.master), ptr(u.exp_avg),
                    ptr(u.exp_avg_sq), u.shard_off, ptr(u.grad_shard) if not fused else None, h["lr"], h["beta1"], h["beta2"], h["eps"],
                    h["weight_decay"], self.step_count + 1, scale, int(fused), None, u.ch_arrive, u.ch_done, e,
                    self.ws.counter_ptr(u.ch_arrive), self.rs_ctas, ctypes.c_void_p(self.rs_stream.cuda_stream)), "reduce_scatter_adamw")
                u.rs_event = torch.cuda.Event()
                u.rs_event.record(self.rs_stream)
            u.updated_in_backward = fused
        else:
            full = u.grad.to(torch.float32)
            torch.distributed.all_reduce(full, group=self.group)
            lo = self.rank * u.shard_numel
            u.grad_shard = full[lo : lo + u.shard_numel] * scale
            u.updated_in_backward = False
        if self.replicate_group is not None and not fused:
            stream_ctx = torch.cuda.stream(self.rs_stream) if self.use_symm else contextlib.nullcontext()
            with stream_ctx:
                torch.distributed.all_reduce(u.grad_shard, group=self.replicate_group)
                u.grad_shard /= torch.distributed.get_world_size(self.replicate_group)
                if self.use_symm:
                    u.rs_event = torch.cuda.Event()
                    u.rs_event.record(self.rs_stream)

    # ------------------------------------------------------------------ optimizer integration
    def bind_optimizer(self, accelerated_optimizer):
        """Take over the parameter update of ``accelerated_optimizer`` (sharded state lives in this engine)."""
        inner = accelerated_optimizer.optimizer
        self._optimizer = inner
        accelerated_optimizer.attach_engine(self)
        name = type(inner).__name__
        g = inner.param_groups[0] if inner.param_groups else {}
        self._hyper_ok = name in ("AdamW", "FusedAdamW", "Adam") and not g.get("amsgrad", False) and not g.get("maximize", False) and len(inner.param_groups) == 1
        if not self._hyper_ok:
            raise NotImplementedError(
                f"the native sharded engine implements (fused) AdamW on its shards; got optimizer {name} with {len(inner.param_groups)} param groups. "
                "Use torch.optim.AdamW / FusedAdamW with a single param group."
            )
        if name == "Adam":
            g.setdefault("weight_decay", 0.0)
        allow = os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto")
        self._fused_allowed = allow != "0" and self.use_symm and self.replicate_group is None
        self._fused = allow == "1" and self._fused_allowed
        # the user's optimizer never sees real tensors again: drop its (empty) state
        inner.state.clear()

    def zero_grad(self):
        self._clip_requested_this_step = False

    def clip_grad_norm_(self, max_norm, norm_type=2):
        """Global gradient norm over all shards; the clip coefficient is applied inside the update kernel."""
        if norm_type != 2:
            raise NotImplementedError("only the L2 norm is implemented for sharded gradients")
        self._clip_requested_this_step = True
        self._saw_clip = True
        if any(getattr(u, "updated_in_backward", False) for u in self._units):
            raise RuntimeError(
                "clip_grad_norm_ was called after the fused reduce-scatter+AdamW already consumed the gradients in backward. "
                "Set ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD=0 when clipping is only used on some steps."
            )
        acc = torch.zeros(2, dtype=torch.float32, device=self.device)
        ctx = torch.cuda.stream(self.rs_stream) if self.use_symm else contextlib.nullcontext()
        if self.use_symm:
            self.rs_stream.wait_stream(torch.cuda.current_stream(self.device))
        with ctx:
            for u in self._units:
                if u.grad_shard is not None:
                    F.sumsq_(u.grad_shard, acc)
            torch.distributed.all_reduce(acc, group=self.group)
            norm = torch.empty(1, dtype=torch.float32, device=self.device)
            coef = torch.ones(1, dtype=torch.float32, device=self.device)
            F.clip_coef_(acc, float(max_norm), norm, coef)
        if self.use_symm:
            torch.cuda.current_stream(self.device).wait_stream(self.rs_stream)
        self._clip_coef = coef
        return norm.reshape(())

    def optimizer_step(self, optimizer, scaler=None) -> bool:
        """Called by ``AcceleratedOptimizer.step``. Returns True if the step was skipped (never, for bf16)."""
        h = self._adam_hyper()
        self.step_count += 1
        ctx = torch.cuda.stream(self.rs_stream) if self.use_symm else contextlib.nullcontext()
        if self.use_symm:
            self.rs_stream.wait_stream(torch.cuda.current_stream(self.device))
        with ctx:
            for u in self._units:
                if getattr(u, "updated_in_backward", False):
                    u.updated_in_backward = False
                    continue
                if u.grad_shard is None:
                    continue
                F.adamw_flat_(u.master, u.exp_avg, u.exp_avg_sq, u.grad_shard, u.shard if u.shard.dtype == BF16 else None,
                              step=self.step_count, clip_coef=self._clip_coef, **h)
                if u.shard.dtype != BF16:
                    u.shard.copy_(u.master)
            if self.use_symm:
                # every rank's shards now hold version `step_count`: one barrier, then gathers need no handshake
                self.ws.barrier(self.ch_step, stream=self.rs_stream)
                self.comm_stream.wait_stream(self.rs_stream)
        self._clip_coef = None
        # first step ran in the safe mode; switch to the fused update if nobody asked for the gradients
        if self._fused_allowed and not self._fused and not self._saw_clip and os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto") == "auto":
            self._fused = True
            for u in self._units:
                u.grad_shard = None
        return False

    # ------------------------------------------------------------------ module surface
    def forward(self, *args, **kwargs):
        # root-level parameters (embedding, head, final norm) are gathered for the whole forward+backward
        for u in self._root_units:
            self.unshard(u)
            if torch.is_grad_enabled():
                self._queue_root_backward(u)
        if self._order_locked and self._order:
            self._launch_gather(self._order[0])
        out = self.module(*args, **kwargs)
        if not torch.is_grad_enabled():
            for u in self._root_units:
                self.reshard(u)
        return out

    def _queue_root_backward(self, u: _Unit):
        u.in_backward = True
        self._attach_grad_buffer(u)
        u._root_pending = True

    def set_requires_gradient_sync(self, flag: bool):
        self.requires_gradient_sync = bool(flag)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.requires_gradient_sync
        self.requires_gradient_sync = False
        try:
            yield
        finally:
            self.requires_gradient_sync = old

    def parameters_full(self):
        """Context-free helper for tools: un-shard everything and return the module."""
        self.unshard()
        return self.module

    # ------------------------------------------------------------------ state dicts / checkpoints
    def _unit_tensors(self, u: _Unit, flat: torch.Tensor):
        return {n: flat[off : off + _numel(p._b200_full_shape)].view(p._b200_full_shape) for n, p, off in zip(u.param_names, u.params, u.offsets)}

    def exposed_comm_ms(self, by_tag: bool = False):
        """Milliseconds (since the last call) the compute stream spent waiting for all-gathers / reduce-scatters that were not
        hidden behind compute.  Needs ``self.trace.enabled = True`` (or ACCELERATE_B200_TRACE_COMM=1); synchronises."""
        return self.trace.exposed_ms(by_tag=by_tag)

    def full_state_dict(self, cpu: bool = True) -> dict:
        """Gather every unit (one at a time) and return name -> full tensor (on CPU by default)."""
        out = {}
        for u in self._units:
            was = u.unsharded
            self.unshard(u)
            if self.use_symm:
                torch.cuda.current_stream(self.device).synchronize()
            for n, t in self._unit_tensors(u, u.full).items():
                out[n] = t.detach().to("cpu", copy=True) if cpu else t.detach().clone()
            if not was:
                self.reshard(u)
        for n, b in self.module.named_buffers():
            out[n] = b.detach().cpu() if cpu else b.detach()
        for p in self._ignored_params:
            for n, q in self.module.named_parameters():
                if q is p:
                    out[n] = p.detach().cpu() if cpu else p.detach()
        return out

    def state_dict(self, *args, **kwargs):
        return self.full_state_dict()

    @torch.no_grad()
    def load_full_state_dict(self, sd: dict):
        for u in self._units:
            flat = torch.zeros(u.padded, dtype=torch.float32)
            for n, p, off in zip(u.param_names, u.params, u.offsets):
                flat[off : off + _numel(p._b200_full_shape)].copy_(sd[n].reshape(-1).float())
            lo = self.rank * u.shard_numel
            u.master.copy_(flat[lo : lo + u.shard_numel].to(self.device))
            u.shard.copy_(u.master)
        for n, b in self.module.named_buffers():
            if n in sd:
                b.copy_(sd[n])
        self._sync_after_external_update()

    def load_state_dict(self, sd, strict: bool = True):
        self.load_full_state_dict(sd)

    def _sync_after_external_update(self):
        if self.use_symm:
            torch.cuda.synchronize(self.device)
            self.ws.barrier(self.ch_step)
            torch.cuda.synchronize(self.device)
        else:
            torch.distributed.barrier(group=self.group)

    def _meta(self) -> dict:
        return {"world": self.world, "units": [{"name": u.name, "padded": u.padded, "shard_numel": u.shard_numel,
                                                 "params": [{"name": n, "shape": list(p._b200_full_shape), "offset": off}
                                                            for n, p, off in zip(u.param_names, u.params, u.offsets)]} for u in self._units]}

    def save_sharded_state(self, output_dir: str, index: int, plugin=None):
        """``pytorch_model_fsdp_{i}/`` (one file per rank + metadata) or ``pytorch_model_fsdp.bin`` (full)."""
        sdt = getattr(plugin or self.plugin, "state_dict_type", "SHARDED_STATE_DICT")
        if sdt == "FULL_STATE_DICT":
            sd = self.full_state_dict()
            if torch.distributed.get_rank() == 0:
                name = f"{FSDP_MODEL_NAME}.bin" if index == 0 else f"{FSDP_MODEL_NAME}_{index}.bin"
                torch.save(sd, os.path.join(output_dir, name))
            torch.distributed.barrier()
            return
        d = os.path.join(output_dir, f"{FSDP_MODEL_NAME}_{index}")
        os.makedirs(d, exist_ok=True)
        torch.save({u.name: u.master.detach().cpu() for u in self._units}, os.path.join(d, f"__{self.rank}_0.distcp"))
        if self.rank == 0:
            with open(os.path.join(d, ".metadata"), "w") as f:
                json.dump(self._meta(), f)
        torch.distributed.barrier()

    @torch.no_grad()
    def load_sharded_state(self, input_dir: str, index: int, plugin=None):
        sdt = getattr(plugin or self.plugin, "state_dict_type", "SHARDED_STATE_DICT")
        full_name = os.path.join(input_dir, f"{FSDP_MODEL_NAME}.bin" if index == 0 else f"{FSDP_MODEL_NAME}_{index}.bin")
        if sdt == "FULL_STATE_DICT" or (os.path.exists(full_name) and not os.path.isdir(os.path.join(input_dir, f"{FSDP_MODEL_NAME}_{index}"))):
            self.load_full_state_dict(torch.load(full_name, map_location="cpu", weights_only=True))
            return
        d = os.path.join(input_dir, f"{FSDP_MODEL_NAME}_{index}")
        with open(os.path.join(d, ".metadata")) as f:
            meta = json.load(f)
        if meta["world"] != self.world:
            sd = merge_sharded_checkpoint(d)
            self.load_full_state_dict(sd)
            return
        mine = torch.load(os.path.join(d, f"__{self.rank}_0.distcp"), map_location="cpu", weights_only=True)
        for u in self._units:
            u.master.copy_(mine[u.name].to(self.device))
            u.shard.copy_(u.master)
        self._sync_after_external_update()

    # optimizer state ------------------------------------------------------------------------------
    def optimizer_state_dict(self, optimizer) -> dict:
        """Full (gathered) AdamW state in torch's layout: usable as ``optimizer.bin``."""
        state, idx = {}, 0
        index_of = {}
        for g in optimizer.param_groups:
            for p in g["params"]:
                index_of[id(p)] = idx
                idx += 1
        for u in self._units:
            full_m = self._gather_fp32(u.exp_avg)
            full_v = self._gather_fp32(u.exp_avg_sq)
            for p, off in zip(u.params, u.offsets):
                n = _numel(p._b200_full_shape)
                if id(p) in index_of:
                    state[index_of[id(p)]] = {"step": torch.tensor(float(self.step_count)), "exp_avg": full_m[off : off + n].view(p._b200_full_shape).cpu(),
                                              "exp_avg_sq": full_v[off : off + n].view(p._b200_full_shape).cpu()}
        groups = [{k: (list(range(len(v))) if k == "params" else v) for k, v in g.items()} for g in optimizer.param_groups]
        # re-number params across groups
        c = 0
        for g in groups:
            g["params"] = list(range(c, c + len(g["params"])))
            c += len(g["params"])
        return {"state": state, "param_groups": groups}

    def load_optimizer_state_dict(self, optimizer, sd: dict):
        idx = 0
        index_of = {}
        for g in optimizer.param_groups:
            for p in g["params"]:
                index_of[id(p)] = idx
                idx += 1
        for u in self._units:
            fm = torch.zeros(u.padded, dtype=torch.float32)
            fv = torch.zeros(u.padded, dtype=torch.float32)
            for p, off in zip(u.params, u.offsets):
                st = sd["state"].get(index_of.get(id(p), -1))
                if st is None:
                    continue
                n = _numel(p._b200_full_shape)
                fm[off : off + n] = st["exp_avg"].reshape(-1).float()
                fv[off : off + n] = st["exp_avg_sq"].reshape(-1).float()
                self.step_count = int(float(st["step"]))
            lo = self.rank * u.shard_numel
            u.exp_avg.copy_(fm[lo : lo + u.shard_numel].to(self.device))
            u.exp_avg_sq.copy_(fv[lo : lo + u.shard_numel].to(self.device))
        for g_new, g_old in zip(optimizer.param_groups, sd["param_groups"]):
            for k, v in g_old.items():
                if k != "params":
                    g_new[k] = v

    def _gather_fp32(self, shard: torch.Tensor) -> torch.Tensor:
        parts = [torch.empty_like(shard) for _ in range(self.world)]
        torch.distributed.all_gather(parts, shard, group=self.group)
        return torch.cat(parts)

    def save_sharded_optimizer(self, optimizer, output_dir: str, index: int, plugin=None):
        sdt = getattr(plugin or self.plugin, "state_dict_type", "SHARDED_STATE_DICT")
        if sdt == "FULL_STATE_DICT":
            sd = self.optimizer_state_dict(optimizer.optimizer if hasattr(optimizer, "optimizer") else optimizer)
            if torch.distributed.get_rank() == 0:
                torch.save(sd, os.path.join(output_dir, f"{OPTIMIZER_NAME}.bin" if index == 0 else f"{OPTIMIZER_NAME}_{index}.bin"))
            torch.distributed.barrier()
            return
        d = os.path.join(output_dir, f"{OPTIMIZER_NAME}_{index}")
        os.makedirs(d, exist_ok=True)
        inner = optimizer.optimizer if hasattr(optimizer, "optimizer") else optimizer
        torch.save({"step": self.step_count, "hyper": [{k: v for k, v in g.items() if k != "params"} for g in inner.param_groups],
                    "exp_avg": {u.name: u.exp_avg.cpu() for u in self._units}, "exp_avg_sq": {u.name: u.exp_avg_sq.cpu() for u in self._units}},
                   os.path.join(d, f"__{self.rank}_0.distcp"))
        if self.rank == 0:
            with open(os.path.join(d, ".metadata"), "w") as f:
                json.dump(self._meta(), f)
        torch.distributed.barrier()

    def load_sharded_optimizer(self, optimizer, input_dir: str, index: int, plugin=None):
        inner = optimizer.optimizer if hasattr(optimizer, "optimizer") else optimizer
        d = os.path.join(input_dir, f"{OPTIMIZER_NAME}_{index}")
        if not os.path.isdir(d):
            f = os.path.join(input_dir, f"{OPTIMIZER_NAME}.bin" if index == 0 else f"{OPTIMIZER_NAME}_{index}.bin")
            self.load_optimizer_state_dict(inner, torch.load(f, map_location="cpu", weights_only=False))
            return
        mine = torch.load(os.path.join(d, f"__{self.rank}_0.distcp"), map_location="cpu", weights_only=False)
        self.step_count = int(mine["step"])
        for u in self._units:
            u.exp_avg.copy_(mine["exp_avg"][u.name].to(self.device))
            u.exp_avg_sq.copy_(mine["exp_avg_sq"][u.name].to(self.device))
        for g_new, g_old in zip(inner.param_groups, mine["hyper"]):
            g_new.update(g_old)

    def release(self):
        for h in self._handles:
            h.remove()
        self._handles = []
        if self.ws is not None:
            self.ws.close()

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)


def merge_sharded_checkpoint(checkpoint_dir: str) -> dict:
    """Rebuild the full name -> tensor dict from a ``pytorch_model_fsdp_{i}/`` directory (no process group needed)."""
    with open(os.path.join(checkpoint_dir, ".metadata")) as f:
        meta = json.load(f)
    shards = [torch.load(os.path.join(checkpoint_dir, f"__{r}_0.distcp"), map_location="cpu", weights_only=True) for r in range(meta["world"])]
    out = {}
    for u in meta["units"]:
        flat = torch.cat([s[u["name"]] for s in shards])
        for prm in u["params"]:
            n = 1
            for s in prm["shape"]:
                n *= s
            out[prm["name"]] = flat[prm["offset"] : prm["offset"] + n].view(prm["shape"]).clone()
    return out


def _numel(shape) -> int:
    n = 1
    for s in shape:
        n *= int(s)
    return n


def _flatten(x):
    if isinstance(x, torch.Tensor):
        yield x
    elif isinstance(x, (list, tuple)):
        for y in x:
            yield from _flatten(y)
    elif isinstance(x, dict):
        for y in x.values():
            yield from _flatten(y)
    elif hasattr(x, "__dict__") and not isinstance(x, nn.Module):
        for y in vars(x).values():
            yield from _flatten(y)
