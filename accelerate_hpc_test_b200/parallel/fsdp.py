"""``B200ShardedModule``: fully sharded data parallelism (FSDP2 semantics) on peer-mapped memory.

The reference hands the model to torch's ``fully_shard`` (reference src/accelerate/utils/fsdp_utils.py:
621-737) and lets FSDP2 issue ``ncclAllGather`` / ``ncclReduceScatter`` per wrapped module.  Here:

* **Per-parameter dim-0 shards** (the FSDP2 layout).  Every wrapped unit (the modules picked by the plugin's
  auto-wrap policy, plus one unit per large leftover tensor) keeps, for each of its parameters, rows
  ``[r * N/W, (r+1) * N/W)`` on rank ``r`` (``W`` = shard-group size; parameters whose first dimension does not
  divide are cut as flat 1-D tensors).  The pieces of one unit are concatenated into one *shard* buffer in
  symmetric (peer-mapped) memory, next to local fp32 master weights and optimizer moments.  In the gathered
  ("full") buffer every parameter sits in its natural ``[N, K]`` layout, rank ``q``'s piece at row ``q * N/W``.
* **All-gather fused into GEMMs.**  A weight that the model consumes as the B operand of the tcgen05 GEMM is not
  gathered by a separate collective: the GEMM that first needs it pulls the peers' row blocks over NVLink with a
  gather warp while its tensor cores start on this rank's own rows (forward, n-gated), and in backward the
  weight-gradient GEMMs carry the gather of the next weights the stream will need (dgrad GEMMs would have to wait
  for the whole reduction dimension, so their weights travel one GEMM ahead).  Everything else (norm weights,
  biases, parameters other ops consume) is gathered by a TMA pull kernel on a side stream one unit ahead.
* **Reduce-scatter + optimizer in one kernel**: when a unit's backward is complete each rank sums its pieces of
  all ranks' flat gradients through NVLink loads, applies AdamW to its fp32 shard and writes the new bf16 shard.
  With hybrid sharding (``dp_replicate`` x ``dp_shard``) the same kernel reads the pieces of *all* data-parallel
  ranks, so the replicate all-reduce needs no second collective and the fused update stays on.  If gradients must
  survive until ``optimizer.step()`` (gradient clipping, any optimizer other than single-group AdamW) the kernel
  only reduces and the update runs at ``step()`` — for foreign optimizers through shard-shaped parameter views
  handed to the user's own optimizer (the reference's ``data_ptr`` swap, accelerator.py:1690-1744).
* Native weight-gradient GEMMs write straight into the unit's flat gradient buffer.

On CPU (gloo) the same engine runs with c10d collectives — that path is the test oracle.
Checkpoints: ``pytorch_model_fsdp_{i}/`` and ``optimizer_{i}/`` are ``torch.distributed.checkpoint`` directories
with one ``Shard(0)`` DTensor per parameter (interchangeable with the reference's, utils/fsdp_utils.py:147-158,
270-278); ``pytorch_model_fsdp.bin`` / ``optimizer.bin`` for FULL_STATE_DICT.
"""

from __future__ import annotations

import contextlib
import ctypes
import os
import re
from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as F
from ..utils.constants import FSDP_MODEL_NAME, OPTIMIZER_NAME
from ..utils.diagnostics import shadow_check
from ..utils.other import get_module_children_bottom_up

BF16 = torch.bfloat16
SEGS_PER_LAUNCH = 32  # csrc/comm.cu MAX_SEGS


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


class _Seg:
    """One parameter of a unit: where its per-rank piece sits in the shard and where it sits in the gathered buffer."""

    __slots__ = ("param", "name", "shape", "numel", "rows", "piece", "full_off", "shard_off", "job")

    def __init__(self, param, name):
        self.param = param
        self.name = name
        self.shape = tuple(param.shape)
        self.numel = param.numel()
        self.rows = 0        # rows per rank when the parameter is cut along dim 0 (0: cut as a flat tensor)
        self.piece = 0       # elements per rank
        self.full_off = 0    # element offset of the parameter in the gathered buffer
        self.shard_off = 0   # element offset of my piece in the shard
        self.job = None      # _GatherJob when this weight's gather rides on GEMMs

    def layout(self, world: int):
        d0 = self.shape[0] if len(self.shape) >= 2 else 0
        row_len = self.numel // d0 if d0 else 0
        if d0 and d0 % world == 0 and ((d0 // world) * row_len) % 8 == 0:
            self.rows = d0 // world
            self.piece = self.rows * row_len
        else:
            self.rows = 0
            self.piece = ((self.numel + world - 1) // world + 7) // 8 * 8
        return self.piece

    @property
    def piece_shape(self):
        """Shape of my piece when it is a dim-0 block of the parameter (rows > 0, or an evenly cut 1-D tensor)."""
        if self.rows:
            return (self.rows,) + self.shape[1:]
        return None


class _GatherJob:
    """Gather of one dim-0-sharded weight performed by the gather warp of a GEMM (ops.functional.gemm)."""

    __slots__ = ("engine", "unit", "seg", "pending", "queued", "calls", "ready", "src_off", "piece_bytes", "rows")

    def __init__(self, engine, unit, seg):
        self.engine, self.unit, self.seg = engine, unit, seg
        self.pending = False   # the gathered buffer does not hold the peers' pieces yet
        self.queued = False    # waiting in the engine's carry queue
        self.calls = 0         # completed gathers: the byte counters are monotonic
        self.ready = torch.zeros(8, dtype=torch.int64, device=engine.device)
        self.src_off = unit.shard_off + seg.shard_off * 2
        self.piece_bytes = seg.piece * 2
        self.rows = seg.rows

    # called by ops.functional.gemm ---------------------------------------------------------------
    def launch_args(self):
        """(peer array, rank, world, src_off, dst pointer, piece_bytes, rows, ready pointer, expect)."""
        e = self.engine
        self.calls += 1
        self.pending = False
        if self.queued:
            self.queued = False
            try:
                e._carry.remove(self)
            except ValueError:
                pass
        dst = self.unit.full.data_ptr() + self.seg.full_off * 2  # (the parameter's .data is only re-pointed at unshard time)
        return (e._ag_peers, e.rank, e.world, self.src_off, ctypes.c_void_p(dst), self.piece_bytes, self.rows,
                ctypes.c_void_p(self.ready.data_ptr()), self.calls * self.piece_bytes)

    def materialize(self):
        """Stand-alone gather on the current stream, for consumers that are not a GEMM."""
        if self.pending:
            self.engine._pull_segments(self.unit, [self.seg], stream=None)
            self.pending = False
            if self.queued:
                self.queued = False
                try:
                    self.engine._carry.remove(self)
                except ValueError:
                    pass


class _Unit:
    """One sharded group of parameters (a wrapped module, or a leftover group of the root)."""

    def __init__(self, name: str, module: Optional[nn.Module], params: list, names: list):
        self.name = name
        self.module = module
        self.params = params
        self.param_names = names
        self.segs = [_Seg(p, n) for p, n in zip(params, names)]
        self.offsets = []          # element offset of every parameter in the gathered buffer
        self.numel = sum(p.numel() for p in params)
        self.padded = 0
        self.shard_numel = 0
        self.full: Optional[torch.Tensor] = None  # pooled full buffer while unsharded
        self.grad: Optional[torch.Tensor] = None  # pooled flat gradient buffer during backward
        self.grad_off = 0
        self.ag_event = None
        self.unsharded = False
        self.in_backward = False
        self.ready = {}
        self.expected = {}
        self.has_partial = False
        self.rs_launched = False
        self.rs_event = None
        self.updated_in_backward = False

    def layout(self, world: int):
        full, shard = 0, 0
        self.offsets = []
        for s in self.segs:
            piece = s.layout(world)
            s.full_off, s.shard_off = full, shard
            self.offsets.append(full)
            full += piece * world
            shard += piece
        self.padded, self.shard_numel = full, shard

    def piece_view(self, flat: torch.Tensor, seg: _Seg) -> torch.Tensor:
        return flat[seg.shard_off : seg.shard_off + seg.piece]


def _seg_table(rows):
    arr = (ctypes.c_longlong * (3 * len(rows)))()
    for i, r in enumerate(rows):
        arr[3 * i], arr[3 * i + 1], arr[3 * i + 2] = r
    return arr


class B200ShardedModule(nn.Module):
    """The sharded engine: wraps ``module``, keeps this rank's dim-0 shard of every parameter (fp32 master + compute copy) in peer-mapped memory, gathers a unit's parameters just before it runs, and reduce-scatters — with ``FusedAdamW``, also updates — its gradients as soon as they are complete."""

    def __init__(self, module: nn.Module, plugin, device, shard_group=None, replicate_group=None, gradient_state=None,
                 shard_group_is_world: bool = True, dp_group=None):
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
        self.dp_group = dp_group  # shard x replicate ranks (hybrid sharding); None: the shard group is all there is
        self.world = torch.distributed.get_world_size(self.group)
        self.rank = torch.distributed.get_rank(self.group)
        self.use_symm = self.device.type == "cuda"
        self.reshard_after_forward = plugin.effective_reshard_after_forward
        self.requires_gradient_sync = True
        first = next(module.parameters())
        pol = plugin.mixed_precision_policy
        if self.use_symm:
            wants_bf16 = first.dtype == BF16 or (pol is not None and getattr(pol, "param_dtype", None) == BF16)
            if not wants_bf16:
                raise ValueError(
                    "the native sharded engine computes in bf16 (bf16 parameters, or mixed_precision='bf16' on fp32 parameters); "
                    f"got {first.dtype} parameters with mixed-precision policy {pol}. Accelerator.prepare falls back to torch FSDP2 "
                    "for that case; constructing B200ShardedModule directly needs bf16."
                )
            self.param_dtype = BF16
        else:
            self.param_dtype = first.dtype
        self.step_count = 0
        self._optimizer = None
        self._fused = False
        self._hyper_ok = False
        self._fused_allowed = False
        self._generic = False          # foreign optimizer: stepped on shard-shaped parameter views
        self._shard_params = {}        # id(original param) -> shard-shaped nn.Parameter (generic optimizers)
        self._clip_requested_this_step = False
        self._saw_clip = False
        self._synced_backwards = 0     # synchronised backward passes since the last optimizer step
        self._clean_steps = 0          # consecutive optimizer steps that followed exactly one synchronised backward
        self._clip_coef = None
        self._callback_queued = False
        self._units: list[_Unit] = []
        self._order: list[_Unit] = []  # forward execution order, learnt on the first iteration
        self._order_locked = False
        self._carry: list[_GatherJob] = []  # gathers waiting for a GEMM to ride on
        self._fuse_gather = os.environ.get("ACCELERATE_B200_FSDP_FUSED_GATHER", "1") != "0"
        self._fwd_gather = os.environ.get("ACCELERATE_B200_FSDP_FWD_GATHER", "jit")  # jit | carry
        self._early_root_rs = os.environ.get("ACCELERATE_B200_FSDP_EARLY_ROOT_RS", "1") != "0"
        self._build_units()
        self._allocate()
        self._install_hooks()
        if self.use_symm:
            F.register_carry_source(self)

    # ------------------------------------------------------------------ construction
    def _build_units(self):
        plugin = self.plugin
        claimed = set()
        ignored = set()
        if isinstance(plugin.ignored_modules, str):
            # a string is a regular expression on module names (FSDP_IGNORED_MODULES); what it names stays whole on every rank
            pattern = re.compile(plugin.ignored_modules)
            plugin.ignored_modules = [m for name, m in self.module.named_modules() if pattern.fullmatch(name)]
        if plugin.ignored_modules is not None:
            for m in plugin.ignored_modules:
                ignored.update(id(p) for p in m.parameters())
        names = {id(p): n for n, p in self.module.named_parameters()}
        # a parameter registered in several modules (tied embeddings) receives one gradient per use
        uses = {}
        for _, p in self.module.named_parameters(remove_duplicate=False):
            uses[id(p)] = uses.get(id(p), 0) + 1
        self._uses = uses
        if plugin.activation_checkpointing:
            apply_activation_checkpointing(self.module, plugin)
        # a module marked `_z3_leaf` (DeepSpeedPlugin.set_moe_leaf_modules: mixture-of-experts blocks, whose experts run in a
        # data-dependent order) is one unit whatever the policy says, and nothing inside it becomes a unit of its own
        inside_leaf = {id(m) for marked in self.module.modules() if getattr(marked, "_z3_leaf", False)
                       for m in marked.modules() if m is not marked}
        for fqn, sub in get_module_children_bottom_up(self.module, return_fqns=True):
            if sub is self.module or id(sub) in inside_leaf:
                continue
            if not (getattr(sub, "_z3_leaf", False) or plugin.should_wrap(sub)):
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
        self._seg_of_param = {id(s.param): s for u in self._units for s in u.segs}
        self._ignored_params = [p for p in self.module.parameters() if id(p) in ignored]
        self._frozen = [p for p in self.module.parameters() if not p.requires_grad and id(p) not in ignored]

    def _allocate(self):
        world = self.world
        for u in self._units:
            u.layout(world)
        layer_units = [u for u in self._units if u.module is not None]
        root_units = [u for u in self._units if u.module is None]
        max_layer = max([u.padded for u in layer_units], default=0)
        total_shard = sum(u.shard_numel for u in self._units)
        esz = 2 if self.param_dtype == BF16 else 4
        n_grad_bufs = int(os.environ.get("ACCELERATE_B200_FSDP_GRAD_BUFFERS", "3"))
        n_full_bufs = int(os.environ.get("ACCELERATE_B200_FSDP_PARAM_BUFFERS", "3"))
        self._ws_group = self.dp_group if (self.dp_group is not None and self.replicate_group is not None) else self.group
        if self.use_symm:
            from .symm import SymmetricWorkspace

            # shards + dedicated root gradient buffers + pooled layer gradient buffers (+ slack for pool growth)
            need = total_shard * esz + sum(u.padded for u in root_units) * esz + (n_grad_bufs + 2) * max_layer * esz + 64 * 2**20
            self.ws = SymmetricWorkspace(self._ws_group, need, self.device)
            # gathers and reductions get their own streams: a reduce-scatter waits for its peers, and a
            # prefetching all-gather must not queue behind that wait
            self.comm_stream = torch.cuda.Stream(device=self.device)
            self.rs_stream = torch.cuda.Stream(device=self.device)
            self.ag_ctas = int(os.environ.get("ACCELERATE_B200_AG_CTAS", "148"))
            # the reduce-scatter (+AdamW) runs beside the backward GEMMs and each rank owns 1/world of a layer: it gets just enough
            # CTAs to finish under the next layer's backward and leaves the other SM slots to the GEMMs.  Measured ms/step:
            # 2 GPUs 218.1 / 218.8 / 223.0 / 226.9 with 148 / 96 / 64 / 48 CTAs; 4 GPUs 216.6 / 206.9 / 216.8 with 148 / 64 / 32;
            # 8 GPUs (one box, one model instance) 213.1 / 211.7 / 205.9 / 206.4 / 208.2 with 148 / 96 / 64 / 48 / 32
            self.rs_ctas = int(os.environ.get("ACCELERATE_B200_RS_CTAS", "0")) or (128 if self.ws.world <= 2 else 64)
            self.ch_step = self.ws.new_channel()
            # all-gathers stay inside the shard group; reductions span shard x replicate (hybrid sharding)
            ws_ranks = torch.distributed.get_process_group_ranks(self._ws_group)
            shard_ranks = torch.distributed.get_process_group_ranks(self.group)
            self._ag_peers = (ctypes.c_void_p * self.world)(*[self.ws.peer_bases[ws_ranks.index(g)] for g in shard_ranks])
            self.rs_world = self.ws.world
            self.rs_rank = self.ws.rank
            self._nvls_rs = self.ws.has_multicast and os.environ.get("ACCELERATE_B200_FSDP_NVLS_RS", "1") != "0"
            # a shard group of one (pure tensor / context parallelism, dp_shard=1): the shard IS the parameter — nothing to gather,
            # nothing to reduce; the engine only keeps the fp32 master / moments and updates the bf16 copy in place
            self._solo = self.world == 1 and self.rs_world == 1 and os.environ.get("ACCELERATE_B200_FSDP_SOLO", "1") != "0"
        else:
            self.ws = None
            self.comm_stream = None
            self.rs_stream = None
            self._ag_peers = None
            self._nvls_rs = False
            self._solo = False
            self.rs_world = self.world * (torch.distributed.get_world_size(self.replicate_group) if self.replicate_group is not None else 1)
        meta_init = any(p.device.type == "meta" for u in self._units for p in u.params)
        src = torch.distributed.get_global_rank(self.group, 0)
        # shards: parameter shard (symmetric), fp32 master + moments (local)
        for u in self._units:
            if self.use_symm:
                u.shard_off, u.shard = self.ws.alloc_tensor((u.shard_numel,), self.param_dtype)
                u.ch_arrive, u.ch_done = self.ws.new_channel(), self.ws.new_channel()
            else:
                u.shard_off, u.shard = 0, torch.zeros(u.shard_numel, dtype=self.param_dtype, device=self.device)
            u.master = torch.zeros(u.shard_numel, dtype=torch.float32, device=self.device)
            for s in u.segs:
                p = s.param
                if meta_init or p.device.type == "meta":
                    # meta-device construction (cpu_ram_efficient_loading): values arrive later through load_full_state_dict
                    pass
                else:
                    # replicas must start identical: the values of the group's first rank win, one parameter at a time so
                    # that no rank ever holds a second full copy of the model (reference: fsdp2_load_full_state_dict)
                    full = p.detach().to(self.device, torch.float32).reshape(-1)
                    if self.world > 1:
                        torch.distributed.broadcast(full, src=src, group=self.group)
                    lo = self.rank * s.piece
                    n = max(0, min(s.piece, s.numel - lo))
                    if n > 0:
                        u.master[s.shard_off : s.shard_off + n].copy_(full[lo : lo + n])
                    del full
                p._b200_full_shape = s.shape
                p.data = torch.empty(0, dtype=self.param_dtype, device=self.device)
            u.shard.copy_(u.master)
            u.exp_avg = torch.zeros_like(u.master)
            u.exp_avg_sq = torch.zeros_like(u.master)
            u.grad_shard = None
            for p in u.params:
                u.expected[id(p)] = self._uses.get(id(p), 1)
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
        for p in self._frozen:
            if p.device != self.device and p.device.type != "meta":
                p.data = p.data.to(self.device)
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
    def _rides_on_gemm(self, s: _Seg) -> bool:
        """Weights the model feeds to the tcgen05 GEMM as its B operand are gathered by that GEMM (or one before it)."""
        return (self._fuse_gather and self.use_symm and self.world > 1 and s.rows > 0 and len(s.shape) == 2 and s.piece >= 2**16
                and getattr(s.param, "_b200_gemm_b_seen", False) and not getattr(s.param, "_b200_no_fused_gather", False))

    def _pull_segments(self, u: _Unit, segs, stream):
        """TMA pull kernel for `segs` of `u` into u.full (stream None: the current stream)."""
        lib = self.ws.lib
        from ..ops._lib import check, cur_stream, ptr

        st = cur_stream() if stream is None else ctypes.c_void_p(stream.cuda_stream)
        for i in range(0, len(segs), SEGS_PER_LAUNCH):
            part = segs[i : i + SEGS_PER_LAUNCH]
            tab = _seg_table([(s.shard_off * 2, s.full_off * 2, s.piece * 2) for s in part])
            check(lib.b200_allgather_pull_segs(self._ag_peers, self.rank, self.world, u.shard_off, tab, len(part), ptr(u.full), -1, 0,
                                               self.ag_ctas, st), "allgather_pull_segs")

    def _launch_gather(self, u: _Unit, carry: bool = False):
        """Give ``u`` a full buffer and start filling it; idempotent.  Segments that ride on GEMMs get a pending job (queued for
        the next GEMMs when ``carry``), the rest is pulled on the comm stream."""
        if u.full is not None:
            return
        if self._solo:
            u.full, u.full_idx, u.ag_event = u.shard, None, None  # same layout when world == 1: the parameters view the shard
            return
        idx = u.fpool.take()
        u.full_idx = idx
        u.full = u.fpool.bufs[idx][1][: u.padded]
        if self.use_symm:
            ev = u.fpool.events[idx]
            riding = [s for s in u.segs if self._rides_on_gemm(s)]
            pulled = [s for s in u.segs if not self._rides_on_gemm(s)]
            cur = torch.cuda.current_stream(self.device)
            if riding:
                if ev is not None:
                    cur.wait_event(ev)
                for s in riding:
                    if s.job is None:
                        s.job = _GatherJob(self, u, s)
                    # the GEMM's gather warp brings every piece, this rank's own included (a local copy out of the shard)
                    s.job.pending = True
                    if carry and not s.job.queued:
                        s.job.queued = True
                        self._carry.append(s.job)
            if pulled:
                if ev is not None:
                    self.comm_stream.wait_event(ev)
                self._pull_segments(u, pulled, self.comm_stream)
                u.ag_event = torch.cuda.Event()
                u.ag_event.record(self.comm_stream)
            else:
                u.ag_event = None
        else:
            parts = [torch.empty_like(u.shard) for _ in range(self.world)]
            torch.distributed.all_gather(parts, u.shard, group=self.group)
            for s in u.segs:
                for q in range(self.world):
                    u.full[s.full_off + q * s.piece : s.full_off + (q + 1) * s.piece].copy_(parts[q][s.shard_off : s.shard_off + s.piece])

    def pop_carry_job(self) -> Optional[_GatherJob]:
        """Next weight gather waiting for a GEMM to ride on (ops.functional.gemm asks before every plain launch)."""
        while self._carry:
            job = self._carry.pop(0)
            job.queued = False
            if job.pending:
                return job
        return None

    def unshard(self, u: Optional[_Unit] = None, complete: bool = False):
        """Make the parameters of ``u`` (or of every unit) whole: ``param.data`` becomes a view of the gathered buffer.
        ``complete`` also runs, right away, the gathers that would otherwise ride on later GEMMs."""
        if u is None:
            for x in self._units:
                self.unshard(x, complete)
            return
        if not u.unsharded:
            self._launch_gather(u)
            if self.use_symm and u.ag_event is not None:
                self.trace.wait_event(torch.cuda.current_stream(self.device), u.ag_event, "all_gather")
            for s in u.segs:
                s.param.data = u.full[s.full_off : s.full_off + s.numel].view(s.shape)
                s.param._b200_ag = s.job if (s.job is not None and s.job.pending) else None
            u.unsharded = True
        if complete:
            for s in u.segs:
                if s.job is not None and s.job.pending:
                    s.job.materialize()
                s.param._b200_ag = None
            if self.use_symm and self._shadow:
                # debug: the same gather through torch.distributed must give the same bytes
                ref = torch.empty(self.world * u.shard_numel, dtype=u.shard.dtype, device=self.device)
                torch.distributed.all_gather_into_tensor(ref, u.shard, group=self.group)
                for s in u.segs:
                    for q in range(self.world):
                        shadow_check(f"all_gather[{u.name}:{s.name}:{q}]", u.full[s.full_off + q * s.piece : s.full_off + (q + 1) * s.piece],
                                     ref[q * u.shard_numel + s.shard_off : q * u.shard_numel + s.shard_off + s.piece])

    def reshard(self, u: Optional[_Unit] = None, write_back: bool = False):
        if u is None:
            for x in self._units:
                self.reshard(x, write_back)
            return
        if not u.unsharded:
            return
        if write_back:
            for s in u.segs:
                if s.job is not None and s.job.pending:
                    s.job.materialize()
                lo = s.full_off + self.rank * s.piece
                u.master[s.shard_off : s.shard_off + s.piece].copy_(u.full[lo : lo + s.piece])
            u.shard.copy_(u.master)
        for s in u.segs:
            s.param.data = torch.empty(0, dtype=self.param_dtype, device=self.device)
            s.param._b200_ag = None
            if s.job is not None:
                s.job.pending = False
        ev = None
        if u.full_idx is not None:
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
            self._launch_gather(nxt, carry=self._fwd_gather == "carry")
        if torch.is_grad_enabled() and not u.in_backward:
            u._want_pre_backward = True
        return None

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
            # the weight-gradient GEMMs of this unit carry the gathers of the unit that runs next
            self._launch_gather(prv, carry=True)
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
            # elements that pad a parameter up to a whole number of pieces belong to no gradient: keep them zero so that the
            # reduced shard (and the gradient norm taken over it) never sees what an earlier user of the buffer left there
            for s in u.segs:
                tail = s.piece * self.world - s.numel
                if tail > 0:
                    u.grad[s.full_off + s.numel : s.full_off + s.numel + tail].zero_()
        u.ready = {}
        u.rs_launched = False
        for s in u.segs:
            s.param._b200_grad_buf = u.grad[s.full_off : s.full_off + s.numel].view(s.shape)
            s.param._b200_has_partial = u.has_partial
            s.param._b200_grad_seen = False

    def _native_grad_ready(self, p):
        u = self._unit_of_param[id(p)]
        self._mark_ready(u, p)

    def _autograd_grad_ready(self, p):
        """Gradient produced by a foreign op landed in ``p.grad``: move it into the flat buffer."""
        u = self._unit_of_param[id(p)]
        g = p.grad
        if g is None:
            return  # the native op already wrote this gradient into the flat buffer and reported it (deliver_weight_grad)
        if u.grad is None:
            # root-level parameters have no pre-backward hook of their own
            self._queue_final_callback()
            u.in_backward = True
            self._attach_grad_buffer(u)
        if g is not None:
            view = p._b200_grad_buf
            if u.has_partial or id(p) in u.ready:
                view.add_(g.to(view.dtype))
            else:
                view.copy_(g)
            p.grad = None
            # autograd has already summed every use of a shared parameter into this one gradient
            u.ready[id(p)] = u.expected.get(id(p), 1) - 1
        self._mark_ready(u, p)

    def _mark_ready(self, u: _Unit, p):
        u.ready[id(p)] = u.ready.get(id(p), 0) + 1
        if all(u.ready.get(id(q), 0) >= u.expected.get(id(q), 1) for q in u.params):
            # root units too: the LM head's reduce-scatter starts at the top of backward instead of after it
            if u.module is not None or self._early_root_rs:
                self._post_backward(u)

    def _post_backward(self, u: _Unit):
        """All gradients of ``u`` are in its flat buffer: reduce (+ update) and let go of the parameters."""
        if u.rs_launched:
            return
        u.rs_launched = True
        if u.grad is not None:
            missing = [s for s in u.segs if id(s.param) not in u.ready]
            if missing and not u.has_partial:
                for s in missing:
                    u.grad[s.full_off : s.full_off + s.numel].zero_()
            if self.requires_gradient_sync:
                self._reduce_model_parallel_replicas(u)
                self._reduce_unit(u)
                self._release_grad(u)
            else:
                u.has_partial = True
        if u.unsharded:
            self.reshard(u)
        u.in_backward = False

    def _reduce_model_parallel_replicas(self, u: _Unit):
        """Parameters that a tensor-parallel group keeps whole (norms, LM head) only saw that rank's token shard: their
        gradients are summed over the tp group before the data-parallel reduce-scatter (tensor_parallel._mark_replicated
        leaves this to the engine when the parameter belongs to a sharded unit).  One all-reduce per unit and group: the
        replicated parameters of a unit are copied into one staging tensor."""
        by_group = {}
        for s in u.segs:
            g = getattr(s.param, "_b200_tp_group", None)
            if g is None or torch.distributed.get_world_size(g) == 1:
                continue
            by_group.setdefault(g, []).append(s)
        for g, segs in by_group.items():
            if len(segs) == 1:
                s = segs[0]
                torch.distributed.all_reduce(u.grad[s.full_off : s.full_off + s.numel], group=g)
                continue
            stage = torch.cat([u.grad[s.full_off : s.full_off + s.numel] for s in segs])
            torch.distributed.all_reduce(stage, group=g)
            o = 0
            for s in segs:
                u.grad[s.full_off : s.full_off + s.numel].copy_(stage[o : o + s.numel])
                o += s.numel

    def _release_grad(self, u: _Unit):
        for p in u.params:
            p._b200_grad_buf = None
            p._b200_has_partial = False
        u.gpool.give(u.grad_idx, u.rs_event if self.use_symm else None)
        u.grad = None
        u.has_partial = False

    def _finish_backward(self):
        self._callback_queued = False
        if self.requires_gradient_sync:
            self._synced_backwards += 1
            if self._fused and self._hyper_ok and self._synced_backwards > 1:
                raise RuntimeError(
                    "a second synchronised backward pass ran before optimizer.step(), but the fused reduce-scatter + AdamW already "
                    "applied the first one's gradients. Accumulate under `accelerator.accumulate(model)` / `no_sync`, or set "
                    "ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD=0."
                )
        if self._order and not self._order_locked:
            self._order_locked = True
        for u in self._units:
            if u.grad is not None and not u.rs_launched:
                self._post_backward(u)
            elif u.in_backward:
                if u.unsharded:
                    self.reshard(u)
                u.in_backward = False
        # gathers nobody picked up (a backward that ended early) must not leak into the next step
        for job in self._carry:
            job.queued = False
        self._carry.clear()
        if self.use_symm:
            cur = torch.cuda.current_stream(self.device)
            for u in self._units:
                if u.rs_event is not None:
                    self.trace.wait_event(cur, u.rs_event, "reduce_scatter_tail")

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
        scale = 1.0 / self.rs_world
        if self.use_symm:
            from ..ops._lib import check, ptr

            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.rs_stream.wait_event(ev)
            if not fused and u.grad_shard is None:
                u.grad_shard = torch.empty(u.shard_numel, dtype=torch.float32, device=self.device)
            h = self._adam_hyper() if fused else dict(lr=0.0, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0)
            if self._solo:
                # nobody to reduce with: the flat gradient buffer already is the gradient shard (same layout at world 1)
                with torch.cuda.stream(self.rs_stream):
                    if fused:
                        F.adamw_flat_(u.master, u.exp_avg, u.exp_avg_sq, u.grad, u.shard, step=self.step_count + 1, **h)
                    else:
                        F.scale_cast(u.grad, u.grad_shard, 1.0)
                    u.rs_event = torch.cuda.Event()
                    u.rs_event.record(self.rs_stream)
                u.updated_in_backward = fused
                return
            with torch.cuda.stream(self.rs_stream):
                for i in range(0, len(u.segs), SEGS_PER_LAUNCH):
                    part = u.segs[i : i + SEGS_PER_LAUNCH]
                    base = part[0].shard_off
                    # hybrid sharding: the pieces I own are indexed by my SHARD rank, read from every rank of shard x replicate
                    tab = _seg_table([(s.shard_off - base, s.full_off, s.piece) for s in part])
                    e = self.ws.next_epoch(u.ch_arrive)
                    self.ws._epochs[u.ch_done] = e
                    fo = base * 4
                    if self._nvls_rs:
                        # the switch adds the ranks' gradient pieces: one incoming stream instead of rs_world
                        check(self.ws.lib.b200_reduce_scatter_adamw_nvls(
                            self.ws.peer_array, ctypes.c_void_p(self.ws.mc_base), self.rs_rank, self.rs_world, self.rank, u.grad_off, tab, len(part),
                            ctypes.c_void_p(u.master.data_ptr() + fo), ctypes.c_void_p(u.exp_avg.data_ptr() + fo), ctypes.c_void_p(u.exp_avg_sq.data_ptr() + fo),
                            u.shard_off + base * 2, ctypes.c_void_p(u.grad_shard.data_ptr() + fo) if not fused else None,
                            h["lr"], h["beta1"], h["beta2"], h["eps"], h["weight_decay"], self.step_count + 1, scale, int(fused), None,
                            u.ch_arrive, u.ch_done, e, self.ws.counter_ptr(u.ch_arrive), self.rs_ctas,
                            ctypes.c_void_p(self.rs_stream.cuda_stream)), "reduce_scatter_adamw_nvls")
                        continue
                    check(self.ws.lib.b200_reduce_scatter_adamw_segs(
                        self.ws.peer_array, self.rs_rank, self.rs_world, self.rank, u.grad_off, tab, len(part),
                        ctypes.c_void_p(u.master.data_ptr() + fo), ctypes.c_void_p(u.exp_avg.data_ptr() + fo), ctypes.c_void_p(u.exp_avg_sq.data_ptr() + fo),
                        u.shard_off + base * 2, ctypes.c_void_p(u.grad_shard.data_ptr() + fo) if not fused else None,
                        h["lr"], h["beta1"], h["beta2"], h["eps"], h["weight_decay"], self.step_count + 1, scale, int(fused), None,
                        u.ch_arrive, u.ch_done, e, self.ws.counter_ptr(u.ch_arrive), self.rs_ctas,
                        ctypes.c_void_p(self.rs_stream.cuda_stream)), "reduce_scatter_adamw_segs")
                u.rs_event = torch.cuda.Event()
                u.rs_event.record(self.rs_stream)
            u.updated_in_backward = fused
        else:
            full = u.grad.to(torch.float32)
            torch.distributed.all_reduce(full, group=self.group)
            if self.replicate_group is not None:
                torch.distributed.all_reduce(full, group=self.replicate_group)
            u.grad_shard = torch.cat([full[s.full_off + self.rank * s.piece : s.full_off + (self.rank + 1) * s.piece] for s in u.segs]) * scale
            u.updated_in_backward = False

    # ------------------------------------------------------------------ optimizer integration
    def bind_optimizer(self, accelerated_optimizer):
        """Take over the parameter update of ``accelerated_optimizer`` (sharded state lives in this engine).

        Single-group Adam/AdamW runs on the engine's own kernels (and may fuse with the reduce-scatter).  Any other optimizer
        keeps working the way the reference keeps it working after ``fully_shard``: its parameter references are swapped for
        shard-shaped fp32 views of the master weights, the reduced gradient shard becomes their ``.grad`` and the user's
        ``optimizer.step()`` runs on them.
        """
        inner = accelerated_optimizer.optimizer
        self._optimizer = inner
        accelerated_optimizer.attach_engine(self)
        name = type(inner).__name__
        g = inner.param_groups[0] if inner.param_groups else {}
        self._hyper_ok = (name in ("AdamW", "FusedAdamW", "Adam") and not g.get("amsgrad", False) and not g.get("maximize", False)
                          and len(inner.param_groups) == 1 and not (name == "Adam" and float(g.get("weight_decay", 0.0)) != 0.0))
        self._generic = not self._hyper_ok
        if name == "Adam":
            g.setdefault("weight_decay", 0.0)
        allow = os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto")
        self._fused_allowed = allow != "0" and self.use_symm and self._hyper_ok
        self._fused = allow == "1" and self._fused_allowed
        # the user's optimizer never sees the full-size tensors again: drop their (empty) state
        inner.state.clear()
        if self._generic:
            self._swap_in_shard_params(inner)

    def _swap_in_shard_params(self, inner):
        mapping = {}
        for u in self._units:
            for s in u.segs:
                view = u.master[s.shard_off : s.shard_off + s.piece]
                if s.piece_shape is not None:
                    view = view.view(s.piece_shape)
                sp = nn.Parameter(view, requires_grad=True)
                sp._b200_origin = s.name
                self._shard_params[id(s.param)] = sp
                mapping[id(s.param)] = sp
        for group in inner.param_groups:
            group["params"] = [mapping.get(id(p), p) for p in group["params"]]

    def zero_grad(self):
        self._clip_requested_this_step = False
        for sp in self._shard_params.values():
            sp.grad = None

    def clip_grad_norm_(self, max_norm, norm_type=2):
        """Global gradient norm over all shards; the clip coefficient is applied inside the update kernel."""
        if norm_type != 2:
            raise NotImplementedError("only the L2 norm is implemented for sharded gradients")
        self._clip_requested_this_step = True
        self._saw_clip = True
        if any(u.updated_in_backward for u in self._units):
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
        self.step_count += 1
        ctx = torch.cuda.stream(self.rs_stream) if self.use_symm else contextlib.nullcontext()
        if self.use_symm:
            self.rs_stream.wait_stream(torch.cuda.current_stream(self.device))
        with ctx:
            if self._generic:
                self._generic_step(optimizer)
            else:
                h = self._adam_hyper()
                for u in self._units:
                    if u.updated_in_backward:
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
                torch.cuda.current_stream(self.device).wait_stream(self.rs_stream)
        self._clip_coef = None
        # the first steps run in the safe mode; the fused update is switched on once two optimizer steps in a row each followed
        # exactly one synchronised backward (a loop that accumulates by hand steps after several) and nobody asked for the gradients
        self._clean_steps = self._clean_steps + 1 if self._synced_backwards == 1 else 0
        self._synced_backwards = 0
        if (self._fused_allowed and not self._fused and not self._saw_clip and self._clean_steps >= 2
                and os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto") == "auto"):
            self._fused = True
            for u in self._units:
                u.grad_shard = None
        return False

    def _generic_step(self, optimizer):
        """The user's optimizer on shard-shaped views: grads are the reduced fp32 shards (clipped if asked)."""
        any_grad = False
        for u in self._units:
            if u.grad_shard is None:
                continue
            if self._clip_coef is not None:
                u.grad_shard.mul_(self._clip_coef)
            for s in u.segs:
                sp = self._shard_params[id(s.param)]
                g = u.grad_shard[s.shard_off : s.shard_off + s.piece]
                sp.grad = g.view(sp.shape)
                any_grad = True
        if any_grad:
            optimizer.step()
        for u in self._units:
            if u.shard.dtype == BF16 and self.use_symm:
                F.scale_cast(u.master, u.shard, 1.0)
            else:
                u.shard.copy_(u.master)

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
        self.unshard(complete=True)
        return self.module

    def exposed_comm_ms(self, by_tag: bool = False):
        """Milliseconds (since the last call) the compute stream spent waiting for all-gathers / reduce-scatters that were not
        hidden behind compute.  Needs ``self.trace.enabled = True`` (or ACCELERATE_B200_TRACE_COMM=1); synchronises."""
        return self.trace.exposed_ms(by_tag=by_tag)

    # ------------------------------------------------------------------ state dicts / checkpoints
    def _gather_flat(self, u: _Unit, shard: torch.Tensor) -> torch.Tensor:
        """All-gather a shard-layout tensor (any dtype) into the full layout, through torch.distributed (off the hot path)."""
        parts = [torch.empty_like(shard) for _ in range(self.world)]
        torch.distributed.all_gather(parts, shard.contiguous(), group=self.group)
        full = torch.empty(u.padded, dtype=shard.dtype, device=shard.device)
        for s in u.segs:
            for q in range(self.world):
                full[s.full_off + q * s.piece : s.full_off + (q + 1) * s.piece].copy_(parts[q][s.shard_off : s.shard_off + s.piece])
        return full

    def _unit_tensors(self, u: _Unit, flat: torch.Tensor):
        return {s.name: flat[s.full_off : s.full_off + s.numel].view(s.shape) for s in u.segs}

    def full_state_dict(self, cpu: bool = True, dtype: Optional[torch.dtype] = None) -> dict:
        """name -> full tensor, gathered one unit at a time from the fp32 master shards (so a save / load round trip is exact);
        ``dtype`` down-casts on request.  On CPU by default."""
        out = {}
        for u in self._units:
            full = self._gather_flat(u, u.master)
            for n, t in self._unit_tensors(u, full).items():
                t = t if dtype is None else t.to(dtype)
                out[n] = t.detach().to("cpu", copy=True) if cpu else t.detach().clone()
            del full
        for n, b in self.module.named_buffers():
            out[n] = b.detach().cpu() if cpu else b.detach()
        extra = {id(p) for p in self._ignored_params} | {id(p) for p in self._frozen}
        for n, q in self.module.named_parameters():
            if id(q) in extra:
                out[n] = q.detach().cpu() if cpu else q.detach()
        return out

    def state_dict(self, *args, **kwargs):
        return self.full_state_dict()

    @torch.no_grad()
    def load_full_state_dict(self, sd: dict, strict: bool = True):
        """Every rank passes the same full dict — or only rank 0 passes values, the others an empty dict, and every tensor rank 0
        holds is broadcast (``cpu_ram_efficient_loading``, ``load_checkpoint_in_model(..., broadcast_from_rank0=True)``).  ``sd`` may
        hold only part of the model (one file of a sharded checkpoint) when ``strict`` is off: what it does not name stays as it is."""
        have = torch.tensor([1 if sd else 0], device=self.device)
        counts = [torch.zeros_like(have) for _ in range(self.world)]
        torch.distributed.all_gather(counts, have, group=self.group)
        bcast = any(int(c) == 0 for c in counts) and any(int(c) for c in counts)
        src = torch.distributed.get_global_rank(self.group, 0)
        if bcast:
            names = [sorted(sd)] if self.rank == 0 else [None]
            torch.distributed.broadcast_object_list(names, src=src, group=self.group)
            present = set(names[0])
        else:
            present = set(sd)
        missing = [s.name for u in self._units for s in u.segs if s.name not in present]
        if missing and strict:
            raise KeyError(f"missing key {missing[0]!r} in state dict ({len(missing)} missing in all)")
        for u in self._units:
            touched = False
            for s in u.segs:
                if s.name not in present:
                    continue
                touched = True
                if bcast:
                    full = torch.empty(s.numel, dtype=torch.float32, device=self.device)
                    if self.rank == 0:
                        full.copy_(sd[s.name].reshape(-1).float())
                    torch.distributed.broadcast(full, src=src, group=self.group)
                else:
                    full = sd[s.name].reshape(-1)
                lo = self.rank * s.piece
                n = max(0, min(s.piece, s.numel - lo))
                dst = u.master[s.shard_off : s.shard_off + s.piece]
                dst.zero_()
                if n > 0:
                    dst[:n].copy_(full[lo : lo + n].to(self.device, torch.float32))
            if touched:
                u.shard.copy_(u.master)
        for n, b in self.module.named_buffers():
            if n not in present:
                continue
            if bcast:
                value = b.detach().clone()
                if self.rank == 0:
                    value.copy_(sd[n])
                torch.distributed.broadcast(value, src=src, group=self.group)
                b.copy_(value)
            else:
                b.copy_(sd[n])
        self._sync_after_external_update()

    def load_state_dict(self, sd, strict: bool = True):
        self.load_full_state_dict(sd, strict=strict)

    def _sync_after_external_update(self):
        if self.use_symm:
            torch.cuda.synchronize(self.device)
            self.ws.barrier(self.ch_step)
            torch.cuda.synchronize(self.device)
        else:
            torch.distributed.barrier(group=self.group)

    # ---- torch.distributed.checkpoint (DCP) views -------------------------------------------------
    def _mesh(self):
        m = getattr(self, "_dcp_mesh", None)
        if m is None:
            from torch.distributed.device_mesh import DeviceMesh

            m = DeviceMesh.from_group(self.group, "cuda" if self.use_symm else "cpu")
            self._dcp_mesh = m
        return m

    def _as_dtensor(self, u: _Unit, flat: torch.Tensor, s: _Seg):
        """Shard(0) DTensor over my piece of parameter `s` taken from the shard-layout tensor `flat`; parameters that are not cut
        along dim 0 in whole rows (odd shapes) are gathered and returned as plain (replicated) tensors."""
        from torch.distributed.tensor import DTensor, Shard

        even_1d = s.rows == 0 and len(s.shape) == 1 and s.numel % self.world == 0 and s.piece == s.numel // self.world
        if s.rows or even_1d:
            shape = s.piece_shape if s.rows else (s.piece,)
            local = flat[s.shard_off : s.shard_off + s.piece].view(shape)
            return DTensor.from_local(local, self._mesh(), [Shard(0)], run_check=False)
        return None

    def sharded_state_dict(self, which: str = "master") -> dict:
        """FQN -> DTensor (or full tensor for oddly shaped parameters) over ``master`` / ``exp_avg`` / ``exp_avg_sq``."""
        out = {}
        for u in self._units:
            flat = getattr(u, which)
            gathered = None
            for s in u.segs:
                dt = self._as_dtensor(u, flat, s)
                if dt is None:
                    if gathered is None:
                        gathered = self._gather_flat(u, flat)
                    dt = gathered[s.full_off : s.full_off + s.numel].view(s.shape).clone()
                out[s.name] = dt
        return out

    def _load_sharded_into(self, which: str, loaded: dict):
        for u in self._units:
            flat = getattr(u, which)
            for s in u.segs:
                t = loaded[s.name]
                if hasattr(t, "to_local"):
                    flat[s.shard_off : s.shard_off + s.piece].copy_(t.to_local().reshape(-1))
                else:
                    lo = self.rank * s.piece
                    n = max(0, min(s.piece, s.numel - lo))
                    if n > 0:
                        flat[s.shard_off : s.shard_off + n].copy_(t.reshape(-1)[lo : lo + n])

    def save_sharded_state(self, output_dir: str, index: int, plugin=None):
        """``pytorch_model_fsdp_{i}/`` (DCP directory, ``{"model": state_dict}``) or ``pytorch_model_fsdp.bin`` (full)."""
        sdt = str(getattr(plugin or self.plugin, "state_dict_type", "SHARDED_STATE_DICT")).split(".")[-1]
        if sdt == "FULL_STATE_DICT":
            sd = self.full_state_dict()
            if torch.distributed.get_rank() == 0:
                name = f"{FSDP_MODEL_NAME}.bin" if index == 0 else f"{FSDP_MODEL_NAME}_{index}.bin"
                torch.save(sd, os.path.join(output_dir, name))
            torch.distributed.barrier()
            return
        import torch.distributed.checkpoint as dist_cp

        d = os.path.join(output_dir, f"{FSDP_MODEL_NAME}_{index}")
        os.makedirs(d, exist_ok=True)
        sd = self.sharded_state_dict("master")
        for n, b in self.module.named_buffers():
            sd[n] = b.detach()
        dist_cp.save({"model": sd}, storage_writer=dist_cp.FileSystemWriter(d), process_group=self.group if self.dp_group is None else None)

    @torch.no_grad()
    def load_sharded_state(self, input_dir: str, index: int, plugin=None):
        sdt = str(getattr(plugin or self.plugin, "state_dict_type", "SHARDED_STATE_DICT")).split(".")[-1]
        full_name = os.path.join(input_dir, f"{FSDP_MODEL_NAME}.bin" if index == 0 else f"{FSDP_MODEL_NAME}_{index}.bin")
        d = os.path.join(input_dir, f"{FSDP_MODEL_NAME}_{index}")
        if sdt == "FULL_STATE_DICT" or (os.path.exists(full_name) and not os.path.isdir(d)):
            self.load_full_state_dict(torch.load(full_name, map_location="cpu", weights_only=True))
            return
        import torch.distributed.checkpoint as dist_cp

        target = self.sharded_state_dict("master")  # DCP loads in place into these views of the master shards
        holder = {"model": target}
        dist_cp.load(holder, storage_reader=dist_cp.FileSystemReader(d), process_group=self.group if self.dp_group is None else None)
        self._load_sharded_into("master", holder["model"])
        for u in self._units:
            u.shard.copy_(u.master)
        self._sync_after_external_update()

    # optimizer state ------------------------------------------------------------------------------
    def _param_index(self, optimizer):
        index_of, idx = {}, 0
        reverse = {id(sp): pid for pid, sp in self._shard_params.items()}
        for g in optimizer.param_groups:
            for p in g["params"]:
                index_of[reverse.get(id(p), id(p))] = idx
                idx += 1
        return index_of

    def optimizer_state_dict(self, optimizer) -> dict:
        """Full (gathered) optimizer state in torch's layout: usable as ``optimizer.bin``."""
        index_of = self._param_index(optimizer)
        state = {}
        if self._generic:
            for u in self._units:
                for s in u.segs:
                    sp = self._shard_params[id(s.param)]
                    st = optimizer.state.get(sp, {})
                    entry = {}
                    for k, v in st.items():
                        if torch.is_tensor(v) and v.numel() == s.piece and s.piece > 1:
                            flat = torch.zeros(u.shard_numel, dtype=v.dtype, device=self.device)
                            flat[s.shard_off : s.shard_off + s.piece].copy_(v.reshape(-1))
                            full = self._gather_flat(u, flat)
                            entry[k] = full[s.full_off : s.full_off + s.numel].view(s.shape).cpu()
                        else:
                            entry[k] = v.cpu() if torch.is_tensor(v) else v
                    if entry and id(s.param) in index_of:
                        state[index_of[id(s.param)]] = entry
        else:
            for u in self._units:
                full_m = self._gather_flat(u, u.exp_avg)
                full_v = self._gather_flat(u, u.exp_avg_sq)
                for s in u.segs:
                    if id(s.param) in index_of:
                        state[index_of[id(s.param)]] = {"step": torch.tensor(float(self.step_count)),
                                                        "exp_avg": full_m[s.full_off : s.full_off + s.numel].view(s.shape).cpu(),
                                                        "exp_avg_sq": full_v[s.full_off : s.full_off + s.numel].view(s.shape).cpu()}
        groups, c = [], 0
        for g in optimizer.param_groups:
            ng = {k: v for k, v in g.items() if k != "params"}
            ng["params"] = list(range(c, c + len(g["params"])))
            c += len(g["params"])
            groups.append(ng)
        return {"state": state, "param_groups": groups}

    def load_optimizer_state_dict(self, optimizer, sd: dict):
        index_of = self._param_index(optimizer)
        for u in self._units:
            for s in u.segs:
                st = sd["state"].get(index_of.get(id(s.param), -1))
                if st is None:
                    continue
                lo = self.rank * s.piece
                n = max(0, min(s.piece, s.numel - lo))
                if self._generic:
                    sp = self._shard_params[id(s.param)]
                    entry = {}
                    for k, v in st.items():
                        if torch.is_tensor(v) and v.numel() == s.numel and s.numel > 1:
                            piece = torch.zeros(s.piece, dtype=v.dtype, device=self.device)
                            if n > 0:
                                piece[:n].copy_(v.reshape(-1)[lo : lo + n])
                            entry[k] = piece.view(sp.shape)
                        else:
                            entry[k] = v.to(self.device) if torch.is_tensor(v) and v.dim() > 0 else v
                    optimizer.state[sp] = entry
                else:
                    for key, flat in (("exp_avg", u.exp_avg), ("exp_avg_sq", u.exp_avg_sq)):
                        dst = flat[s.shard_off : s.shard_off + s.piece]
                        dst.zero_()
                        if n > 0:
                            dst[:n].copy_(st[key].reshape(-1)[lo : lo + n].to(self.device, torch.float32))
                    self.step_count = int(float(st["step"]))
        for g_new, g_old in zip(optimizer.param_groups, sd["param_groups"]):
            for k, v in g_old.items():
                if k != "params":
                    g_new[k] = v

    def _sharded_optimizer_state(self, inner) -> dict:
        """{"state": {fqn: {...DTensors...}}, "param_groups": [...]} keyed by FQN like torch's get_optimizer_state_dict."""
        state = {}
        if self._generic:
            for u in self._units:
                for s in u.segs:
                    sp = self._shard_params[id(s.param)]
                    entry = {}
                    for k, v in inner.state.get(sp, {}).items():
                        if torch.is_tensor(v) and v.numel() == s.piece and s.piece > 1:
                            flat = torch.zeros(u.shard_numel, dtype=v.dtype, device=self.device)
                            flat[s.shard_off : s.shard_off + s.piece].copy_(v.reshape(-1))
                            dt = self._as_dtensor(u, flat, s)
                            entry[k] = dt if dt is not None else self._gather_flat(u, flat)[s.full_off : s.full_off + s.numel].view(s.shape).clone()
                        elif torch.is_tensor(v):
                            entry[k] = v.detach().clone()
                        else:
                            entry[k] = v
                    if entry:
                        state[s.name] = entry
        else:
            m, v = self.sharded_state_dict("exp_avg"), self.sharded_state_dict("exp_avg_sq")
            for u in self._units:
                for s in u.segs:
                    state[s.name] = {"step": torch.tensor(float(self.step_count)), "exp_avg": m[s.name], "exp_avg_sq": v[s.name]}
        name_of = {id(s.param): s.name for u in self._units for s in u.segs}
        name_of.update({id(sp): name_of[pid] for pid, sp in self._shard_params.items()})
        groups = []
        for g in inner.param_groups:
            ng = {k: v for k, v in g.items() if k != "params"}
            ng["params"] = [name_of.get(id(p), str(i)) for i, p in enumerate(g["params"])]
            groups.append(ng)
        return {"state": state, "param_groups": groups}

    def save_sharded_optimizer(self, optimizer, output_dir: str, index: int, plugin=None):
        sdt = str(getattr(plugin or self.plugin, "state_dict_type", "SHARDED_STATE_DICT")).split(".")[-1]
        inner = optimizer.optimizer if hasattr(optimizer, "optimizer") else optimizer
        if sdt == "FULL_STATE_DICT":
            sd = self.optimizer_state_dict(inner)
            if torch.distributed.get_rank() == 0:
                torch.save(sd, os.path.join(output_dir, f"{OPTIMIZER_NAME}.bin" if index == 0 else f"{OPTIMIZER_NAME}_{index}.bin"))
            torch.distributed.barrier()
            return
        import torch.distributed.checkpoint as dist_cp

        d = os.path.join(output_dir, f"{OPTIMIZER_NAME}_{index}")
        os.makedirs(d, exist_ok=True)
        dist_cp.save({"optimizer": self._sharded_optimizer_state(inner)}, storage_writer=dist_cp.FileSystemWriter(d),
                     process_group=self.group if self.dp_group is None else None)

    def load_sharded_optimizer(self, optimizer, input_dir: str, index: int, plugin=None):
        inner = optimizer.optimizer if hasattr(optimizer, "optimizer") else optimizer
        d = os.path.join(input_dir, f"{OPTIMIZER_NAME}_{index}")
        if not os.path.isdir(d):
            f = os.path.join(input_dir, f"{OPTIMIZER_NAME}.bin" if index == 0 else f"{OPTIMIZER_NAME}_{index}.bin")
            self.load_optimizer_state_dict(inner, torch.load(f, map_location="cpu", weights_only=False))
            return
        import torch.distributed.checkpoint as dist_cp

        if self._generic:
            # make sure the user's optimizer has state tensors to load into (zero gradients: no parameter moves for the usual rules)
            if not inner.state:
                for sp in self._shard_params.values():
                    sp.grad = torch.zeros_like(sp)
                saved = [u.master.clone() for u in self._units]
                inner.step()
                for u, m in zip(self._units, saved):
                    u.master.copy_(m)
                for sp in self._shard_params.values():
                    sp.grad = None
        target = self._sharded_optimizer_state(inner)
        holder = {"optimizer": target}
        # shards written at another world size are re-cut by DCP's load planner
        dist_cp.load(holder, storage_reader=dist_cp.FileSystemReader(d), process_group=self.group if self.dp_group is None else None)
        got = holder["optimizer"]
        if self._generic:
            for u in self._units:
                for s in u.segs:
                    sp = self._shard_params[id(s.param)]
                    for k, v in got["state"].get(s.name, {}).items():
                        cur = inner.state[sp].get(k)
                        if hasattr(v, "to_local"):
                            cur.copy_(v.to_local().reshape(cur.shape))
                        elif torch.is_tensor(v) and torch.is_tensor(cur) and v.numel() == s.numel and s.numel > 1:
                            lo = self.rank * s.piece
                            n = max(0, min(s.piece, s.numel - lo))
                            cur.reshape(-1)[:n].copy_(v.reshape(-1)[lo : lo + n])
                        elif torch.is_tensor(cur):
                            cur.copy_(v)
                        else:
                            inner.state[sp][k] = v
        else:
            self._load_sharded_into("exp_avg", {n: st["exp_avg"] for n, st in got["state"].items()})
            self._load_sharded_into("exp_avg_sq", {n: st["exp_avg_sq"] for n, st in got["state"].items()})
            steps = [float(st["step"]) for st in got["state"].values() if "step" in st]
            if steps:
                self.step_count = int(max(steps))
        for g_new, g_old in zip(inner.param_groups, got["param_groups"]):
            for k, v in g_old.items():
                if k != "params":
                    g_new[k] = v

    def release(self):
        F.unregister_carry_source(self)
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
    """Rebuild the full name -> tensor dict from a DCP ``pytorch_model_fsdp_{i}/`` directory without a process group
    (what ``accelerate merge-weights`` runs; reads checkpoints written by this engine or by the reference's FSDP2)."""
    import torch.distributed.checkpoint as dist_cp
    from torch.distributed.checkpoint.default_planner import _EmptyStateDictLoadPlanner
    from torch.distributed.checkpoint.state_dict_loader import _load_state_dict

    sd: dict = {}
    _load_state_dict(sd, storage_reader=dist_cp.FileSystemReader(checkpoint_dir), planner=_EmptyStateDictLoadPlanner(), no_dist=True)
    model = sd.get("model", sd)
    return dict(model)


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
