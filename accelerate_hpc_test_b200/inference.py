"""Pipeline-parallel inference (``prepare_pippy``): every process keeps one contiguous slice of the model and micro-batches
stream through the slices.

Two back-ends behind the same call (API parity: reference ``src/accelerate/inference.py:126-186``):

* **native stages** — for models that can list their blocks in execution order (``nn.Sequential``, or anything with a
  ``pipeline_blocks()`` method such as the native Llama).  No tracing: rank *r* simply owns blocks ``[lo_r, hi_r)``.
  Activations travel as point-to-point messages (NCCL p2p on GPUs, gloo on CPU) and are **overlapped with compute**: all
  receives of a call are posted up front into per-micro-batch buffers, each send is asynchronous and only waited for at the
  end, so while stage *r* computes micro-batch *i* its send of *i-1* and its receive of *i+1* are in flight
  (this path stays on the library's p2p by specification, the overlap is ours).
* **traced stages** — any other model is cut with ``torch.distributed.pipelining`` at the split points and run with its
  GPipe schedule, as the reference does.

In both cases rank 0 feeds the inputs and the last rank holds the output; ``gather_output=True`` copies it to every rank.
"""

from __future__ import annotations

import math
from typing import Callable, Optional, Sequence, Union

import torch
import torch.distributed as dist
import torch.nn as nn

from .state import PartialState
from .utils.modeling import calculate_maximum_sizes, infer_auto_device_map
from .utils.operations import copy_tensor_to_devices, ignorant_find_batch_size, pad_input_tensors, send_to_device
from .utils.other import convert_bytes

_DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32, torch.bool, torch.float64, torch.uint8]


# =========================================================================== where to cut
def _even_budget(model, parts: int) -> dict:
    """Per-part memory budget: an even share of the model plus room for its largest indivisible layer, rounded up 10 %."""
    total, (largest, _names) = calculate_maximum_sizes(model)
    amount, unit = convert_bytes((total + largest) / parts).split(" ")
    per_part = f"{math.ceil(float(amount)) * 1.1} {unit}"
    return dict.fromkeys(range(parts), per_part)


def generate_device_map(model, num_processes: int = 1, no_split_module_classes=None, max_memory: dict = None):
    """Device map that spreads ``model`` over ``num_processes`` virtual devices of equal size."""
    if num_processes > 1 and max_memory is None:
        max_memory = _even_budget(model, num_processes)
    kwargs = {} if num_processes == 1 else {"max_memory": max_memory}
    return infer_auto_device_map(model, no_split_module_classes=no_split_module_classes, clean_result=False, **kwargs)


def _auto_split_points(model, parts: int, no_split_module_classes) -> list:
    placement = generate_device_map(model, parts, no_split_module_classes=no_split_module_classes)
    firsts = []
    for part in range(1, parts):
        firsts.append(next(name for name, where in placement.items() if where == part))
    return firsts


def find_pippy_batch_size(args, kwargs):
    """Batch size of the first tensor found in the positional, then the keyword arguments."""
    for value in list(args or ()) + list((kwargs or {}).values()):
        n = ignorant_find_batch_size(value)
        if n is not None:
            return n
    return None


# =========================================================================== native stages (no tracing)
class Block:
    """One step of a stageable model: ``fn(*tensors) -> tuple of tensors``; ``path`` is the module path it starts at."""

    def __init__(self, path: str, fn: Callable, module: Optional[nn.Module] = None):
        self.path, self.fn, self.module = path, fn, module


def _blocks_of(model) -> Optional[list]:
    if hasattr(model, "pipeline_blocks"):
        return list(model.pipeline_blocks())
    if isinstance(model, nn.Sequential):
        return [Block(name, (lambda x, _m=m: (_m(x),)), m) for name, m in model.named_children()]
    return None


def _cut(blocks: list, split_points: Sequence[str], parts: int) -> list:
    """Index ranges ``[(lo, hi), ...]`` of the blocks owned by each part; a split point names the first block of a part (a
    path inside a block counts for that block)."""
    starts = [0]
    for point in split_points:
        hit = next((i for i, b in enumerate(blocks) if b.path == point or point.startswith(b.path + ".") or b.path.startswith(point + ".")), None)
        if hit is None or hit <= starts[-1]:
            hit = min(len(blocks) - (parts - len(starts)), starts[-1] + 1)
        starts.append(hit)
    while len(starts) < parts:
        starts.append(min(len(blocks), starts[-1] + 1))
    bounds = starts + [len(blocks)]
    return [(bounds[i], bounds[i + 1]) for i in range(parts)]


def _balanced_cut(blocks: list, parts: int) -> list:
    """Contiguous ranges with (nearly) equal parameter bytes."""
    weights = []
    for b in blocks:
        mods = b.module if isinstance(b.module, (list, tuple)) else [b.module]
        weights.append(sum(p.numel() * p.element_size() for m in mods if m is not None for p in m.parameters()) + 1)
    total, bounds, acc, want = sum(weights), [0], 0, 1
    for i, w in enumerate(weights):
        acc += w
        remaining_blocks, remaining_parts = len(blocks) - (i + 1), parts - want
        if want < parts and (acc >= total * want / parts or remaining_blocks == remaining_parts):
            bounds.append(i + 1)
            want += 1
    bounds = bounds[:parts] + [len(blocks)]
    return [(bounds[i], bounds[i + 1]) for i in range(parts)]


class NativePipeline:
    """This rank's slice of the blocks plus the overlapped send / receive machinery."""

    def __init__(self, blocks: list, ranges: list, num_chunks: int, state: PartialState):
        self.state = state
        self.rank, self.world = state.process_index, state.num_processes
        self.lo, self.hi = ranges[self.rank]
        self.blocks = blocks[self.lo : self.hi]
        self.num_chunks = num_chunks
        self.device = state.device
        for b in self.blocks:
            mods = b.module if isinstance(b.module, (list, tuple)) else [b.module]
            for m in mods:
                if m is not None:
                    m.to(self.device)

    # -- wire format: one int64 header per call (tensor count, then ndim / dtype / dims of each), then the tensors per chunk
    def _send_header(self, sample: tuple) -> None:
        fields = [len(sample)]
        for t in sample:
            fields += [t.dim(), _DTYPES.index(t.dtype), *t.shape]
        head = torch.zeros(64, dtype=torch.int64)
        head[0] = len(fields)
        head[1 : 1 + len(fields)] = torch.tensor(fields)
        dist.send(head.to(self._wire_device()), dst=self.rank + 1)

    def _recv_header(self) -> list:
        head = torch.zeros(64, dtype=torch.int64, device=self._wire_device())
        dist.recv(head, src=self.rank - 1)
        fields = head.cpu().tolist()
        fields = fields[1 : 1 + fields[0]]
        specs, i = [], 1
        for _ in range(fields[0]):
            nd, code = fields[i], fields[i + 1]
            specs.append((tuple(fields[i + 2 : i + 2 + nd]), _DTYPES[code]))
            i += 2 + nd
        return specs

    def _wire_device(self):
        return self.device if dist.get_backend() == "nccl" else torch.device("cpu")

    def _run_blocks(self, tensors: tuple) -> tuple:
        for b in self.blocks:
            out = b.fn(*tensors)
            tensors = out if isinstance(out, tuple) else (out,)
        return tensors

    @torch.no_grad()
    def step(self, *args, **kwargs):
        first, last = self.rank == 0, self.rank == self.world - 1
        n = self.num_chunks
        wire = self._wire_device()
        inbox = None
        if first:
            chunks = _split_inputs(args, kwargs, n)
        else:
            specs = self._recv_header()
            inbox = [[torch.empty(shape, dtype=dtype, device=wire) for shape, dtype in specs] for _ in range(n)]
            pending_in = [[dist.irecv(t, src=self.rank - 1) for t in bufs] for bufs in inbox]  # all receives in flight at once
        pending_out, outputs, keep = [], [], []
        for i in range(n):
            if first:
                a, kw = chunks[i]
                tensors = self.blocks[0].fn(*send_to_device(a, self.device), **send_to_device(kw, self.device))
                tensors = tensors if isinstance(tensors, tuple) else (tensors,)
                for b in self.blocks[1:]:
                    out = b.fn(*tensors)
                    tensors = out if isinstance(out, tuple) else (out,)
            else:
                for req in pending_in[i]:
                    req.wait()
                tensors = self._run_blocks(tuple(t.to(self.device) for t in inbox[i]))
            if last:
                outputs.append(tensors)
                continue
            if i == 0:
                self._send_header(tensors)
            payload = [t.contiguous().to(wire) for t in tensors]
            keep.append(payload)  # buffers must outlive their asynchronous sends
            pending_out += [dist.isend(t, dst=self.rank + 1) for t in payload]
        for req in pending_out:
            req.wait()
        if not last:
            return None
        merged = tuple(torch.cat([o[j] for o in outputs], dim=0) for j in range(len(outputs[0])))
        return merged[0] if len(merged) == 1 else merged


def _split_inputs(args, kwargs, n: int) -> list:
    """Cut every tensor argument into ``n`` micro-batches along dim 0 (non-tensors are repeated)."""

    def piece(value, i):
        if torch.is_tensor(value) and value.dim() > 0 and value.shape[0] >= n:
            return value.chunk(n, dim=0)[i]
        return value

    return [(tuple(piece(a, i) for a in args), {k: piece(v, i) for k, v in kwargs.items()}) for i in range(n)]


# =========================================================================== traced stages (generic models)
def build_pipeline(model, split_points, args, kwargs, num_chunks):
    """Trace ``model``, cut it in front of every module named in ``split_points`` and return a GPipe schedule over this
    rank's stage (``schedule.step`` runs one call)."""
    from torch.distributed.pipelining import ScheduleGPipe, SplitPoint, pipeline

    me = PartialState()
    cuts = dict.fromkeys(split_points, SplitPoint.BEGINNING)
    traced = pipeline(model, mb_args=args, mb_kwargs=kwargs, split_spec=cuts)
    return ScheduleGPipe(traced.build_stage(me.local_process_index, device=me.device), num_chunks)


# =========================================================================== the call a prepared model makes
def pippy_forward(forward, num_chunks, gather_output, *args, **kwargs):
    """Role-dependent call of one pipeline step: the first rank pads the batch to a multiple the schedule can chunk and feeds
    it, the last rank returns the result, the ranks in between just take part."""
    me = PartialState()
    result = None
    if me.num_processes == 1:
        result = forward(*args, **kwargs)
    elif me.is_local_main_process:
        batch = find_pippy_batch_size(args, kwargs)
        if batch is None:
            raise ValueError("Could not find batch size from args or kwargs")
        if batch != num_chunks:
            args = pad_input_tensors(args, batch, num_chunks)
            kwargs = pad_input_tensors(kwargs, batch, num_chunks)
        forward(*args, **kwargs)
    else:
        out = forward()
        if me.is_last_process:
            result = out
    if gather_output:
        result = copy_tensor_to_devices(result)  # broadcast from the last stage
    return result


def prepare_pippy(model, split_points: Optional[Union[str, list]] = "auto", no_split_module_classes: Optional[list] = None,
                  example_args: Optional[tuple] = (), example_kwargs: Optional[dict] = None, num_chunks: Optional[int] = None,
                  gather_output: Optional[bool] = False):
    """Turn ``model`` into this process's pipeline stage; afterwards ``model(...)`` runs one pipelined forward.

    ``split_points``: ``"auto"`` (even spread by parameter size) or the module paths that start stages 1, 2, ….
    ``num_chunks``: micro-batches per call (default: one per process).  ``example_args`` / ``example_kwargs`` are only needed
    for models that have to be traced.
    """
    me = PartialState()
    parts = me.num_processes
    if num_chunks is None:
        num_chunks = parts
    blocks = _blocks_of(model)
    if blocks is not None and len(blocks) >= parts:
        if split_points == "auto":
            ranges = _balanced_cut(blocks, parts)
            split_points = [blocks[lo].path for lo, _ in ranges[1:]]
        else:
            ranges = _cut(blocks, split_points, parts)
        runner = NativePipeline(blocks, ranges, num_chunks, me)
        step = runner.step
        model.pippy_stage = runner
    else:
        example_args = send_to_device(example_args, "cpu")
        example_kwargs = send_to_device(example_kwargs, "cpu")
        if split_points == "auto":
            split_points = _auto_split_points(model, num_chunks, no_split_module_classes)
        schedule = build_pipeline(model, split_points, example_args, example_kwargs, num_chunks)
        step = schedule.step
        model.pippy_stage = schedule
    model.hf_split_points = split_points
    model._original_forward = model.forward
    model._original_call = model.__call__

    def forward(*args, **kwargs):
        return pippy_forward(step, num_chunks, gather_output, *args, **kwargs)

    forward.__wrapped__ = model._original_forward
    model.forward = forward
    return model
