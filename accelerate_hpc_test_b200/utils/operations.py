"""Nested-structure helpers and the small collectives the framework itself issues.

Contract parity with reference src/accelerate/utils/operations.py: ``recursively_apply`` :84-133,
``send_to_device`` :136-195, ``gather`` :419-432 (``all_gather_into_tensor`` on NCCL, list gather on
gloo :332-350), ``gather_object`` :435-462, ``broadcast`` :539-560, ``broadcast_object_list``
:563-580, ``slice_tensors`` / ``concatenate`` / ``pad_across_processes`` :583-695,
``pad_input_tensors`` :698-724, ``reduce`` :727-766, ``convert_outputs_to_fp32`` :769-828 and the
debug-mode shape verification :364-416.  These carry metrics, RNG states and batch structures —
kilobytes — so they stay on torch.distributed; the bulk traffic (gradients, parameters,
activations) goes through the fused kernels in :mod:`..parallel`.
"""

from __future__ import annotations

import pickle
import warnings
from collections.abc import Mapping
from contextlib import contextmanager
from functools import update_wrapper, wraps
from typing import Any, Callable

import torch

from ._lazy_state import PartialState
from .dataclasses import DistributedType, TensorInformation


class DistributedOperationException(Exception):
    """Raised (in debug mode) when ranks disagree on the shapes going into a collective."""


def is_torch_tensor(tensor) -> bool:
    return isinstance(tensor, torch.Tensor)


def is_tensor_information(tensor_info) -> bool:
    return isinstance(tensor_info, TensorInformation)


def is_namedtuple(data) -> bool:
    return isinstance(data, tuple) and hasattr(data, "_asdict") and hasattr(data, "_fields")


def honor_type(obj, generator):
    """A sequence of ``obj``'s own type (list, tuple, namedtuple) holding the items of ``generator``."""
    items = list(generator)
    return type(obj)(*items) if is_namedtuple(obj) else type(obj)(items)


# The framework's batches, metrics and states are *trees*: lists / tuples / namedtuples / mappings nested to any depth with tensors
# (and arbitrary other objects) at the leaves.  Everything below is written on two primitives: rebuild a tree with its leaves mapped,
# and walk the leaves in order.
def _map_leaves(tree, on_leaf: Callable, skip_keys=()):
    if isinstance(tree, (list, tuple)):
        return honor_type(tree, (_map_leaves(item, on_leaf, skip_keys) for item in tree))
    if isinstance(tree, Mapping):
        return type(tree)({key: value if key in skip_keys else _map_leaves(value, on_leaf, skip_keys) for key, value in tree.items()})
    return on_leaf(tree)


def _leaves(tree):
    if isinstance(tree, (list, tuple)):
        for item in tree:
            yield from _leaves(item)
    elif isinstance(tree, Mapping):
        for value in tree.values():
            yield from _leaves(value)
    else:
        yield tree


def recursively_apply(func, data, *args, test_type=is_torch_tensor, error_on_other_type=False, **kwargs):
    """``data`` with ``func(leaf, *args, **kwargs)`` in place of every leaf that passes ``test_type``; other leaves are kept, or
    rejected with ``error_on_other_type``."""

    def on_leaf(leaf):
        if test_type(leaf):
            return func(leaf, *args, **kwargs)
        if error_on_other_type:
            raise TypeError(f"Unsupported types ({type(leaf)}) passed to `{func.__name__}`. Only nested list/tuple/dicts of objects that "
                            f"are valid for `{test_type.__name__}` should be passed.")
        return leaf

    return _map_leaves(data, on_leaf)


def send_to_device(tensor, device, non_blocking: bool = False, skip_keys=None):
    """The tree with everything that has a ``.to`` moved to ``device`` (a whole ``BatchEncoding``-like container moves itself);
    mapping entries named in ``skip_keys`` stay where they are."""
    if isinstance(skip_keys, str):
        skip_keys = [skip_keys]

    def move(leaf):
        if not hasattr(leaf, "to"):
            return leaf
        try:
            return leaf.to(device, non_blocking=non_blocking)
        except TypeError:  # a `.to` that does not know `non_blocking`
            return leaf.to(device)

    if not isinstance(tensor, (list, tuple, Mapping)) or hasattr(tensor, "to"):
        return move(tensor)
    return _map_leaves(tensor, move, tuple(skip_keys or ()))


def get_data_structure(data):
    """The same tree with a ``TensorInformation`` (shape, dtype) in place of every tensor — what rank 0 announces before a broadcast."""
    return recursively_apply(lambda t: TensorInformation(shape=t.shape, dtype=t.dtype), data)


def get_shape(data):
    return recursively_apply(lambda t: list(t.shape), data)


def initialize_tensors(data_structure):
    """Inverse of ``get_data_structure``: uninitialised tensors of the announced shapes and dtypes."""
    return recursively_apply(lambda info: torch.empty(*info.shape, dtype=info.dtype), data_structure, test_type=is_tensor_information)


def find_batch_size(data):
    """Dim 0 of the first leaf of a batch tree; that leaf has to be a tensor."""
    if isinstance(data, (list, tuple, Mapping)) and len(data) == 0:
        raise ValueError(f"Cannot find the batch size from empty {type(data)}.")
    first = next(_leaves(data), None)
    if not isinstance(first, torch.Tensor):
        raise TypeError(f"Can only find the batch size of tensors but got {type(first)}.")
    return first.shape[0]


def ignorant_find_batch_size(data):
    """``find_batch_size`` or None."""
    try:
        return find_batch_size(data)
    except (ValueError, TypeError):
        return None


def listify(data):
    """Tensors → nested Python lists (bf16 through fp32: ``tolist`` does not know bf16 on every build)."""

    def to_list(t):
        t = t.detach().cpu()
        return (t.float() if t.dtype == torch.bfloat16 else t).tolist()

    return recursively_apply(to_list, data)


# ------------------------------------------------------------------------------- debug verification
def verify_operation(function):
    """In ``ACCELERATE_DEBUG_MODE`` all-gather the operand shapes first and raise on mismatch."""

    @wraps(function)
    def wrapper(*args, **kwargs):
        state = PartialState()
        if state.distributed_type == DistributedType.NO or not state.debug:
            return function(*args, **kwargs)
        operation = f"{function.__module__}.{function.__name__}"
        tensor = kwargs["tensor"] if "tensor" in kwargs else args[0]
        if state.device.type != find_device(tensor).type:
            raise DistributedOperationException(
                f"One or more of the tensors passed to {operation} were not on the {tensor.device.type} while the "
                f"`Accelerator` is configured for {state.device.type}. Please move it to the {state.device.type} before calling {operation}."
            )
        shapes = get_shape(tensor)
        output = gather_object([shapes])
        if output[0] is not None and output.count(output[0]) != len(output):
            listing = "\n  - ".join(f"Process {i}: {s}" for i, s in enumerate(output))
            raise DistributedOperationException(
                f"Cannot apply desired operation due to shape mismatches. All shapes across devices must be valid.\n\n"
                f"Operation: `{operation}`\nInput shapes:\n  - {listing}"
            )
        return function(*args, **kwargs)

    return wrapper


def chained_operation(function):
    """Re-raise ``DistributedOperationException`` with the outer operation named."""

    @wraps(function)
    def wrapper(*args, **kwargs):
        try:
            return function(*args, **kwargs)
        except DistributedOperationException as e:
            operation = f"{function.__module__}.{function.__name__}"
            raise DistributedOperationException(f"Error found while calling `{operation}`. Please see the earlier error for more details.") from e

    return wrapper


def find_device(data):
    """Device of the first tensor in the tree (None when it holds none)."""
    return next((leaf.device for leaf in _leaves(data) if isinstance(leaf, torch.Tensor)), None)


# ------------------------------------------------------------------------------- gather
def _backend_has_into_tensor() -> bool:
    backend = str(torch.distributed.get_backend())
    return "nccl" in backend


def _gather_one(tensor: torch.Tensor) -> torch.Tensor:
    state = PartialState()
    if tensor.ndim == 0:
        tensor = tensor.clone()[None]
    if not tensor.is_contiguous():
        tensor = tensor.contiguous()
    world = state.num_processes
    if _backend_has_into_tensor() and tensor.is_cuda:
        out = torch.empty(world * tensor.numel(), dtype=tensor.dtype, device=tensor.device)
        torch.distributed.all_gather_into_tensor(out, tensor)
        return out.view(-1, *tensor.size()[1:])
    outs = [torch.empty_like(tensor) for _ in range(world)]
    torch.distributed.all_gather(outs, tensor)
    return torch.cat(outs, dim=0)


@verify_operation
def gather(tensor):
    """Concatenate ``tensor`` (or every tensor of a nested structure) from all ranks along dim 0."""
    if PartialState().distributed_type == DistributedType.NO:
        return tensor
    return recursively_apply(_gather_one, tensor, error_on_other_type=True)


def gather_object(object: Any):
    """Gather picklable python lists from all ranks into one flat list."""
    state = PartialState()
    if state.distributed_type == DistributedType.NO:
        return object
    output = [None for _ in range(state.num_processes)]
    torch.distributed.all_gather_object(output, object)
    return [x for y in output for x in y]


def _broadcast_one(tensor, src=0):
    torch.distributed.broadcast(tensor, src=src)
    return tensor


@verify_operation
def broadcast(tensor, from_process: int = 0):
    """In-place broadcast of a (nested) tensor from ``from_process``."""
    if PartialState().distributed_type == DistributedType.NO:
        return tensor
    return recursively_apply(_broadcast_one, tensor, error_on_other_type=True, src=from_process)


def broadcast_object_list(object_list, from_process: int = 0):
    if PartialState().distributed_type != DistributedType.NO:
        torch.distributed.broadcast_object_list(object_list, src=from_process)
    return object_list


def copy_tensor_to_devices(tensor=None) -> torch.Tensor:
    """Send a tensor that exists on one rank only to every rank (used by pipeline inference)."""
    state = PartialState()
    if state.num_processes == 1:
        return tensor
    if tensor is not None:
        tensor_shape, tensor_dtype = tensor.shape, tensor.dtype
        flag = torch.tensor([state.process_index + 1], device=state.device)
    else:
        tensor_shape, tensor_dtype = None, None
        flag = torch.tensor([0], device=state.device)
    torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MAX)
    src = int(flag.item()) - 1
    meta = [tensor_shape, tensor_dtype]
    torch.distributed.broadcast_object_list(meta, src=src)
    if tensor is None:
        tensor = torch.empty(meta[0], dtype=meta[1], device=state.device)
    else:
        tensor = tensor.to(state.device)
    torch.distributed.broadcast(tensor, src=src)
    return tensor


# dtype codes for announcing a tensor that only one rank holds (shape and code travel in one integer vector)
TENSOR_TYPE_TO_INT = {dtype: code for code, dtype in enumerate(
    (torch.float, torch.double, torch.half, torch.bfloat16, torch.uint8, torch.int8, torch.int16, torch.int32, torch.int64, torch.bool), start=1)}
TENSOR_INT_TO_DTYPE = {code: dtype for dtype, code in TENSOR_TYPE_TO_INT.items()}


def gather_tensor_shape(tensor):
    """``(shape, dtype code)`` of a tensor that exists on one rank only, known to every rank afterwards: the holder writes its
    dimensions followed by the dtype code into a zero vector, a sum over ranks spreads it, the non-zero entries are the answer.
    (Dimensions of size zero cannot be told from padding — as in the reference, utils/operations.py:496-518.)"""
    state = PartialState()
    announced = torch.zeros(64, dtype=torch.int64, device=state.device)
    if tensor is not None:
        fields = [*tensor.shape, TENSOR_TYPE_TO_INT[tensor.dtype]]
        announced[: len(fields)] = torch.tensor(fields, dtype=torch.int64)
    announced = reduce(announced, reduction="sum")
    filled = announced[announced.nonzero()].flatten()
    return filled[:-1], int(filled[-1])


# ------------------------------------------------------------------------------- slicing / padding
def slice_tensors(data, tensor_slice, process_index=None, num_processes=None):
    return recursively_apply(lambda t: t[tensor_slice], data)


def concatenate(data, dim: int = 0):
    """``data`` is a list of identically shaped trees: one tree whose leaves are the concatenations along ``dim``."""
    head = data[0]
    if isinstance(head, (list, tuple)):
        return honor_type(head, (concatenate([tree[i] for tree in data], dim=dim) for i in range(len(head))))
    if isinstance(head, Mapping):
        return type(head)({key: concatenate([tree[key] for tree in data], dim=dim) for key in head})
    if not isinstance(head, torch.Tensor):
        raise TypeError(f"Can only concatenate tensors but got {type(head)}")
    return torch.cat(data, dim=dim)


class CannotPadNestedTensorWarning(UserWarning):
    pass


@chained_operation
def pad_across_processes(tensor, dim: int = 0, pad_index: int = 0, pad_first: bool = False):
    """Pad every rank's tensor along ``dim`` to the longest one so they can be gathered."""

    def _pad_one(t, dim=0, pad_index=0, pad_first=False):
        if getattr(t, "is_nested", False):
            warnings.warn("Cannot pad nested tensors without more information. Leaving unprocessed.", CannotPadNestedTensorWarning)
            return t
        if not -t.ndim <= dim < t.ndim:
            return t
        dim %= t.ndim
        # the whole shape goes round, not just `dim`: in debug mode a rank whose tensor has another number of dimensions is caught here
        longest = int(gather(torch.tensor(t.shape, device=t.device)[None])[:, dim].max())
        if longest == t.shape[dim]:
            return t
        shape = list(t.shape)
        shape[dim] = longest
        out = t.new_full(shape, pad_index)
        out.narrow(dim, longest - t.shape[dim] if pad_first else 0, t.shape[dim]).copy_(t)
        return out

    return recursively_apply(_pad_one, tensor, error_on_other_type=True, dim=dim, pad_index=pad_index, pad_first=pad_first)


def pad_input_tensors(tensor, batch_size: int, num_processes: int, dim: int = 0):
    """Pad the batch dim with copies of the last row so it divides evenly over ``num_processes``."""

    def _pad_one(t, batch_size, num_processes, dim=0):
        short_by = -batch_size % num_processes
        if short_by == 0:
            return t
        return torch.cat([t, t[-1:].expand(short_by, *t.shape[1:])], dim=0)

    return recursively_apply(_pad_one, tensor, error_on_other_type=True, batch_size=batch_size, num_processes=num_processes, dim=dim)


# ------------------------------------------------------------------------------- reduce
@verify_operation
def reduce(tensor, reduction: str = "mean", scale: float = 1.0):
    """Sum (or mean) a (nested) tensor over all ranks; returns new tensors."""

    def _reduce_one(t, reduction="mean", scale=1.0):
        state = PartialState()
        cloned = t.clone()
        if state.distributed_type == DistributedType.NO:
            return cloned
        if scale != 1.0 and cloned.is_floating_point():
            cloned *= scale
        torch.distributed.all_reduce(cloned, torch.distributed.ReduceOp.SUM)
        if reduction == "mean":
            cloned = cloned / state.num_processes
        return cloned

    return recursively_apply(_reduce_one, tensor, error_on_other_type=True, reduction=reduction, scale=scale)


# ------------------------------------------------------------------------------- precision of outputs
def convert_to_fp32(tensor):
    """fp16/bf16 leaves -> fp32 (other leaves untouched)."""
    def _is_low(t):
        return (is_torch_tensor(t) or hasattr(t, "dtype")) and getattr(t, "dtype", None) in (torch.float16, torch.bfloat16)

    return recursively_apply(lambda t: t.float(), tensor, test_type=_is_low)


class ConvertOutputsToFp32:
    """Picklable decorator: call ``model_forward`` and up-cast its floating outputs to fp32."""

    def __init__(self, model_forward):
        self.model_forward = model_forward
        update_wrapper(self, model_forward)

    def __call__(self, *args, **kwargs):
        return convert_to_fp32(self.model_forward(*args, **kwargs))

    def __getstate__(self):
        raise pickle.PicklingError(
            "Cannot pickle a prepared model with automatic mixed precision, please unwrap the model with "
            "`Accelerator.unwrap_model(model)` before pickling it."
        )


def convert_outputs_to_fp32(model_forward):
    model_forward = ConvertOutputsToFp32(model_forward)

    def forward(*args, **kwargs):
        return model_forward(*args, **kwargs)

    forward.__wrapped__ = model_forward
    return forward


@contextmanager
def GatheredParameters(params, modifier_rank=None, fwd_module=None, enabled=True):
    """Context in which sharded parameters are whole.  With the native FSDP engine the owning unit is
    un-sharded for the duration; without sharding it is a no-op (reference :852-871 defers to DeepSpeed)."""
    units = []
    if enabled:
        for p in params if isinstance(params, (list, tuple)) else [params]:
            unit = getattr(p, "_b200_fsdp_unit", None)
            if unit is not None and unit not in units:
                units.append(unit)
    for u in units:
        u.engine.unshard(u)
    try:
        yield
    finally:
        for u in units:
            u.engine.reshard(u, write_back=modifier_rank is not None)
