"""Model surgery for models that do not fit one device: where each tensor lives, how big things are, which device gets what,
and how a checkpoint is streamed in without ever holding the whole model twice.

The pieces, bottom up:

* sizes — ``convert_file_size_to_int``, ``dtype_byte_size``, ``compute_module_sizes`` (bytes under every dotted module name),
  ``get_max_layer_size`` (the largest block that must stay in one piece);
* placement — ``set_module_tensor_to_device`` replaces one parameter / buffer by a copy elsewhere (or by a checkpoint value) keeping
  its class, ``requires_grad`` and ties; ``align_module_device`` does it temporarily for a whole module;
* ties — ``find_tied_parameters`` / ``retie_parameters`` / ``check_tied_parameters_on_same_device``;
* budgets and maps — ``get_max_memory``, ``get_balanced_memory``, ``infer_auto_device_map`` (greedy packer), ``clean_device_map``,
  ``check_device_map``;
* checkpoints — ``load_state_dict`` (safetensors land directly on their target devices), ``shard_checkpoint``,
  ``load_checkpoint_in_model`` (gpu / cpu / disk per the device map), ``get_state_dict_offloaded_model``.

Behavioural parity: reference ``src/accelerate/utils/modeling.py`` (``set_module_tensor_to_device`` :217-424, ``named_module_tensors``
:427-474, ties :477-648, sizes :651-741, ``get_max_memory`` :744-852, ``clean_device_map`` :855-877, ``get_balanced_memory`` :918-1049,
``infer_auto_device_map`` :1278-1583, ``check_device_map`` :1586-1617, ``load_state_dict`` :1620-1712, ``load_checkpoint_in_model``
:1788-2046, autocast / scaler helpers :2049-2131).  Devices here are CUDA GPUs (integers), ``"cpu"`` and ``"disk"``.
"""

from __future__ import annotations

import contextlib
import gc
import itertools
import json
import logging
import os
import re
import shutil
import tempfile
import warnings
from collections import OrderedDict, defaultdict
from typing import Optional, Union

import torch
import torch.nn as nn

from .constants import SAFE_WEIGHTS_NAME, WEIGHTS_NAME
from .dataclasses import AutocastKwargs, CustomDtype, DistributedType
from .memory import clear_device_cache  # noqa: F401 - re-exported
from .offload import load_offloaded_weight, offload_weight, save_offload_index

logger = logging.getLogger(__name__)

WEIGHTS_INDEX_NAME = "pytorch_model.bin.index.json"
_META = torch.device("meta")
_QUANTIZED_PARAM_CLASSES = ("Int8Params", "FP4Params", "Params4bit")
_TENSOR_SUBCLASS_PARAMS = ("QTensor", "QBitsTensor", "AffineQuantizedTensor")


# ------------------------------------------------------------------------------------------------- sizes
_SIZE_PATTERN = re.compile(r"^\s*(\d+(?:\.\d+)?)\s*([KMG])(I?)([Bb])\s*$", re.IGNORECASE)


def convert_file_size_to_int(size: Union[int, str]) -> int:
    """``"5GB"`` → 5·10⁹, ``"2MiB"`` → 2·2²⁰; a lower-case ``b`` means bits (``"8Gb"`` is one gigabyte).  Integers pass through."""
    if isinstance(size, int):
        return size
    match = _SIZE_PATTERN.match(size)
    if match is None:
        raise ValueError("`size` is not in a valid format. Use an integer followed by the unit, e.g., '5GB'.")
    amount, prefix, binary, unit = match.groups()
    step = 1024 if binary else 1000
    scale = step ** ("KMG".index(prefix.upper()) + 1)
    total = int(float(amount) * scale) if "." in amount else int(amount) * scale
    return total // 8 if unit == "b" else total


_CUSTOM_DTYPE_BYTES = {CustomDtype.INT2: 0.25, CustomDtype.INT4: 0.5, CustomDtype.FP8: 1}


def dtype_byte_size(dtype) -> float:
    """Bytes per element; booleans count as one bit and the sub-byte ``CustomDtype`` s as what they are."""
    if dtype == torch.bool:
        return 1 / 8
    if dtype in _CUSTOM_DTYPE_BYTES:
        return _CUSTOM_DTYPE_BYTES[dtype]
    if isinstance(dtype, torch.dtype):
        info = torch.finfo(dtype) if (dtype.is_floating_point or dtype.is_complex) else torch.iinfo(dtype)
        return info.bits // 8 * (2 if dtype.is_complex else 1)
    digits = re.search(r"(\d+)(?:_\w+)?$", str(dtype))
    if digits is None:
        raise ValueError(f"`dtype` is not a valid dtype: {dtype}.")
    return int(digits.group(1)) // 8


def _is_integer_like(dtype: torch.dtype) -> bool:
    """Integer, unsigned and boolean tensors are never cast when a floating dtype is requested."""
    return not (dtype.is_floating_point or dtype.is_complex)


def _as_torch_dtype(dtype):
    return getattr(torch, dtype.replace("torch.", "")) if isinstance(dtype, str) else dtype


def id_tensor_storage(tensor: torch.Tensor):
    """``(device, address, bytes)`` of the storage behind ``tensor``: equal for views of one allocation, ``(dev, 0, 0)`` when the
    tensor has no accessible storage (meta, some subclasses)."""
    try:
        storage = tensor.untyped_storage()
        return tensor.device, storage.data_ptr(), storage.nbytes()
    except Exception:
        return tensor.device, 0, 0


def is_peft_model(model) -> bool:
    try:
        from peft import PeftModel
    except ImportError:
        return False
    return isinstance(model, PeftModel)


def check_device_same(first_device, second_device) -> bool:
    """Same physical device?  An accelerator device without an index means index 0 (``cuda`` is ``cuda:0``)."""

    def canonical(device):
        if device.type != "cpu" and device.index is None:
            return torch.device(device.type, 0)
        return device

    return first_device.type == second_device.type and canonical(first_device) == canonical(second_device)


def named_module_tensors(module: nn.Module, include_buffers: bool = True, recurse: bool = False, remove_non_persistent: bool = False):
    """``(name, tensor)`` over the parameters, then (optionally) the buffers — all of them or only those saved in a state dict."""
    yield from module.named_parameters(recurse=recurse)
    if not include_buffers:
        return
    transient = get_non_persistent_buffers(module, recurse=recurse) if remove_non_persistent else ()
    yield from ((name, buf) for name, buf in module.named_buffers(recurse=recurse) if name not in transient)


def get_non_persistent_buffers(module: nn.Module, recurse: bool = False, fqns: bool = False) -> set:
    """Names of the buffers ``register_buffer(..., persistent=False)`` created; ``fqns`` qualifies the names of sub-modules."""
    names = set(module._non_persistent_buffers_set)
    if recurse:
        for prefix, child in module.named_modules():
            names.update(f"{prefix}.{b}" if fqns and prefix else b for b in child._non_persistent_buffers_set)
    return names


def compute_module_sizes(model: nn.Module, dtype=None, special_dtypes: Optional[dict] = None, buffers_only: bool = False):
    """``{dotted name: bytes}`` for every tensor, every module that contains one, and ``""`` (the whole model).

    ``dtype`` is the dtype the floating tensors will be loaded in (a tensor never grows: the smaller of the two counts);
    ``special_dtypes`` overrides it per tensor name.
    """
    target_bytes = None if dtype is None else dtype_byte_size(_as_torch_dtype(dtype))
    override_bytes = {name: dtype_byte_size(_as_torch_dtype(d)) for name, d in (special_dtypes or {}).items()}

    def bytes_per_element(name, tensor):
        own = dtype_byte_size(tensor.dtype)
        if name in override_bytes:
            return override_bytes[name]
        if target_bytes is None or _is_integer_like(tensor.dtype):
            return own
        return min(own, target_bytes)

    sizes = defaultdict(int)
    tensors = model.named_buffers(recurse=True) if buffers_only else named_module_tensors(model, recurse=True)
    for name, tensor in tensors:
        nbytes = tensor.numel() * bytes_per_element(name, tensor)
        sizes[""] += nbytes
        for prefix in itertools.accumulate(name.split("."), lambda a, b: f"{a}.{b}"):
            sizes[prefix] += nbytes
    return sizes


def compute_module_total_buffer_size(model: nn.Module, dtype=None, special_dtypes=None):
    return compute_module_sizes(model, dtype=dtype, special_dtypes=special_dtypes, buffers_only=True).get("", 0)


def _top_level_entries(model: nn.Module) -> list:
    """What a packer starts from: the model's own parameters, its children, its own buffers — in that order."""
    return [*model.named_parameters(recurse=False), *model.named_children(), *model.named_buffers(recurse=False)]


def get_max_layer_size(modules: list, module_sizes: dict, no_split_module_classes: list):
    """``(bytes, [names])`` of the largest piece that cannot be divided: a tensor, a module without children, or a module whose
    class is in ``no_split_module_classes``.  ``modules`` is a list of ``(name, module or tensor)``; visiting is depth first, in order."""
    best, holders = 0, []
    stack = list(reversed(modules))
    while stack:
        name, entry = stack.pop()
        children = list(entry.named_children()) if isinstance(entry, nn.Module) else []
        if children and type(entry).__name__ not in no_split_module_classes:
            stack.extend((f"{name}.{child_name}", child) for child_name, child in reversed(children))
            continue
        size = module_sizes[name]
        if size > best:
            best, holders = size, [name]
        elif size == best:
            holders.append(name)
    return best, holders


def calculate_maximum_sizes(model: nn.Module):
    """``(total bytes, (largest undividable layer bytes, its names))`` — what ``accelerate estimate-memory`` prints."""
    sizes = compute_module_sizes(model)
    keep_whole = getattr(model, "_no_split_modules", None) or []
    return sizes[""], get_max_layer_size(_top_level_entries(model), sizes, keep_whole)


def _holds(owner: str, tensor_name: str) -> bool:
    """Whether ``tensor_name`` lives at or below ``owner`` ("1" holds "1.weight" but not "10.weight")."""
    return tensor_name == owner or tensor_name.startswith(owner + ".")


def _entries_below(name: str, entry) -> list:
    """``entry`` taken apart one level: its own parameters, then its children, under their full names."""
    if not isinstance(entry, nn.Module):
        return []
    return [(f"{name}.{n}", v) for n, v in [*entry.named_parameters(recurse=False), *entry.named_children()]]


def get_module_size_with_ties(tied_params: list, module_size: int, module_sizes: dict, modules_to_treat: list):
    """Bytes needed to place a module *together with* the entries of ``modules_to_treat`` that hold the parameters tied to it
    (``tied_params``: names outside the module).  A tied parameter shares storage with one already counted in ``module_size``, so
    each is charged its holder minus itself.  Returns ``(bytes, holder names, holders)``, holders in first-seen order
    (reference utils/modeling.py:1134-1167)."""
    names, holders, total = [], [], module_size
    for tied in tied_params:
        holder = next(((n, m) for n, m in modules_to_treat if _holds(n, tied)), None)
        if holder is None:
            raise ValueError(f"tied parameter {tied!r} is not below any of the entries still to be placed")
        if holder[0] not in names:
            names.append(holder[0])
            holders.append(holder[1])
            total += module_sizes[holder[0]]
        total -= module_sizes[tied]
    return total, names, holders


def _open_ancestors(modules: list, name: str) -> list:
    """``modules`` with every ancestor of ``name`` taken apart, outermost first, so that ``name`` is an entry of its own."""
    opened = list(modules)
    steps = name.split(".")
    for depth in range(1, len(steps)):
        ancestor = ".".join(steps[:depth])
        at = next((i for i, (n, _) in enumerate(opened) if n == ancestor), None)
        if at is not None:
            opened[at : at + 1] = _entries_below(*opened[at])
    return opened


def fallback_allocate(modules: list, module_sizes: dict, size_limit, no_split_module_classes: Optional[list] = None,
                      tied_parameters: Optional[list] = None):
    """The first piece of ``modules`` that still fits ``size_limit`` — used by the device-map packer to fill the room a device has
    left when the module whose turn it is does not fit and cannot be divided.

    ``modules`` is searched in order; an entry that is too big is taken apart (unless its class is in ``no_split_module_classes``)
    and its pieces are searched before the entries after it.  "Fits" counts the entries that would have to come along because a
    parameter inside the piece is tied to one inside them.  Returns ``(name, piece, rest)``; ``rest`` is ``modules`` with the piece's
    ancestors taken apart and the piece removed.  ``(None, None, modules)`` when nothing fits (reference utils/modeling.py:1170-1275).
    """
    try:
        limit = convert_file_size_to_int(size_limit)
    except ValueError:
        return None, None, modules
    keep_whole = set(no_split_module_classes or [])
    groups = tied_parameters or []

    def needed(name: str) -> int:
        crossing = [g for g in groups if any(_holds(name, p) for p in g) and not all(_holds(name, p) for p in g)]
        outside = [p for g in crossing for p in g if not _holds(name, p)]
        if not outside:
            return module_sizes[name]
        others = [(n, m) for n, m in _open_ancestors(modules, name) if n != name]
        try:
            return get_module_size_with_ties(outside, module_sizes[name], module_sizes, others)[0]
        except ValueError:  # the tied partner was placed already: nothing more to bring
            return module_sizes[name]

    queue = list(modules)
    while queue:
        name, piece = queue.pop(0)
        if needed(name) <= limit:
            rest = [(n, m) for n, m in _open_ancestors(modules, name) if n != name]
            return name, piece, rest
        if isinstance(piece, nn.Module) and type(piece).__name__ not in keep_whole and any(True for _ in piece.children()):
            queue = _entries_below(name, piece) + queue
    return None, None, modules


# ------------------------------------------------------------------------------------------------- tensor placement
def _owner_of(module: nn.Module, dotted: str) -> tuple[nn.Module, str]:
    """The sub-module that directly holds ``dotted`` and the tensor's own name."""
    *path, leaf = dotted.split(".")
    for step in path:
        child = getattr(module, step)
        if child is None:
            raise ValueError(f"{module} has no attribute {step}.")
        module = child
    return module, leaf


def _tied_copy(tied_params_map, tensor, device):
    """The already materialised copy on ``device`` of a tensor registered in ``tied_params_map`` (else None)."""
    if tied_params_map is None or tensor is None:
        return None
    return tied_params_map.get(tensor.data_ptr(), {}).get(device)


def _rewrap_parameter(template: nn.Parameter, data: torch.Tensor, device, requires_grad: bool, non_blocking: bool):
    """A parameter of ``template``'s class around ``data`` on ``device`` (bitsandbytes / quanto / torchao classes need their own treatment)."""
    kind = type(template)
    requires_grad = requires_grad and (data.is_floating_point() or data.is_complex())  # integer storage (pre-quantised weights) cannot
    if kind.__name__ in _QUANTIZED_PARAM_CLASSES:
        if kind.__name__ == "Int8Params" and data.dtype == torch.float32:
            data = data.to(torch.float16, non_blocking=non_blocking)  # int8 quantisation starts from half precision
        return kind(data, requires_grad=requires_grad, **template.__dict__).to(device)
    if kind.__name__ in _TENSOR_SUBCLASS_PARAMS:
        return nn.Parameter(data, requires_grad=requires_grad).to(device)
    return kind(data, requires_grad=requires_grad).to(device, non_blocking=non_blocking)


def set_module_tensor_to_device(module: nn.Module, tensor_name: str, device, value: Optional[torch.Tensor] = None, dtype=None,
                                fp16_statistics=None, tied_params_map=None, non_blocking: bool = False, clear_cache: bool = True):
    """Put the parameter or buffer ``tensor_name`` (dotted names allowed) of ``module`` on ``device``, optionally with a new ``value``.

    Leaving the meta device needs a ``value``.  The tensor keeps its kind (parameter class, ``requires_grad``, buffer-ness) and, when
    ``dtype`` is None, its dtype; integer tensors are never cast.  ``tied_params_map`` ``{data_ptr: {device: tensor}}`` makes every
    member of a tie group resolve to one tensor per device instead of one copy each.
    """
    owner, leaf = _owner_of(module, tensor_name)
    is_parameter, is_buffer = leaf in owner._parameters, leaf in owner._buffers
    if not (is_parameter or is_buffer):
        raise ValueError(f"{owner} does not have a parameter or a buffer named {leaf}.")
    current = getattr(owner, leaf)

    shared = _tied_copy(tied_params_map, value, device)
    if shared is None:
        shared = _tied_copy(tied_params_map, current, device)
    if shared is not None:
        owner._parameters[leaf] = shared
        return

    to_meta = device in ("meta", _META)
    if current.device == _META and value is None and not to_meta:
        raise ValueError(f"{leaf} is on the meta device, we need a `value` to put in on {device}.")
    template = owner._parameters[leaf] if is_parameter else None
    if value is not None:
        if value.shape != current.shape and type(template).__name__ != "Params4bit":
            raise ValueError(f'Trying to set a tensor of shape {value.shape} in "{leaf}" (which has shape {current.shape}), this looks incorrect.')
        if dtype is None:
            value = value.to(current.dtype, non_blocking=non_blocking)
        elif not _is_integer_like(value.dtype):
            value = value.to(dtype, non_blocking=non_blocking)
    requested = device  # the key callers use in `tied_params_map`
    if isinstance(device, int):
        device = torch.device("cuda", device)

    with torch.no_grad():
        if value is None:
            placed = current.to(device, non_blocking=non_blocking)
            if to_meta and dtype is not None:
                if not _is_integer_like(current.dtype):
                    placed = placed.to(dtype, non_blocking=non_blocking)
                if is_parameter:
                    owner._parameters[leaf] = type(template)(placed, requires_grad=current.requires_grad)
        elif isinstance(value, torch.Tensor):
            placed = value.to(device, non_blocking=non_blocking)
        else:
            placed = torch.tensor(value, device=device)

        if is_buffer:
            owner._buffers[leaf] = placed
        elif value is not None or not check_device_same(torch.device(device), owner._parameters[leaf].device):
            placed = _rewrap_parameter(owner._parameters[leaf], placed, device, current.requires_grad, non_blocking)
            owner._parameters[leaf] = placed
            if fp16_statistics is not None:
                placed.SCB = fp16_statistics.to(device, non_blocking=non_blocking)

    if tied_params_map is not None:
        for source in (current, value):
            copies = tied_params_map.get(source.data_ptr()) if source is not None else None
            if copies is not None and requested not in copies:
                copies[requested] = placed
                break


def has_offloaded_params(module: nn.Module) -> bool:
    """Is ``module`` run through an ``AlignDevicesHook`` that keeps its weights somewhere else between forwards?"""
    from ..hooks import AlignDevicesHook

    hook = getattr(module, "_hf_hook", None)
    return isinstance(hook, AlignDevicesHook) and hook.offload


@contextlib.contextmanager
def align_module_device(module: nn.Module, execution_device=None):
    """Within the block, ``module``'s own parameters are on ``execution_device`` (offloaded ones are brought in through their hook);
    afterwards everything is back where it was."""
    if has_offloaded_params(module):
        hook = module._hf_hook
        saved = hook.execution_device
        if execution_device is not None:
            hook.execution_device = execution_device
        try:
            hook.pre_forward(module)
            yield
        finally:
            hook.post_forward(module, None)
            hook.execution_device = saved
        return
    if execution_device is None:
        yield
        return
    home = {name: param.device for name, param in module.named_parameters(recurse=False)}
    try:
        for name in home:
            set_module_tensor_to_device(module, name, execution_device)
        yield
    finally:
        for name, device in home.items():
            set_module_tensor_to_device(module, name, device)


# ------------------------------------------------------------------------------------------------- tied parameters
class FindTiedParametersResult(list):
    """List of tie groups (each a sorted list of names).  ``values()`` is what the old dict-shaped result offered."""

    def values(self):
        warnings.warn("The 'values' method of FindTiedParametersResult is deprecated and will be removed in Accelerate v1.3.0. ", FutureWarning)
        return [name for group in self for name in group[1:]]


def find_tied_parameters(model: nn.Module, **kwargs) -> list:
    """Groups of parameter names bound to one and the same ``nn.Parameter`` object, e.g. ``[["embed.weight", "lm_head.weight"]]``."""
    names_of = defaultdict(list)
    for name, param in model.named_parameters(remove_duplicate=False):
        names_of[id(param)].append(name)
    return FindTiedParametersResult(sorted(set(names)) for names in names_of.values() if len(names) > 1)


def check_tied_parameters_in_config(model: nn.Module) -> bool:
    """Does the model's (transformers-style) config promise tied weights — so that finding none means ``tie_weights()`` was forgotten?"""
    config = getattr(model, "config", None)
    if not any(cls.__name__ == "PreTrainedModel" for cls in type(model).__mro__):
        return bool(getattr(config, "tie_word_embeddings", False))
    promises = [any(hasattr(sub, "_tie_weights") for sub in model.modules())]
    if config is not None:
        try:
            text_config = config.get_text_config(decoder=True)
        except Exception:
            text_config = config
        promises.append(bool(getattr(text_config, "tie_word_embeddings", False) and model.get_output_embeddings()))
        promises.append(bool(getattr(config, "is_encoder_decoder", False) and getattr(config, "tie_encoder_decoder", False)))
    return any(promises)


def _get_param_device(param: str, device_map: dict):
    """Device of the closest enclosing entry of ``device_map`` (``""`` encloses everything)."""
    name = param
    while name and name not in device_map:
        name = name.rpartition(".")[0]
    if name not in device_map:
        raise ValueError(f"The `device_map` does not contain the module {param}.")
    return device_map[name]


def check_tied_parameters_on_same_device(tied_params, device_map) -> None:
    """Warn about tie groups the device map tears apart: each member would get its own copy."""
    for group in tied_params:
        placement = {name: _get_param_device(name, device_map) for name in group}
        if len(set(placement.values())) > 1:
            logger.warning(f"Tied parameters are on different devices: {placement}. Please modify your custom device map or set `device_map='auto'`. ")


def retie_parameters(model: nn.Module, tied_params) -> None:
    """Loading gives every name its own tensor; make each group share the first member that holds real data again."""
    for group in tied_params:
        holders = [_owner_of(model, name) for name in group]
        keeper = next((getattr(owner, leaf) for owner, leaf in holders if getattr(owner, leaf).device != _META), None)
        if keeper is None:
            continue
        for owner, leaf in holders:
            setattr(owner, leaf, keeper)


# ------------------------------------------------------------------------------------------------- budgets
def get_max_memory(max_memory: Optional[dict] = None) -> dict:
    """Byte budget per device.  Not given: what is free on every visible GPU plus the available host RAM.  Given: sizes parsed
    (``"10GiB"``), keys validated and ordered GPUs first, then ``cpu``, then ``disk``."""
    gpu_count = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if max_memory is None:
        import psutil

        budget = {}
        for index in range(gpu_count):
            try:
                torch.zeros(1, device=index)  # forces the context; a dead device raises here
                budget[index] = torch.cuda.mem_get_info(index)[0]
            except Exception:
                logger.info(f"Device {index} seems unavailable, Proceeding to check subsequent devices.")
        budget["cpu"] = psutil.virtual_memory().available
        return budget
    parsed = {key: convert_file_size_to_int(amount) if isinstance(amount, str) else amount for key, amount in max_memory.items()}
    max_memory.update(parsed)  # callers see their strings parsed, as with the reference
    gpus = sorted(key for key in parsed if isinstance(key, int))
    named = [key for key in ("mps", "cpu", "disk") if key in parsed]
    strangers = [key for key in parsed if key not in gpus and key not in named]
    if strangers:
        raise ValueError(f"Device {strangers[0]} is not recognized, available devices are integers(for GPU/XPU), 'mps', 'cpu' and 'disk'")
    for index in gpus:
        if not 0 <= index < gpu_count:
            logger.warning(f"Device {index} is not available, available devices are {list(range(gpu_count))}")
    return {key: parsed[key] for key in gpus + named}


def get_module_leaves(module_sizes) -> list:
    """Names in ``module_sizes`` nothing else in it sits under, weights and biases themselves excluded."""
    parents = {name.rsplit(".", 1)[0] for name in module_sizes if "." in name}
    return [name for name in module_sizes if name not in parents and not name.endswith(("weight", "bias"))]


def _first_size_per_class(model: nn.Module, module_sizes: dict, class_names) -> dict:
    """Bytes of the first module of each of ``class_names`` met in ``module_sizes`` order."""
    wanted, found = set(class_names), {}
    for name, size in module_sizes.items():
        if name == "" or len(found) == len(wanted):
            continue
        try:
            kind = type(model.get_submodule(name)).__name__
        except AttributeError:  # a tensor's name, not a module's
            continue
        if kind in wanted:
            found.setdefault(kind, size)
    return found


def get_balanced_memory(model: nn.Module, max_memory: Optional[dict] = None, no_split_module_classes: Optional[list] = None,
                        dtype=None, special_dtypes: Optional[dict] = None, low_zero: bool = False) -> dict:
    """Budgets that make the greedy packer spread ``model`` evenly over the GPUs instead of filling GPU 0 first: every GPU but the
    last is capped at an equal share plus slack (1.25 x the larger of the biggest undividable block and the mean leaf module).
    ``low_zero`` keeps GPU 0 as empty as the others allow (room for ``generate``'s activations)."""
    defaults_used = max_memory is None
    max_memory = get_max_memory(max_memory)
    gpus = sorted(key for key, room in max_memory.items() if isinstance(key, int) and room > 0) if torch.cuda.is_available() else []
    if not gpus:
        return max_memory
    if len(gpus) == 1:
        low_zero = False
        if defaults_used:
            max_memory[gpus[0]] *= 0.9
            logger.info(f"We will use 90% of the memory on device {gpus[0]} for storing the model, and 10% for the buffer to avoid OOM. "
                        "You can set `max_memory` in to a higher value to use more memory (at your own risk).")
    sizes = compute_module_sizes(model, dtype=dtype, special_dtypes=special_dtypes)
    share = sizes[""] // (len(gpus) - 1 if low_zero else len(gpus))

    if no_split_module_classes is None:
        no_split_module_classes = []
    elif not isinstance(no_split_module_classes, (list, tuple)):
        no_split_module_classes = [no_split_module_classes]
    undividable = max(_first_size_per_class(model, sizes, no_split_module_classes).values(), default=0)
    # mean size of the smallest *modules* (drop the tensor-level entries first, then take what is a leaf among the rest)
    tensor_level = set(get_module_leaves(sizes))
    module_level = {name: size for name, size in sizes.items() if name not in tensor_level}
    leaves = get_module_leaves(module_level)
    mean_leaf = int(sum(module_level[name] for name in leaves) / max(len(leaves), 1))
    share += int(1.25 * max(undividable, mean_leaf))

    for index in gpus[:-1]:
        cap = max_memory[0] if (low_zero and index == 0) else share
        max_memory[index] = min(cap, max_memory[index])
    if low_zero:
        overflow = sizes[""] - sum(max_memory[i] for i in range(1, len(gpus)))
        max_memory[0] = min(max(0, overflow), max_memory[0])
    return max_memory


def clean_device_map(device_map: dict, module_name: str = "") -> dict:
    """Replace the entries below a module by one entry for the module wherever they all name the same device; works on
    ``device_map`` in place, deepest modules first, so whole sub-trees fold up as far as they can."""
    root = module_name

    def below(prefix):
        dotted = f"{prefix}." if prefix else ""
        return [key for key in device_map if key.startswith(dotted)]

    ancestors = set()
    for key in below(root):
        parts = key.split(".")
        ancestors.update(".".join(parts[:depth]) for depth in range(len(root.split(".")) if root else 0, len(parts)))
    ancestors.add(root)
    for prefix in sorted(ancestors, key=lambda name: (-(name.count(".") + 1 if name else 0), name)):
        members = below(prefix)
        places = {device_map[key] for key in members}
        if len(members) > 1 and len(places) == 1:
            for key in members:
                del device_map[key]
            device_map[prefix] = places.pop()
    return device_map


# ------------------------------------------------------------------------------------------------- device map inference
def infer_auto_device_map(model: nn.Module, max_memory: Optional[dict] = None, no_split_module_classes: Optional[list] = None,
                          dtype=None, special_dtypes: Optional[dict] = None, verbose: bool = False, clean_result: bool = True,
                          offload_buffers: bool = False, fallback_allocation: bool = False):
    """Greedy placement of ``model``'s modules on GPU0..N, then cpu, then disk.

    Modules are visited in definition order.  A module that does not fit the current device is split
    into its children (unless its class is in ``no_split_module_classes``); if it cannot be split,
    the packer moves on to the next device — or, with ``fallback_allocation``, first looks for a
    later module small enough to use the remaining room.  On every accelerator except the last one
    room is reserved for the largest un-splittable layer, so weights offloaded to cpu/disk can be
    brought in one layer at a time.  Groups of tied parameters are kept on one device.
    """
    max_memory = get_max_memory(max_memory)
    if no_split_module_classes is None:
        no_split_module_classes = []
    elif not isinstance(no_split_module_classes, (list, tuple)):
        no_split_module_classes = [no_split_module_classes]
    devices = list(max_memory.keys())
    if "disk" not in devices:
        devices.append("disk")
    gpus = [d for d in devices if d not in ("cpu", "disk")]
    main_devices = [gpus[0], "cpu"] if len(gpus) > 0 else ["cpu"]
    module_sizes = compute_module_sizes(model, dtype=dtype, special_dtypes=special_dtypes)
    tied_parameters = find_tied_parameters(model)
    if check_tied_parameters_in_config(model) and len(tied_parameters) == 0:
        logger.warn("The model weights are not tied. Please use the `tie_weights` method before using the `infer_auto_device` function.")
    modules_to_treat = _top_level_entries(model)
    max_layer_size, max_layer_names = get_max_layer_size(modules_to_treat, module_sizes, no_split_module_classes)

    device_map: "OrderedDict[str, Union[int, str]]" = OrderedDict()
    used = {d: 0 for d in devices}
    buffers_on_device = {d: 0 for d in devices}
    dev_idx = 0

    def budget(idx: int) -> Optional[int]:
        d = devices[idx]
        if d == "disk":
            return None
        b = max_memory[d]
        if d in main_devices and idx < len(devices) - 1:
            # leave room to page the biggest layer in
            b = b - max_layer_size
        return b

    def children_of(name, module):
        if not isinstance(module, nn.Module) or len(list(module.named_children())) == 0:
            return []  # tensors and leaf modules are placed whole
        kids = list(module.named_parameters(recurse=False)) + list(module.named_children())
        return [(f"{name}.{n}", v) for n, v in kids]

    def tied_group_of(name):
        """Names (other than ``name``'s own) of parameters tied to something inside ``name``."""
        others = []
        for group in tied_parameters:
            inside = [p for p in group if p == name or p.startswith(name + ".")]
            if inside and len(inside) < len(group):
                others += [p for p in group if p not in inside]
        return others

    while len(modules_to_treat) > 0:
        name, module = modules_to_treat.pop(0)
        if verbose:
            print(f"\nTreating module {name}.")
        module_size = module_sizes[name]
        # tied partners not yet placed must come along
        partners = [p for p in tied_group_of(name) if not any(p == k or p.startswith(k + ".") for k in device_map)]
        placeable = [p for p in partners if any(_holds(n, p) for n, _ in modules_to_treat)]
        total_size, partner_names, partner_entries = get_module_size_with_ties(placeable, module_size, module_sizes, modules_to_treat)
        partner_modules = list(zip(partner_names, partner_entries))
        device = devices[dev_idx]
        cap = budget(dev_idx)
        if cap is None or used[device] + total_size <= cap:
            device_map[name] = device
            used[device] += total_size
            for n, m in partner_modules:
                device_map[n] = device
                modules_to_treat.remove((n, m))
            if isinstance(module, nn.Module):
                buffers_on_device[device] += compute_module_total_buffer_size(module, dtype=dtype, special_dtypes=special_dtypes)
            if verbose:
                print(f"Putting {name} (size={module_size}) on {device}.")
            continue
        # does not fit: split if allowed
        kids = children_of(name, module)
        if partner_modules and not kids:
            # try splitting a partner instead
            split_done = False
            for n, m in partner_modules:
                pk = children_of(n, m)
                if pk and m.__class__.__name__ not in no_split_module_classes:
                    i = modules_to_treat.index((n, m))
                    modules_to_treat = modules_to_treat[:i] + pk + modules_to_treat[i + 1:]
                    modules_to_treat.insert(0, (name, module))
                    split_done = True
                    break
            if split_done:
                continue
        if kids and module.__class__.__name__ not in no_split_module_classes:
            if verbose:
                print(f"Splitting {name}.")
            modules_to_treat = kids + modules_to_treat
            max_layer_size, max_layer_names = get_max_layer_size(
                [(n, m) for n, m in modules_to_treat if isinstance(m, torch.nn.Module)], module_sizes, no_split_module_classes)
            continue
        # cannot split: optionally fill the room this device has left with a later piece that fits, then come back to this module
        if fallback_allocation and cap is not None and cap - used[device] > 0:
            pick_name, pick, rest = fallback_allocate(modules_to_treat, module_sizes, cap - used[device], no_split_module_classes,
                                                      tied_parameters)
            if pick is not None:
                modules_to_treat = [(pick_name, pick), (name, module)] + rest
                continue
        if verbose:
            print(f"Not enough space on {device} to put {name} (space available {None if cap is None else cap - used[device]}, module size {total_size}).")
        if dev_idx >= len(devices) - 1:
            raise ValueError(f"The model does not fit the given `max_memory`: {name} ({total_size} bytes) has nowhere to go.")
        dev_idx += 1
        modules_to_treat.insert(0, (name, module))

    # offloaded modules keep their buffers on the main GPU unless told otherwise: check they fit
    if clean_result:
        device_map = clean_device_map(device_map)
    non_gpu_buffer_size = buffers_on_device.get("cpu", 0) + buffers_on_device.get("disk", 0)
    if non_gpu_buffer_size > 0 and not offload_buffers and len(gpus) > 0:
        is_buffer_fit_any_gpu = any(max_memory[g] - used[g] >= non_gpu_buffer_size for g in gpus)
        if len(gpus) > 0 and not is_buffer_fit_any_gpu:
            warnings.warn(
                f"Current model requires {non_gpu_buffer_size} bytes of buffer for offloaded layers, which seems does "
                f"not fit any GPU's remaining memory. If you are experiencing a OOM later, please consider using offload_buffers=True."
            )
    if verbose:
        print({d: used[d] for d in used if used[d]})
    return device_map


def check_device_map(model: nn.Module, device_map: dict) -> None:
    """Raise when some tensor of the state dict has no entry above it in ``device_map``; warn about entries that name nothing."""
    known = set(dict(model.named_modules())) | {name for name, _ in named_module_tensors(model, recurse=True)}
    strangers = [key for key in device_map if key != "" and key not in known]
    if strangers:
        warnings.warn(f"The following device_map keys do not match any submodules in the model: {strangers}", UserWarning)
    if "" in device_map:
        return

    def covered(tensor_name):
        return any(tensor_name == key or tensor_name.startswith(f"{key}.") for key in device_map)

    orphans = [name for name in model.state_dict() if not covered(name)]
    if orphans:
        raise ValueError(f"The device_map provided does not give any device for the following parameters: {', '.join(orphans)}")


# ------------------------------------------------------------------------------------------------- checkpoint IO
def _torch_device_name(device) -> str:
    return f"cuda:{device}" if isinstance(device, int) else device


def _require_pt_safetensors(checkpoint_file, metadata) -> None:
    if metadata is None:
        logger.warning(f"The safetensors archive passed at {checkpoint_file} does not contain metadata. Make sure to save your model "
                       "with the `save_pretrained` method. Defaulting to 'pt' metadata.")
        return
    fmt = metadata.get("format")
    if fmt not in ("pt", "tf", "flax"):
        raise OSError(f"The safetensors archive passed at {checkpoint_file} does not contain the valid metadata. Make sure you save "
                      "your model with the `save_pretrained` method.")
    if fmt != "pt":
        raise ValueError(f"The checkpoint passed was saved with {fmt}, we need a the pt format.")


def load_state_dict(checkpoint_file, device_map=None):
    """Read one checkpoint file.  ``torch.save`` files come back on CPU.  Safetensors files are read tensor by tensor onto the device
    ``device_map`` assigns to each (``disk`` and unassigned weights go to CPU), so GPU weights never make a stop in host memory."""
    if not str(checkpoint_file).endswith(".safetensors"):
        return torch.load(checkpoint_file, map_location=torch.device("cpu"), weights_only=True)
    from safetensors import safe_open
    from safetensors.torch import load_file

    with safe_open(checkpoint_file, framework="pt") as archive:
        _require_pt_safetensors(checkpoint_file, archive.metadata())
        stored = list(archive.keys())
    if device_map is None:
        return load_file(checkpoint_file)
    targets = set(device_map.values())
    if len(targets) == 1:
        (only,) = targets
        return load_file(checkpoint_file, device=_torch_device_name("cpu" if only == "disk" else only))

    def destination(weight_name):
        for entry, device in device_map.items():
            if weight_name == entry or weight_name.startswith(f"{entry}."):
                return "cpu" if device == "disk" else device
        return "cpu"

    per_device = defaultdict(list)
    for weight_name in stored:
        per_device[destination(weight_name)].append(weight_name)
    tensors = {}
    for device, names in per_device.items():
        with safe_open(checkpoint_file, framework="pt", device=_torch_device_name(device)) as archive:
            tensors.update((name, archive.get_tensor(name)) for name in names)
    return tensors


def load_offloaded_weights(model, index, offload_folder) -> None:
    """Bring the weights listed in an offload ``index`` (``offload_weight`` files in ``offload_folder``) back into ``model`` on CPU;
    int8 weights pick up their ``SCB`` statistics on the way."""
    for name, entry in (index or {}).items():
        if "SCB" in name:
            continue
        statistics = None
        companion = name.replace("weight", "SCB")
        if "weight" in name and companion in index:
            statistics = load_offloaded_weight(os.path.join(offload_folder, f"{companion}.dat"), index[companion])
        weight = load_offloaded_weight(os.path.join(offload_folder, f"{name}.dat"), entry)
        set_module_tensor_to_device(model, name, "cpu", value=weight, fp16_statistics=statistics)


def get_state_dict_offloaded_model(model: nn.Module) -> dict:
    """CPU state dict of a model whose weights live behind offload hooks, gathered one sub-module at a time."""
    gathered, still_meta = {}, set()
    for prefix, sub in model.named_modules():
        if prefix == "":
            continue
        try:
            with align_module_device(sub, "cpu"):
                local = sub.state_dict()
        except MemoryError:
            raise MemoryError("Offloaded module must fit in CPU memory to call save_model!") from None
        for key, tensor in local.items():
            full = f"{prefix}.{key}"
            if tensor.device == _META:
                still_meta.add(full)
            else:
                gathered[full] = tensor.to("cpu")
    still_meta -= set(gathered)
    if still_meta:
        logger.warning(f"The following tensors were not saved because they were still on meta device: {still_meta}")
    return gathered


def get_state_dict_from_offload(module: nn.Module, module_name: str, state_dict: dict, device_to_put_offload="cpu") -> dict:
    """Fill the entries of ``state_dict`` that belong to the offloaded ``module`` (``module_name`` is one of its tensors' names)."""
    prefix = module_name[: module_name.rfind(".")]
    with align_module_device(module, device_to_put_offload):
        for key, tensor in module.state_dict().items():
            if f"{prefix}.{key}" in state_dict:
                state_dict[f"{prefix}.{key}"] = tensor
    return state_dict


def shard_checkpoint(state_dict: dict, max_shard_size: Union[int, str] = "10GB", weights_name: str = WEIGHTS_NAME):
    """Cut ``state_dict`` into files of at most ``max_shard_size`` (one oversized tensor still gets a file of its own).  Tensors that
    share storage stay in one file.  Returns ``({file name: tensors}, index)``; a single file keeps ``weights_name`` and has no index,
    several are named ``stem-0000i-of-0000n.ext`` and the index maps every tensor to its file."""
    limit = convert_file_size_to_int(max_shard_size)
    shards: list = [{}]
    filled = total = 0
    home_of_storage = {}
    for key, tensor in state_dict.items():
        if isinstance(tensor, str):
            continue
        storage = id_tensor_storage(tensor)
        if storage[1] != 0 and storage in home_of_storage:
            shards[home_of_storage[storage]][key] = tensor
            continue
        nbytes = tensor.numel() * dtype_byte_size(tensor.dtype)
        if shards[-1] and filled + nbytes > limit:
            shards.append({})
            filled = 0
        shards[-1][key] = tensor
        home_of_storage[storage] = len(shards) - 1
        filled += nbytes
        total += nbytes
    if len(shards) == 1:
        return {weights_name: shards[0]}, None
    stem, extension = os.path.splitext(weights_name)
    files, weight_map = {}, {}
    for number, shard in enumerate(shards, start=1):
        file_name = f"{stem}-{number:05d}-of-{len(shards):05d}{extension}"
        files[file_name] = shard
        weight_map.update(dict.fromkeys(shard, file_name))
    return files, {"metadata": {"total_size": int(total)}, "weight_map": weight_map}


def _checkpoint_files(checkpoint) -> list:
    """Files to read for ``checkpoint``: a weights file, an index (``*.index.json``), or a folder holding either."""
    index_file = None
    if os.path.isfile(checkpoint):
        if not str(checkpoint).endswith(".json"):
            return [checkpoint]
        index_file = checkpoint
    elif os.path.isdir(checkpoint):
        present = os.listdir(checkpoint)
        for whole in (WEIGHTS_NAME, SAFE_WEIGHTS_NAME):
            if whole in present:
                return [os.path.join(checkpoint, whole)]
        indexes = [name for name in present if name.endswith(".index.json")]
        if not indexes:
            raise ValueError(f"{checkpoint} is not a folder containing a `.index.json` file or a {WEIGHTS_NAME} or a {SAFE_WEIGHTS_NAME} file")
        if len(indexes) > 1:
            raise ValueError(f"{checkpoint} containing more than one `.index.json` file, delete the irrelevant ones.")
        index_file = os.path.join(checkpoint, indexes[0])
    else:
        raise ValueError("`checkpoint` should be the path to a file containing a whole state dict, or the index of a sharded checkpoint, "
                         f"or a folder containing a sharded checkpoint or the whole state dict, but got {checkpoint}.")
    with open(index_file) as handle:
        index = json.load(handle)
    files = sorted(set(index.get("weight_map", index).values()))
    return [os.path.join(os.path.dirname(index_file), name) for name in files]


def _device_entry_for(tensor_name: str, device_map: dict):
    try:
        return _get_param_device(tensor_name, device_map)
    except ValueError:
        raise ValueError(f"{tensor_name} doesn't have any device set.") from None


def _stays_fp32(param_name: str, keep_in_fp32_modules) -> bool:
    """fp16 loading keeps the modules named in ``keep_in_fp32_modules`` (matched as a path component, or the full name) in fp32."""
    return any(key == param_name or (key in param_name and f"{key}." in param_name) for key in keep_in_fp32_modules or ())


def _receive_from_rank0(model: nn.Module, weights: dict) -> None:
    """``broadcast_from_rank0``: rank 0 holds ``weights`` (one checkpoint file), every other rank an empty dict.  The names rank 0
    has for this model are announced, then each tensor is broadcast and copied into the model — into the local shard when the
    target is a ``DTensor`` (torch ``fully_shard``), whole otherwise (plain modules, the module inside a DDP wrapper).  The reference
    gets this from ``set_model_state_dict(..., broadcast_from_rank0=True)`` (utils/modeling.py:1934-1953)."""
    import torch.distributed as dist

    targets = model.state_dict()
    announced = [[name for name in weights if name in targets]] if dist.get_rank() == 0 else [None]
    dist.broadcast_object_list(announced, src=0)
    with torch.no_grad():
        for name in announced[0]:
            target = targets[name]
            sharded = hasattr(target, "device_mesh") and hasattr(target, "placements")
            local = target.to_local() if sharded else target
            staging_device = local.device if local.device.type != "meta" else torch.device("cpu")
            full = torch.empty(tuple(target.shape), dtype=target.dtype, device=staging_device)
            if dist.get_rank() == 0:
                full.copy_(weights[name])
            dist.broadcast(full, src=0)
            if sharded:
                from torch.distributed.tensor import distribute_tensor

                local.copy_(distribute_tensor(full, target.device_mesh, target.placements).to_local())
            else:
                target.copy_(full)


def load_checkpoint_in_model(model: nn.Module, checkpoint, device_map: Optional[dict] = None, offload_folder=None, dtype=None,
                             offload_state_dict: bool = False, offload_buffers: bool = False, keep_in_fp32_modules: list = None,
                             offload_8bit_bnb: bool = False, strict: bool = False, full_state_dict: bool = True,
                             broadcast_from_rank0: bool = False) -> None:
    """Stream ``checkpoint`` (file, index or folder; sharded or not) into ``model``, one file at a time.

    With a ``device_map`` every weight goes where its closest enclosing entry says: a GPU or the CPU (set in place), or ``"disk"``
    (written to ``offload_folder`` as a memory-mappable file, the model keeps a meta tensor).  ``offload_state_dict`` parks the CPU
    weights on disk while the other files are read, to bound host memory.  Without a map the files are loaded the ordinary way — or,
    with ``broadcast_from_rank0`` in a distributed run, only rank 0 reads each file and the tensors are broadcast: into the native
    sharded engine through ``load_full_state_dict``, into ``DTensor`` parameters shard by shard, into plain modules whole.
    Tied weights are tied again at the end.
    """
    tied = find_tied_parameters(model)
    if not tied and check_tied_parameters_in_config(model):
        logger.warning("The model weights are not tied. Please use the `tie_weights` method before using the `infer_auto_device` function.")
    spills_to_disk = device_map is not None and "disk" in device_map.values()
    if device_map is not None:
        check_tied_parameters_on_same_device(tied, device_map)
    if spills_to_disk:
        if offload_folder is None:
            raise ValueError("At least one of the model submodule will be offloaded to disk, please pass along an `offload_folder`.")
        os.makedirs(offload_folder, exist_ok=True)
    dtype = _as_torch_dtype(dtype)
    files = _checkpoint_files(checkpoint)

    expected = set(model.state_dict())
    buffers = {name for name, _ in model.named_buffers()}
    real_devices = {t.device for t in model.state_dict().values() if isinstance(t, torch.Tensor)} - {_META}
    disk_index: dict = {}
    parked_index: dict = {}
    parked_folder = tempfile.mkdtemp() if offload_state_dict else None
    unexpected = set()

    from_rank0 = broadcast_from_rank0 and device_map is None and torch.distributed.is_available() and torch.distributed.is_initialized()
    reader = not from_rank0 or torch.distributed.get_rank() == 0

    for file in files:
        weights = load_state_dict(file, device_map=device_map) if reader else {}
        unexpected.update(name for name in weights if name not in expected and "SCB" not in name)
        if device_map is None:
            if hasattr(model, "load_full_state_dict"):
                # native sharded engine: one file may hold part of the model; with `broadcast_from_rank0` only rank 0 has values
                model.load_full_state_dict(weights, strict=False)
            elif from_rank0:
                _receive_from_rank0(model, weights)
            else:
                if len(real_devices) == 1:
                    (where,) = real_devices
                    weights = {name: tensor.to(where) for name, tensor in weights.items()}
                model.load_state_dict(weights, strict=strict)
        else:
            for name, tensor in weights.items():
                if "SCB" in name or (name not in expected and not strict):
                    continue
                target = _device_entry_for(name, device_map)
                wanted = dtype
                if dtype == torch.float16 and torch.is_floating_point(tensor) and _stays_fp32(name, keep_in_fp32_modules):
                    wanted = torch.float32
                statistics = None
                if tensor.dtype == torch.int8 and "weight" in name:
                    statistics = weights.get(name.replace("weight", "SCB"))
                if target == "disk":
                    if offload_buffers or name not in buffers:
                        stored_dtype = wanted or tensor.dtype
                        if offload_8bit_bnb:
                            from .bnb import quantize_and_offload_8bit

                            quantize_and_offload_8bit(model, tensor, name, stored_dtype, offload_folder, disk_index, statistics)
                            continue
                        set_module_tensor_to_device(model, name, "meta", dtype=stored_dtype)
                        offload_weight(tensor.to(stored_dtype) if torch.is_floating_point(tensor) else tensor, name, offload_folder, index=disk_index)
                elif target == "cpu" and offload_state_dict:
                    if offload_8bit_bnb:
                        from .bnb import quantize_and_offload_8bit

                        quantize_and_offload_8bit(model, tensor, name, wanted or tensor.dtype, parked_folder, parked_index, statistics)
                        continue
                    set_module_tensor_to_device(model, name, "meta", dtype=wanted or tensor.dtype)
                    offload_weight(tensor, name, parked_folder, index=parked_index)
                else:
                    set_module_tensor_to_device(model, name, target, value=tensor, dtype=wanted, fp16_statistics=statistics)
        del weights
        gc.collect()

    if unexpected and not strict:
        logger.warning(f"Some weights of the model checkpoint at {checkpoint} were not used when initializing {type(model).__name__}: "
                       f"{unexpected}. This may or may not be an issue - make sure that the checkpoint does not have unnecessary "
                       "parameters, or that the model definition correctly corresponds to the checkpoint.")
    save_offload_index(disk_index, offload_folder)
    if parked_folder is not None:
        load_offloaded_weights(model, parked_index, parked_folder)
        shutil.rmtree(parked_folder)
    retie_parameters(model, tied)


# ------------------------------------------------------------------------------------------------- autocast / scaler
def get_mixed_precision_context_manager(native_amp: bool = False, autocast_kwargs: AutocastKwargs = None):
    """The ``torch.autocast`` context for the active mixed-precision mode; a no-op context when AMP is off.  (The native bf16 engine
    does not need it — its kernels take bf16 in and accumulate in fp32 — this is for arbitrary ``nn.Module`` s.)"""
    if not native_amp:
        return contextlib.nullcontext()
    from ..state import AcceleratorState

    state = AcceleratorState()
    options = autocast_kwargs.to_kwargs() if autocast_kwargs is not None else {}
    on_gpu = state.device.type == "cuda"
    device_type = "cuda" if on_gpu else "cpu"
    low_precision = {"fp16": torch.float16 if on_gpu else torch.bfloat16, "bf16": torch.bfloat16, "fp8": torch.bfloat16}
    if state.mixed_precision in low_precision:
        options = {"dtype": low_precision[state.mixed_precision], **options}
    return torch.autocast(device_type=device_type, **options)


def get_grad_scaler(distributed_type: DistributedType = None, **kwargs):
    """Loss scaler for fp16.  One kind serves every engine here: sharded gradients are unscaled shard by shard and the inf check is
    all-reduced by the engine itself."""
    return torch.amp.GradScaler("cuda" if torch.cuda.is_available() else "cpu", **kwargs)
