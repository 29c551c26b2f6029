"""Model-surgery utilities for big-model inference: tensor placement, size accounting, device maps,
checkpoint streaming.

Behaviour parity with reference src/accelerate/utils/modeling.py: ``set_module_tensor_to_device``
:217-424, ``named_module_tensors`` :427-474, ``find_tied_parameters`` :554-606, ``retie_parameters``
:609-648, ``compute_module_sizes`` :651-702, ``get_max_layer_size`` :705-741, ``get_max_memory``
:744-852, ``clean_device_map`` :855-877, ``get_balanced_memory`` :918-1049,
``infer_auto_device_map`` :1278-1583, ``check_device_map`` :1586-1617, ``load_state_dict``
:1620-1712, ``load_checkpoint_in_model`` :1788-2046, ``get_mixed_precision_context_manager``
:2049-2089, ``get_grad_scaler`` :2092-2131.  CUDA + CPU + disk only; the greedy packer is written
from the behaviour (reserve room for the largest un-splittable layer, keep tied weights together,
optionally look ahead for a module that fits).
"""

from __future__ import annotations

import contextlib
import gc
import json
import logging
import os
import re
import warnings
from collections import OrderedDict, defaultdict
from typing import Optional, Union

import torch
import torch.nn as nn

from .constants import SAFE_WEIGHTS_NAME, WEIGHTS_NAME
from .dataclasses import AutocastKwargs, CustomDtype, DistributedType
from .memory import clear_device_cache
from .offload import load_offloaded_weight, offload_weight, save_offload_index

logger = logging.getLogger(__name__)

WEIGHTS_INDEX_NAME = "pytorch_model.bin.index.json"


# ------------------------------------------------------------------------------------ sizes
def convert_file_size_to_int(size: Union[int, str]) -> int:
    """``"5GB"`` -> 5_000_000_000, ``"2MiB"`` -> 2 * 2**20 (bit units end in lowercase ``b``)."""
    if isinstance(size, int):
        return size
    units = {"GIB": 2**30, "MIB": 2**20, "KIB": 2**10, "GB": 10**9, "MB": 10**6, "KB": 10**3}
    up = size.upper()
    for suffix, mult in units.items():
        if up.endswith(suffix):
            try:
                value = int(float(size[: -len(suffix)]) * mult) if "." in size[: -len(suffix)] else int(size[: -len(suffix)]) * mult
            except ValueError:
                break
            return value // 8 if size.endswith("b") else value
    raise ValueError("`size` is not in a valid format. Use an integer followed by the unit, e.g., '5GB'.")


def dtype_byte_size(dtype) -> float:
    """Bytes per element, including the sub-byte ``CustomDtype`` s."""
    if dtype == torch.bool:
        return 1 / 8
    if dtype == CustomDtype.INT2:
        return 1 / 4
    if dtype == CustomDtype.INT4:
        return 1 / 2
    if dtype == CustomDtype.FP8:
        return 1
    bit_search = re.search(r"[^\d](\d+)(_.*)?$", str(dtype))
    if bit_search is None:
        raise ValueError(f"`dtype` is not a valid dtype: {dtype}.")
    return int(bit_search.groups()[0]) // 8


def id_tensor_storage(tensor: torch.Tensor):
    """(device, storage pointer, storage size): equal for tensors that share memory."""
    try:
        storage_ptr = tensor.untyped_storage().data_ptr()
        storage_size = tensor.untyped_storage().nbytes()
    except Exception:
        storage_ptr, storage_size = 0, 0
    return tensor.device, storage_ptr, storage_size


def is_peft_model(model) -> bool:
    try:
        from peft import PeftModel

        return isinstance(model, PeftModel)
    except ImportError:
        return False


def check_device_same(first_device, second_device) -> bool:
    """``cuda`` == ``cuda:0`` (index defaults to 0); otherwise exact type+index match."""
    if first_device.type != second_device.type:
        return False
    if first_device.type != "cpu" and first_device.index is None:
        first_device = torch.device(first_device.type, index=0)
    if second_device.type != "cpu" and second_device.index is None:
        second_device = torch.device(second_device.type, index=0)
    return first_device == second_device


# ------------------------------------------------------------------------------------ tensor placement
def set_module_tensor_to_device(module: nn.Module, tensor_name: str, device, value: Optional[torch.Tensor] = None, dtype=None,
                                fp16_statistics=None, tied_params_map=None, non_blocking: bool = False, clear_cache: bool = True):
    """Re-home one parameter or buffer of ``module`` (``tensor_name`` may be dotted).

    Moving *off* the meta device needs ``value``.  Parameter-ness / ``requires_grad`` are kept, tied
    parameters are re-used through ``tied_params_map`` ``{data_ptr: {device: tensor}}`` so a shared
    weight is materialised once per device.
    """
    if "." in tensor_name:
        splits = tensor_name.split(".")
        for split in splits[:-1]:
            new_module = getattr(module, split)
            if new_module is None:
                raise ValueError(f"{module} has no attribute {split}.")
            module = new_module
        tensor_name = splits[-1]
    if tensor_name not in module._parameters and tensor_name not in module._buffers:
        raise ValueError(f"{module} does not have a parameter or a buffer named {tensor_name}.")
    is_buffer = tensor_name in module._buffers
    old_value = getattr(module, tensor_name)

    # a tied parameter already materialised on the target device: alias it
    if value is not None and tied_params_map is not None and value.data_ptr() in tied_params_map and device in tied_params_map[value.data_ptr()]:
        module._parameters[tensor_name] = tied_params_map[value.data_ptr()][device]
        return
    elif tied_params_map is not None and old_value.data_ptr() in tied_params_map and device in tied_params_map[old_value.data_ptr()]:
        module._parameters[tensor_name] = tied_params_map[old_value.data_ptr()][device]
        return

    if old_value.device == torch.device("meta") and device not in ["meta", torch.device("meta")] and value is None:
        raise ValueError(f"{tensor_name} is on the meta device, we need a `value` to put in on {device}.")
    param = module._parameters[tensor_name] if tensor_name in module._parameters else None
    param_cls = type(param)

    if value is not None:
        if old_value.shape != value.shape and param_cls.__name__ != "Params4bit":
            raise ValueError(f'Trying to set a tensor of shape {value.shape} in "{tensor_name}" (which has shape {old_value.shape}), this looks incorrect.')
        if dtype is None:
            value = value.to(old_value.dtype, non_blocking=non_blocking)
        elif not str(value.dtype).startswith(("torch.uint", "torch.int", "torch.bool")):
            value = value.to(dtype, non_blocking=non_blocking)

    device_quantization = None
    with torch.no_grad():
        if isinstance(device, int):
            device = torch.device("cuda", device)
        if "cuda" in str(device) and clear_cache and False:
            clear_device_cache()
        if value is None:
            new_value = old_value.to(device, non_blocking=non_blocking)
            if dtype is not None and device in ["meta", torch.device("meta")]:
                if not str(old_value.dtype).startswith(("torch.uint", "torch.int", "torch.bool")):
                    new_value = new_value.to(dtype, non_blocking=non_blocking)
                if not is_buffer:
                    module._parameters[tensor_name] = param_cls(new_value, requires_grad=old_value.requires_grad)
        elif isinstance(value, torch.Tensor):
            new_value = value.to(device, non_blocking=non_blocking)
        else:
            new_value = torch.tensor(value, device=device)
        if is_buffer:
            module._buffers[tensor_name] = new_value
        elif value is not None or not check_device_same(torch.device(device), module._parameters[tensor_name].device):
            param_cls = type(module._parameters[tensor_name])
            kwargs = module._parameters[tensor_name].__dict__
            if param_cls.__name__ in ("Int8Params", "FP4Params", "Params4bit"):
                if param_cls.__name__ == "Int8Params" and new_value.dtype == torch.float32:
                    new_value = new_value.to(torch.float16, non_blocking=non_blocking)
                new_value = param_cls(new_value, requires_grad=old_value.requires_grad, **kwargs).to(device)
            elif param_cls.__name__ in ("QTensor", "QBitsTensor", "AffineQuantizedTensor"):
                new_value = torch.nn.Parameter(new_value, requires_grad=old_value.requires_grad).to(device)
            else:
                new_value = param_cls(new_value, requires_grad=old_value.requires_grad).to(device, non_blocking=non_blocking)
            module._parameters[tensor_name] = new_value
            if fp16_statistics is not None:
                module._parameters[tensor_name].SCB = fp16_statistics.to(device, non_blocking=non_blocking)
                del fp16_statistics
    # remember materialised tied parameters
    if tied_params_map is not None and old_value.data_ptr() in tied_params_map and device not in tied_params_map[old_value.data_ptr()]:
        tied_params_map[old_value.data_ptr()][device] = new_value
    elif value is not None and tied_params_map is not None and value.data_ptr() in tied_params_map and device not in tied_params_map[value.data_ptr()]:
        tied_params_map[value.data_ptr()][device] = new_value


def named_module_tensors(module: nn.Module, include_buffers: bool = True, recurse: bool = False, remove_non_persistent: bool = False):
    """(name, tensor) for parameters and, optionally, buffers (non-persistent ones can be dropped)."""
    yield from module.named_parameters(recurse=recurse)
    if include_buffers:
        non_persistent_buffers = set()
        if remove_non_persistent:
            non_persistent_buffers = get_non_persistent_buffers(module, recurse=recurse)
        for named_buffer in module.named_buffers(recurse=recurse):
            name, _ = named_buffer
            if name not in non_persistent_buffers:
                yield named_buffer


def get_non_persistent_buffers(module: nn.Module, recurse: bool = False, fqns: bool = False):
    non_persistent_buffers_set = module._non_persistent_buffers_set
    if fqns:
        non_persistent_buffers_set = set(non_persistent_buffers_set)
    if recurse:
        for n, m in module.named_modules():
            if fqns:
                non_persistent_buffers_set |= {n + "." + b for b in m._non_persistent_buffers_set}
            else:
                non_persistent_buffers_set |= m._non_persistent_buffers_set
    return non_persistent_buffers_set


class FindTiedParametersResult(list):
    """List of groups of tied-parameter names (with a deprecated dict-style ``values`` accessor)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)

    def values(self):
        warnings.warn("The 'values' method of FindTiedParametersResult is deprecated and will be removed in Accelerate v1.3.0. ", FutureWarning)
        return sum([x[1:] for x in self], [])


def check_tied_parameters_in_config(model: nn.Module) -> bool:
    """Does the (transformers) config say that some weights are tied?"""
    has_tied_word_embedding = False
    has_tied_encoder_decoder = False
    has_tied_module = False
    if "PreTrainedModel" in [c.__name__ for c in model.__class__.__mro__]:
        has_tied_word_embedding = False
        config = getattr(model, "config", None)
        if config is not None:
            try:
                text = config.get_text_config(decoder=True)
            except Exception:
                text = config
            has_tied_word_embedding = bool(getattr(text, "tie_word_embeddings", False) and model.get_output_embeddings())
            has_tied_encoder_decoder = bool(getattr(config, "is_encoder_decoder", False) and getattr(config, "tie_encoder_decoder", False))
        has_tied_module = any(hasattr(module, "_tie_weights") for module in model.modules())
    elif getattr(getattr(model, "config", None), "tie_word_embeddings", False):
        has_tied_word_embedding = True
    return any([has_tied_word_embedding, has_tied_encoder_decoder, has_tied_module])


def _get_param_device(param, device_map):
    if param in device_map:
        return device_map[param]
    parent_param = ".".join(param.split(".")[:-1])
    if parent_param == param:
        raise ValueError(f"The `device_map` does not contain the module {param}.")
    return _get_param_device(parent_param, device_map)


def check_tied_parameters_on_same_device(tied_params, device_map):
    """Warn when tied parameters were mapped to different devices (they will be duplicated)."""
    for tie_param in tied_params:
        tie_param_devices = {}
        for param in tie_param:
            tie_param_devices[param] = _get_param_device(param, device_map)
        if len(set(tie_param_devices.values())) > 1:
            logger.warning(f"Tied parameters are on different devices: {tie_param_devices}. Please modify your custom device map or set `device_map='auto'`. ")


def find_tied_parameters(model: nn.Module, **kwargs) -> list:
    """Groups of parameter names that are the same tensor object."""
    all_named_parameters = dict(model.named_parameters(remove_duplicate=False))
    no_duplicate_named_parameters = dict(model.named_parameters(remove_duplicate=True))
    tied_param_names = set(all_named_parameters.keys()) - set(no_duplicate_named_parameters.keys())
    tied_param_groups = {}
    for tied_param_name in tied_param_names:
        tied_param = all_named_parameters[tied_param_name]
        for param_name, param in no_duplicate_named_parameters.items():
            if param is tied_param:
                tied_param_groups.setdefault(param_name, []).append(tied_param_name)
    return FindTiedParametersResult([sorted([weight] + list(set(tied))) for weight, tied in tied_param_groups.items()])


def retie_parameters(model: nn.Module, tied_params):
    """Make the parameters of each group the same object again (after loading broke the tie)."""
    for tied_group in tied_params:
        param_to_tie = None
        for param_name in tied_group:
            module = model
            splits = param_name.split(".")
            for split in splits[:-1]:
                module = getattr(module, split)
            param = getattr(module, splits[-1])
            if param_to_tie is None and param.device != torch.device("meta"):
                param_to_tie = param
                break
        if param_to_tie is not None:
            for param_name in tied_group:
                module = model
                splits = param_name.split(".")
                for split in splits[:-1]:
                    module = getattr(module, split)
                setattr(module, splits[-1], param_to_tie)


def _get_proper_dtype(dtype):
    if isinstance(dtype, str):
        dtype = dtype.replace("torch.", "")
        dtype = getattr(torch, dtype)
    return dtype


def compute_module_sizes(model: nn.Module, dtype=None, special_dtypes: Optional[dict] = None, buffers_only: bool = False):
    """Bytes of every (sub)module name, ``""`` being the whole model."""
    if dtype is not None:
        dtype = _get_proper_dtype(dtype)
        dtype_size = dtype_byte_size(dtype)
    if special_dtypes is not None:
        special_dtypes = {key: _get_proper_dtype(dtyp) for key, dtyp in special_dtypes.items()}
        special_dtypes_size = {key: dtype_byte_size(dtyp) for key, dtyp in special_dtypes.items()}
    module_sizes = defaultdict(int)
    module_list = []
    if not buffers_only:
        module_list = named_module_tensors(model, recurse=True)
    else:
        module_list = model.named_buffers(recurse=True)
    for name, tensor in module_list:
        if special_dtypes is not None and name in special_dtypes:
            size = tensor.numel() * special_dtypes_size[name]
        elif dtype is None:
            size = tensor.numel() * dtype_byte_size(tensor.dtype)
        elif str(tensor.dtype).startswith(("torch.uint", "torch.int", "torch.bool")):
            size = tensor.numel() * dtype_byte_size(tensor.dtype)
        else:
            size = tensor.numel() * min(dtype_size, dtype_byte_size(tensor.dtype))
        name_parts = name.split(".")
        for idx in range(len(name_parts) + 1):
            module_sizes[".".join(name_parts[:idx])] += size
    return module_sizes


def compute_module_total_buffer_size(model: nn.Module, dtype=None, special_dtypes=None):
    module_sizes = compute_module_sizes(model, dtype=dtype, special_dtypes=special_dtypes, buffers_only=True)
    return module_sizes.get("", 0)


def get_max_layer_size(modules: list, module_sizes: dict, no_split_module_classes: list):
    """(size, names) of the largest module that cannot be split further."""
    max_size = 0
    layer_names = []
    modules_to_treat = modules.copy()
    while len(modules_to_treat) > 0:
        module_name, module = modules_to_treat.pop(0)
        modules_children = list(module.named_children()) if isinstance(module, torch.nn.Module) else []
        if len(modules_children) == 0 or module.__class__.__name__ in no_split_module_classes:
            size = module_sizes[module_name]
            if size > max_size:
                max_size = size
                layer_names = [module_name]
            elif size == max_size:
                layer_names.append(module_name)
        else:
            modules_to_treat = [(f"{module_name}.{n}", v) for n, v in modules_children] + modules_to_treat
    return max_size, layer_names


def get_max_memory(max_memory: Optional[dict] = None):
    """Per-device byte budget: free CUDA memory per GPU plus available CPU RAM when not given."""
    import psutil

    if max_memory is None:
        max_memory = {}
        if torch.cuda.is_available():
            for i in range(torch.cuda.device_count()):
                try:
                    _ = torch.tensor([0], device=i)
                    max_memory[i] = torch.cuda.mem_get_info(i)[0]
                except Exception:
                    logger.info(f"Device {i} seems unavailable, Proceeding to check subsequent devices.")
                    continue
        max_memory["cpu"] = psutil.virtual_memory().available
        return max_memory
    for key in max_memory:
        if isinstance(max_memory[key], str):
            max_memory[key] = convert_file_size_to_int(max_memory[key])
    gpu_devices = sorted([k for k in max_memory.keys() if isinstance(k, int)])
    num_devices = torch.cuda.device_count() if torch.cuda.is_available() else 0
    for device in gpu_devices:
        if device >= num_devices or device < 0:
            logger.warning(f"Device {device} is not available, available devices are {list(range(num_devices))}")
    all_devices = gpu_devices + [k for k in ["mps", "cpu", "disk"] if k in max_memory.keys()]
    for k in max_memory.keys():
        if k not in all_devices:
            raise ValueError(f"Device {k} is not recognized, available devices are integers(for GPU/XPU), 'mps', 'cpu' and 'disk'")
    return {k: max_memory[k] for k in all_devices}


def clean_device_map(device_map: dict, module_name: str = ""):
    """Collapse children that all live on one device into their parent entry."""
    prefix = "" if module_name == "" else f"{module_name}."
    values = [v for k, v in device_map.items() if k.startswith(prefix)]
    if len(set(values)) == 1 and len(values) > 1:
        for k in [k for k in device_map if k.startswith(prefix)]:
            del device_map[k]
        device_map[module_name] = values[0]
    children_modules = [k for k in device_map.keys() if k.startswith(prefix) and len(k) > len(module_name)]
    idx = len(module_name.split(".")) + 1 if len(module_name) > 0 else 1
    children = set(".".join(k.split(".")[:idx]) for k in children_modules)
    for child in children:
        clean_device_map(device_map, module_name=child)
    return device_map


def load_offloaded_weights(model, index, offload_folder):
    """Load disk-offloaded weights listed in ``index`` into ``model`` on CPU."""
    if index is None or len(index) == 0:
        return
    for param_name, metadata in index.items():
        if "SCB" in param_name:
            continue
        fp16_statistics = None
        if "weight" in param_name and param_name.replace("weight", "SCB") in index.keys():
            weight_name = param_name.replace("weight", "SCB")
            fp16_statistics = load_offloaded_weight(os.path.join(offload_folder, f"{weight_name}.dat"), index[weight_name])
        tensor_file = os.path.join(offload_folder, f"{param_name}.dat")
        weight = load_offloaded_weight(tensor_file, metadata)
        set_module_tensor_to_device(model, param_name, "cpu", value=weight, fp16_statistics=fp16_statistics)


def get_module_leaves(module_sizes):
    module_children = {}
    for module in module_sizes:
        if module == "" or "." not in module:
            continue
        parent = module.rsplit(".", 1)[0]
        module_children[parent] = module_children.get(parent, 0) + 1
    return [module for module in module_sizes if module_children.get(module, 0) == 0 and not module.endswith(("weight", "bias"))]


def get_balanced_memory(model: nn.Module, max_memory: Optional[dict] = None, no_split_module_classes: Optional[list] = None,
                        dtype=None, special_dtypes: Optional[dict] = None, low_zero: bool = False):
    """Budgets that spread ``model`` evenly over the GPUs (``low_zero`` keeps GPU 0 nearly empty)."""
    user_not_set_max_memory = max_memory is None
    max_memory = get_max_memory(max_memory)
    num_devices = len([d for d in max_memory if torch.device(d).type == "cuda" and max_memory[d] > 0]) if torch.cuda.is_available() else 0
    if num_devices == 0:
        return max_memory
    if num_devices == 1:
        low_zero = False
        if user_not_set_max_memory:
            for key in max_memory.keys():
                if isinstance(key, int):
                    max_memory[key] *= 0.9
                    logger.info(f"We will use 90% of the memory on device {key} for storing the model, and 10% for the buffer to avoid OOM. You can set `max_memory` in to a higher value to use more memory (at your own risk).")
                    break
    module_sizes = compute_module_sizes(model, dtype=dtype, special_dtypes=special_dtypes)
    per_gpu = module_sizes[""] // (num_devices - 1 if low_zero else num_devices)
    if no_split_module_classes is None:
        no_split_module_classes = []
    elif not isinstance(no_split_module_classes, (list, tuple)):
        no_split_module_classes = [no_split_module_classes]
    if len(no_split_module_classes) > 0:
        no_split_children = {}
        for name, size in module_sizes.items():
            if name == "":
                continue
            submodule = model
            for submodule_name in name.split("."):
                submodule = getattr(submodule, submodule_name)
            class_name = submodule.__class__.__name__
            if class_name in no_split_module_classes and class_name not in no_split_children:
                no_split_children[class_name] = size
            if set(no_split_children.keys()) == set(no_split_module_classes):
                break
        buffer = max(no_split_children.values()) if len(no_split_children) > 0 else 0
    else:
        buffer = 0
    leaves = get_module_leaves(module_sizes)
    module_sizes = {n: v for n, v in module_sizes.items() if n not in leaves}
    leaves = get_module_leaves(module_sizes)
    mean_leaves = int(sum([module_sizes[n] for n in leaves]) / max(len(leaves), 1))
    buffer = int(1.25 * max(buffer, mean_leaves))
    per_gpu += buffer
    gpus_idx_list = list(sorted(device_id for device_id, device_mem in max_memory.items() if isinstance(device_id, int) and device_mem > 0))
    for idx in gpus_idx_list[:-1]:
        max_memory[idx] = min(max_memory[0] if low_zero and idx == 0 else per_gpu, max_memory[idx])
    if low_zero:
        min_zero = max(0, module_sizes[""] - sum([max_memory[i] for i in range(1, num_devices)]))
        max_memory[0] = min(min_zero, max_memory[0])
    return max_memory


def calculate_maximum_sizes(model: torch.nn.Module):
    """(total bytes, (largest layer bytes, names))."""
    sizes = compute_module_sizes(model)
    no_split_modules = getattr(model, "_no_split_modules", None)
    if no_split_modules is None:
        no_split_modules = []
    modules_to_treat = list(model.named_parameters(recurse=False)) + list(model.named_children()) + list(model.named_buffers(recurse=False))
    largest_layer = get_max_layer_size(modules_to_treat, sizes, no_split_modules)
    total_size = sizes[""]
    return total_size, largest_layer


# ------------------------------------------------------------------------------------ device map inference
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
    modules_to_treat = (list(model.named_parameters(recurse=False)) + list(model.named_children()) + list(model.named_buffers(recurse=False)))
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
        partner_modules = []
        for p in partners:
            for cand_name, cand in modules_to_treat:
                if p == cand_name or p.startswith(cand_name + "."):
                    if (cand_name, cand) not in partner_modules:
                        partner_modules.append((cand_name, cand))
        total_size = module_size + sum(module_sizes[n] for n, _ in partner_modules)
        device = devices[dev_idx]
        cap = budget(dev_idx)
        if cap is None or used[device] + total_size <= cap:
            device_map[name] = device
            used[device] += module_size
            for n, m in partner_modules:
                device_map[n] = device
                used[device] += module_sizes[n]
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
        # cannot split: optionally fill the remaining room with a later module that fits
        if fallback_allocation and cap is not None:
            placed = False
            queue = list(modules_to_treat)
            while queue:
                cn, cm = queue.pop(0)
                csize = module_sizes[cn]
                if used[device] + csize <= cap and not tied_group_of(cn):
                    device_map[cn] = device
                    used[device] += csize
                    # remove it (or the ancestor chain it was expanded from) from the work list
                    if (cn, cm) in modules_to_treat:
                        modules_to_treat.remove((cn, cm))
                        placed = True
                        break
                    continue
                if isinstance(cm, nn.Module) and cm.__class__.__name__ not in no_split_module_classes and (cn, cm) in modules_to_treat:
                    sub = children_of(cn, cm)
                    if sub and any(module_sizes[s] + used[device] <= cap for s, _ in sub):
                        i = modules_to_treat.index((cn, cm))
                        modules_to_treat = modules_to_treat[:i] + sub + modules_to_treat[i + 1:]
                        queue = sub + queue
            if placed:
                modules_to_treat.insert(0, (name, module))
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
        is_buffer_fit_any_gpu = any(max_memory[g] - used[g] >= non_gpu_buffer_size for g in gpus if used[g] > 0 or True)
        if len(gpus) > 0 and not is_buffer_fit_any_gpu:
            warnings.warn(
                f"Current model requires {non_gpu_buffer_size} bytes of buffer for offloaded layers, which seems does "
                f"not fit any GPU's remaining memory. If you are experiencing a OOM later, please consider using offload_buffers=True."
            )
    if verbose:
        print({d: used[d] for d in used if used[d]})
    return device_map


def check_device_map(model: nn.Module, device_map: dict):
    """Every parameter must be covered by an entry of ``device_map``."""
    all_module_names = dict(model.named_modules())
    invalid_keys = [k for k in device_map if k != "" and k not in all_module_names and k not in dict(named_module_tensors(model, recurse=True))]
    if invalid_keys:
        warnings.warn(f"The following device_map keys do not match any submodules in the model: {invalid_keys}", UserWarning)
    all_model_tensors = [name for name, _ in model.state_dict().items()]
    for module_name in device_map.keys():
        if module_name == "":
            all_model_tensors.clear()
            break
        all_model_tensors = [name for name in all_model_tensors if not name == module_name and not name.startswith(module_name + ".")]
    if len(all_model_tensors) > 0:
        non_covered_params = ", ".join(all_model_tensors)
        raise ValueError(f"The device_map provided does not give any device for the following parameters: {non_covered_params}")


# ------------------------------------------------------------------------------------ checkpoint IO
def load_state_dict(checkpoint_file, device_map=None):
    """Load a (safetensors or torch) checkpoint file; with a ``device_map`` safetensors go straight to their devices."""
    if str(checkpoint_file).endswith(".safetensors"):
        from safetensors import safe_open
        from safetensors.torch import load_file as safe_load_file

        with safe_open(checkpoint_file, framework="pt") as f:
            metadata = f.metadata()
            weight_names = f.keys()
        if metadata is None:
            logger.warning(f"The safetensors archive passed at {checkpoint_file} does not contain metadata. Make sure to save your model with the `save_pretrained` method. Defaulting to 'pt' metadata.")
            metadata = {"format": "pt"}
        if metadata.get("format") not in ["pt", "tf", "flax"]:
            raise OSError(f"The safetensors archive passed at {checkpoint_file} does not contain the valid metadata. Make sure you save your model with the `save_pretrained` method.")
        elif metadata["format"] != "pt":
            raise ValueError(f"The checkpoint passed was saved with {metadata['format']}, we need a the pt format.")
        if device_map is None:
            return safe_load_file(checkpoint_file)
        if len(set(device_map.values())) == 1:
            device = list(device_map.values())[0]
            target_device = device if device != "disk" else "cpu"
            if isinstance(target_device, int):
                target_device = f"cuda:{target_device}"
            return safe_load_file(checkpoint_file, device=target_device)
        devices = list(set(device_map.values()) - {"disk"})
        if "cpu" not in devices:
            devices.append("cpu")
        device_weights = {device: [] for device in devices}
        for module_name, device in device_map.items():
            if device in devices:
                device_weights[device].extend([k for k in weight_names if k == module_name or k.startswith(module_name + ".")])
        device_weights["cpu"].extend([k for k in weight_names if k not in sum(device_weights.values(), [])])
        tensors = {}
        for device in devices:
            target_device = f"cuda:{device}" if isinstance(device, int) else device
            with safe_open(checkpoint_file, framework="pt", device=target_device) as f:
                for key in device_weights[device]:
                    tensors[key] = f.get_tensor(key)
        return tensors
    return torch.load(checkpoint_file, map_location=torch.device("cpu"), weights_only=True)


def get_state_dict_offloaded_model(model: nn.Module):
    """State dict (on CPU) of a model whose weights are offloaded behind ``AlignDevicesHook`` s."""
    state_dict = {}
    placeholders = set()
    for name, module in model.named_modules():
        if name == "":
            continue
        try:
            with align_module_device(module, "cpu"):
                module_state_dict = module.state_dict()
        except MemoryError:
            raise MemoryError("Offloaded module must fit in CPU memory to call save_model!") from None
        for key in module_state_dict:
            if module_state_dict[key].device == torch.device("meta"):
                placeholders.add(name + f".{key}")
                continue
            params = module_state_dict[key]
            state_dict[name + f".{key}"] = params.to("cpu")
    for key in placeholders.copy():
        if key in state_dict:
            placeholders.remove(key)
    if placeholders:
        logger.warning(f"The following tensors were not saved because they were still on meta device: {placeholders}")
    return state_dict


def get_state_dict_from_offload(module: nn.Module, module_name: str, state_dict: dict, device_to_put_offload="cpu"):
    root = module_name[: module_name.rfind(".")]
    with align_module_device(module, device_to_put_offload):
        for m_key, params in module.state_dict().items():
            if (root + f".{m_key}") in state_dict:
                state_dict[root + f".{m_key}"] = params
    return state_dict


def shard_checkpoint(state_dict: dict, max_shard_size: Union[int, str] = "10GB", weights_name: str = WEIGHTS_NAME):
    """Split ``state_dict`` into shards of at most ``max_shard_size`` (HF naming ``name-0000X-of-0000Y.ext``).

    Returns ``({filename: shard}, index-or-None)``.
    """
    max_shard_size = convert_file_size_to_int(max_shard_size)
    sharded = [{}]
    last_block_size = 0
    total_size = 0
    storage_id_to_block = {}
    for key, weight in state_dict.items():
        if isinstance(weight, str):
            continue
        storage_id = id_tensor_storage(weight)
        if storage_id in storage_id_to_block and storage_id[1] != 0:
            sharded[storage_id_to_block[storage_id]][key] = weight
            continue
        weight_size = weight.numel() * dtype_byte_size(weight.dtype)
        if last_block_size + weight_size > max_shard_size and len(sharded[-1]) > 0:
            sharded.append({})
            last_block_size = 0
        sharded[-1][key] = weight
        last_block_size += weight_size
        total_size += weight_size
        storage_id_to_block[storage_id] = len(sharded) - 1
    if len(sharded) == 1:
        return {weights_name: sharded[0]}, None
    weight_map = {}
    shards = {}
    for idx, shard in enumerate(sharded):
        shard_file = weights_name.replace(".bin", f"-{idx + 1:05d}-of-{len(sharded):05d}.bin")
        shard_file = shard_file.replace(".safetensors", f"-{idx + 1:05d}-of-{len(sharded):05d}.safetensors")
        shards[shard_file] = shard
        for key in shard.keys():
            weight_map[key] = shard_file
    index = {"metadata": {"total_size": int(total_size)}, "weight_map": weight_map}
    return shards, index


def load_checkpoint_in_model(model: nn.Module, checkpoint, device_map: Optional[dict] = None, offload_folder=None, dtype=None,
                             offload_state_dict: bool = False, offload_buffers: bool = False, keep_in_fp32_modules: list = None,
                             offload_8bit_bnb: bool = False, strict: bool = False, full_state_dict: bool = True,
                             broadcast_from_rank0: bool = False):
    """Stream a (possibly sharded) checkpoint into ``model`` honouring ``device_map`` (gpu / cpu / disk)."""
    tied_params = find_tied_parameters(model)
    if check_tied_parameters_in_config(model) and len(tied_params) == 0:
        logger.warning("The model weights are not tied. Please use the `tie_weights` method before using the `infer_auto_device` function.")
    if device_map is not None:
        check_tied_parameters_on_same_device(tied_params, device_map)
    if offload_folder is None and device_map is not None and "disk" in device_map.values():
        raise ValueError("At least one of the model submodule will be offloaded to disk, please pass along an `offload_folder`.")
    elif offload_folder is not None and device_map is not None and "disk" in device_map.values():
        os.makedirs(offload_folder, exist_ok=True)
    if isinstance(dtype, str):
        dtype = dtype.replace("torch.", "")
        dtype = getattr(torch, dtype)

    checkpoint_files = None
    index_filename = None
    if os.path.isfile(checkpoint):
        if str(checkpoint).endswith(".json"):
            index_filename = checkpoint
        else:
            checkpoint_files = [checkpoint]
    elif os.path.isdir(checkpoint):
        potential_state_bin = [f for f in os.listdir(checkpoint) if f == WEIGHTS_NAME]
        potential_state_safetensor = [f for f in os.listdir(checkpoint) if f == SAFE_WEIGHTS_NAME]
        if len(potential_state_bin) == 1:
            checkpoint_files = [os.path.join(checkpoint, potential_state_bin[0])]
        elif len(potential_state_safetensor) == 1:
            checkpoint_files = [os.path.join(checkpoint, potential_state_safetensor[0])]
        else:
            potential_index = [f for f in os.listdir(checkpoint) if f.endswith(".index.json")]
            if len(potential_index) == 0:
                raise ValueError(f"{checkpoint} is not a folder containing a `.index.json` file or a {WEIGHTS_NAME} or a {SAFE_WEIGHTS_NAME} file")
            elif len(potential_index) == 1:
                index_filename = os.path.join(checkpoint, potential_index[0])
            else:
                raise ValueError(f"{checkpoint} containing more than one `.index.json` file, delete the irrelevant ones.")
    else:
        raise ValueError(f"`checkpoint` should be the path to a file containing a whole state dict, or the index of a sharded checkpoint, or a folder containing a sharded checkpoint or the whole state dict, but got {checkpoint}.")
    if index_filename is not None:
        checkpoint_folder = os.path.split(index_filename)[0]
        with open(index_filename) as f:
            index = json.loads(f.read())
        if "weight_map" in index:
            index = index["weight_map"]
        checkpoint_files = sorted(list(set(index.values())))
        checkpoint_files = [os.path.join(checkpoint_folder, f) for f in checkpoint_files]

    offload_index = {}
    if offload_state_dict:
        import tempfile

        state_dict_folder = tempfile.mkdtemp()
        state_dict_index = {}
    unexpected_keys = set()
    model_keys = set(model.state_dict().keys())
    buffer_names = [name for name, _ in model.named_buffers()]
    model_devices = {t.device for t in model.state_dict().values() if isinstance(t, torch.Tensor)}
    model_physical_devices = model_devices - {torch.device("meta")}
    for checkpoint_file in checkpoint_files:
        if device_map is None:
            # plain load (optionally rank-0 broadcast into a sharded engine)
            loaded_checkpoint = load_state_dict(checkpoint_file, device_map=None)
            if hasattr(model, "load_full_state_dict") and broadcast_from_rank0:
                model.load_full_state_dict(loaded_checkpoint)
            else:
                target = list(model_physical_devices)[0] if len(model_physical_devices) == 1 else None
                if target is not None:
                    loaded_checkpoint = {k: v.to(target) for k, v in loaded_checkpoint.items()}
                model.load_state_dict(loaded_checkpoint, strict=strict)
            unexpected_keys.update(set(loaded_checkpoint.keys()) - model_keys)
            del loaded_checkpoint
            gc.collect()
            continue
        loaded_checkpoint = load_state_dict(checkpoint_file, device_map=device_map)
        for param_name, param in loaded_checkpoint.items():
            if "SCB" in param_name:
                continue
            if param_name not in model_keys:
                unexpected_keys.add(param_name)
                if not strict:
                    continue
            module_name = param_name
            while len(module_name) > 0 and module_name not in device_map:
                module_name = ".".join(module_name.split(".")[:-1])
            if module_name == "" and "" not in device_map:
                raise ValueError(f"{param_name} doesn't have any device set.")
            param_device = device_map[module_name]
            new_dtype = dtype
            if dtype is not None and torch.is_floating_point(param):
                if keep_in_fp32_modules is not None and dtype == torch.float16:
                    proceed = False
                    for key in keep_in_fp32_modules:
                        if ((key in param_name) and (key + "." in param_name)) or key == param_name:
                            proceed = True
                            break
                    if proceed:
                        new_dtype = torch.float32
            if "weight" in param_name and param_name.replace("weight", "SCB") in loaded_checkpoint.keys():
                fp16_statistics = loaded_checkpoint[param_name.replace("weight", "SCB")] if param.dtype == torch.int8 else None
            else:
                fp16_statistics = None
            if param_device == "disk":
                if offload_buffers or param_name not in buffer_names:
                    if new_dtype is None:
                        new_dtype = param.dtype
                    set_module_tensor_to_device(model, param_name, "meta", dtype=new_dtype)
                    offload_weight(param.to(new_dtype) if torch.is_floating_point(param) else param, param_name, offload_folder, index=offload_index)
            elif param_device == "cpu" and offload_state_dict:
                if new_dtype is None:
                    new_dtype = param.dtype
                set_module_tensor_to_device(model, param_name, "meta", dtype=new_dtype)
                offload_weight(param, param_name, state_dict_folder, index=state_dict_index)
            else:
                set_module_tensor_to_device(model, param_name, param_device, value=param, dtype=new_dtype, fp16_statistics=fp16_statistics)
        del loaded_checkpoint
        gc.collect()
    if not strict and len(unexpected_keys) > 0:
        logger.warning(f"Some weights of the model checkpoint at {checkpoint} were not used when initializing {model.__class__.__name__}: {unexpected_keys}. This may or may not be an issue - make sure that the checkpoint does not have unnecessary parameters, or that the model definition correctly corresponds to the checkpoint.")
    save_offload_index(offload_index, offload_folder)
    if offload_state_dict:
        import shutil

        load_offloaded_weights(model, state_dict_index, state_dict_folder)
        shutil.rmtree(state_dict_folder)
    retie_parameters(model, tied_params)


# ------------------------------------------------------------------------------------ contexts
def get_mixed_precision_context_manager(native_amp: bool = False, autocast_kwargs: AutocastKwargs = None):
    """``torch.autocast`` for the active mixed-precision mode (nullcontext when none)."""
    from ..state import AcceleratorState

    state = AcceleratorState()
    autocast_kwargs = {} if autocast_kwargs is None else autocast_kwargs.to_kwargs()
    if not native_amp:
        return contextlib.nullcontext()
    device_type = "cuda" if state.device.type == "cuda" else "cpu"
    if state.mixed_precision == "fp16":
        return torch.autocast(device_type=device_type, dtype=torch.float16 if device_type == "cuda" else torch.bfloat16, **autocast_kwargs)
    if state.mixed_precision in ("bf16", "fp8"):
        return torch.autocast(device_type=device_type, dtype=torch.bfloat16, **autocast_kwargs)
    return torch.autocast(device_type=device_type, **autocast_kwargs)


def get_grad_scaler(distributed_type: DistributedType = None, **kwargs):
    """A ``torch.amp.GradScaler`` for the current device type (sharded engines need no special scaler)."""
    device = "cuda" if torch.cuda.is_available() else "cpu"
    return torch.amp.GradScaler(device, **kwargs)


def has_offloaded_params(module: nn.Module) -> bool:
    from ..hooks import AlignDevicesHook

    return hasattr(module, "_hf_hook") and isinstance(module._hf_hook, AlignDevicesHook) and module._hf_hook.offload


@contextlib.contextmanager
def align_module_device(module: nn.Module, execution_device=None):
    """Inside the block the module's (possibly offloaded) parameters live on ``execution_device``."""
    if has_offloaded_params(module):
        if execution_device is not None:
            original_device = module._hf_hook.execution_device
            module._hf_hook.execution_device = execution_device
        try:
            module._hf_hook.pre_forward(module)
            yield
        finally:
            module._hf_hook.post_forward(module, None)
            if execution_device is not None:
                module._hf_hook.execution_device = original_device
    elif execution_device is not None:
        devices = {name: param.device for name, param in module.named_parameters(recurse=False)}
        try:
            for name in devices:
                set_module_tensor_to_device(module, name, execution_device)
            yield
        finally:
            for name, device in devices.items():
                set_module_tensor_to_device(module, name, device)
    else:
        yield
