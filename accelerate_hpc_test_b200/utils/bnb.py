"""bitsandbytes 4/8-bit loading (reference src/accelerate/utils/bnb.py:44-469).  bitsandbytes kernels are
outside the B200 hot path; the loader API is kept behind an import guard."""

import logging
from copy import deepcopy

import torch
import torch.nn as nn

from .dataclasses import BnbQuantizationConfig
from .imports import is_4bit_bnb_available, is_8bit_bnb_available, is_bnb_available

logger = logging.getLogger(__name__)


def _require_bnb(cfg: BnbQuantizationConfig):
    if cfg.load_in_8bit and not is_8bit_bnb_available():
        raise ImportError("You have a version of `bitsandbytes` that is not compatible with 8bit quantization, make sure you have the latest version of `bitsandbytes` installed.")
    if cfg.load_in_4bit and not is_4bit_bnb_available():
        raise ValueError("You have a version of `bitsandbytes` that is not compatible with 4bit quantization, make sure you have the latest version of `bitsandbytes` installed.")


def get_keys_to_not_convert(model):
    """Names of modules to keep in full precision: tied weights and the output head."""
    tied_model = deepcopy(model)
    if hasattr(tied_model, "tie_weights"):
        tied_model.tie_weights()
    from .modeling import find_tied_parameters

    tied_params = find_tied_parameters(tied_model)
    tied_keys = sum(tied_params, []) if len(tied_params) > 0 else []
    has_tied_params = len(tied_keys) > 0
    is_base_model = not hasattr(model, getattr(model, "base_model_prefix", "__none__")) if hasattr(model, "base_model_prefix") else False
    if (not has_tied_params) and is_base_model:
        return []
    list_modules = list(model.named_children())
    list_last_module = [list_modules[-1][0]] if list_modules else []
    intersection = set(list_last_module) - set(tied_keys)
    list_untouched = list(set(tied_keys)) + list(intersection)
    names_to_remove = [".weight", ".bias"]
    filtered = []
    for name in list_untouched:
        for name_to_remove in names_to_remove:
            if name_to_remove in name:
                name = name.replace(name_to_remove, "")
        filtered.append(name)
    return filtered


def replace_with_bnb_layers(model, bnb_quantization_config, modules_to_not_convert=None, current_key_name=None):
    """Replace ``nn.Linear`` modules by ``bnb.nn.Linear8bitLt`` / ``Linear4bit``."""
    import bitsandbytes as bnb

    if modules_to_not_convert is None:
        modules_to_not_convert = []
    has_been_replaced = False
    for name, module in model.named_children():
        if current_key_name is None:
            current_key_name = []
        current_key_name.append(name)
        if isinstance(module, nn.Linear) and name not in modules_to_not_convert:
            current_key_name_str = ".".join(current_key_name)
            proceed = True
            for key in modules_to_not_convert:
                if ((key in current_key_name_str) and (key + "." in current_key_name_str)) or key == current_key_name_str:
                    proceed = False
                    break
            if proceed:
                if bnb_quantization_config.load_in_8bit:
                    bnb_module = bnb.nn.Linear8bitLt(module.in_features, module.out_features, module.bias is not None, has_fp16_weights=False,
                                                     threshold=bnb_quantization_config.llm_int8_threshold)
                elif bnb_quantization_config.load_in_4bit:
                    bnb_module = bnb.nn.Linear4bit(module.in_features, module.out_features, module.bias is not None,
                                                   bnb_quantization_config.bnb_4bit_compute_dtype,
                                                   compress_statistics=bnb_quantization_config.bnb_4bit_use_double_quant,
                                                   quant_type=bnb_quantization_config.bnb_4bit_quant_type)
                else:
                    raise ValueError("load_in_8bit and load_in_4bit can't be both False")
                bnb_module.weight.data = module.weight.data
                if module.bias is not None:
                    bnb_module.bias.data = module.bias.data
                bnb_module.requires_grad_(False)
                setattr(model, name, bnb_module)
                has_been_replaced = True
        if len(list(module.children())) > 0:
            _, _has_been_replaced = replace_with_bnb_layers(module, bnb_quantization_config, modules_to_not_convert, current_key_name)
            has_been_replaced = has_been_replaced | _has_been_replaced
        current_key_name.pop(-1)
    return model, has_been_replaced


def has_4bit_bnb_layers(model) -> bool:
    if not is_bnb_available():
        return False
    import bitsandbytes as bnb

    return any(isinstance(m, bnb.nn.Linear4bit) for m in model.modules())


def load_and_quantize_model(model: torch.nn.Module, bnb_quantization_config: BnbQuantizationConfig, weights_location=None, device_map=None,
                            no_split_module_classes=None, max_memory=None, offload_folder=None, offload_state_dict: bool = False):
    """Quantize ``model`` (possibly created on meta) with bitsandbytes and load / dispatch its weights."""
    _require_bnb(bnb_quantization_config)
    from ..big_modeling import dispatch_model, init_empty_weights
    from .modeling import get_balanced_memory, infer_auto_device_map, load_checkpoint_in_model

    load_in_4bit = bnb_quantization_config.load_in_4bit
    load_in_8bit = bnb_quantization_config.load_in_8bit
    if bnb_quantization_config.skip_modules is None:
        bnb_quantization_config.skip_modules = get_keys_to_not_convert(model)
    if load_in_4bit and isinstance(device_map, dict) and len(set(device_map.values()) & {"cpu", "disk"}) > 0:
        bnb_quantization_config.skip_modules.extend([key for key, value in device_map.items() if value in ["disk", "cpu"]])
    modules_to_not_convert = bnb_quantization_config.skip_modules
    if bnb_quantization_config.keep_in_fp32_modules is None:
        bnb_quantization_config.keep_in_fp32_modules = []
    keep_in_fp32_modules = bnb_quantization_config.keep_in_fp32_modules
    modules_to_not_convert.extend(keep_in_fp32_modules)
    model.is_loaded_in_4bit = load_in_4bit
    model.is_loaded_in_8bit = load_in_8bit
    model_device = next(model.parameters()).device if any(True for _ in model.parameters()) else torch.device("cpu")
    if model_device.type != "meta":
        logger.warning("It is not recommended to quantize a loaded model. The model should be instantiated under the `init_empty_weights` context manager.")
        model, _ = replace_with_bnb_layers(model, bnb_quantization_config, modules_to_not_convert=modules_to_not_convert)
        dtype = bnb_quantization_config.torch_dtype
        for name, param in model.state_dict().items():
            if any(module_to_keep_in_fp32 in name for module_to_keep_in_fp32 in keep_in_fp32_modules):
                param.to(torch.float32)
                if param.dtype != torch.float32:
                    name = name.replace(".weight", "").replace(".bias", "")
                    param = getattr(model, name, None)
                    if param is not None:
                        param.to(torch.float32)
            elif torch.is_floating_point(param):
                param.to(dtype)
        if model_device.type == "cuda":
            model.cuda(torch.cuda.current_device())
            torch.cuda.empty_cache()
        elif torch.cuda.is_available():
            model.to(torch.cuda.current_device())
        else:
            raise RuntimeError("No GPU found. A GPU is needed for quantization.")
        logger.info(f"The model device type is {model_device.type}. However, gpu is needed for quantization. We move the model to gpu.")
        return model
    if weights_location is None:
        raise RuntimeError(f"`weights_location` needs to be the folder path containing the weights of the model, but we found {weights_location} ")
    with init_empty_weights():
        model, _ = replace_with_bnb_layers(model, bnb_quantization_config, modules_to_not_convert=modules_to_not_convert)
    if isinstance(device_map, str):
        if device_map not in ["auto", "balanced", "balanced_low_0", "sequential"]:
            raise ValueError("If passing a string for `device_map`, please choose 'auto', 'balanced', 'balanced_low_0' or 'sequential'.")
        special_dtypes = {name: bnb_quantization_config.torch_dtype for name, _ in model.named_parameters()
                          if any(m in name for m in bnb_quantization_config.skip_modules)}
        special_dtypes.update({name: torch.float32 for name, _ in model.named_parameters()
                               if any(m in name for m in bnb_quantization_config.keep_in_fp32_modules)})
        kwargs = {"special_dtypes": special_dtypes, "no_split_module_classes": no_split_module_classes, "dtype": bnb_quantization_config.target_dtype}
        if device_map != "sequential":
            max_memory = get_balanced_memory(model, low_zero=(device_map == "balanced_low_0"), max_memory=max_memory, **kwargs)
        kwargs["max_memory"] = max_memory
        device_map = infer_auto_device_map(model, **kwargs)
    if device_map is None:
        device_map = {"": torch.cuda.current_device()}
    load_checkpoint_in_model(model, weights_location, device_map, dtype=bnb_quantization_config.torch_dtype, offload_folder=offload_folder,
                             offload_state_dict=offload_state_dict, keep_in_fp32_modules=bnb_quantization_config.keep_in_fp32_modules,
                             offload_8bit_bnb=load_in_8bit and offload_folder is not None)
    return dispatch_model(model, device_map=device_map, offload_dir=offload_folder)
