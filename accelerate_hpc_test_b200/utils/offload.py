"""Disk offload store for big-model inference: raw memory-mapped arrays plus a JSON index.

On-disk format (kept identical to the reference's ``utils/offload.py:25-213`` so an offload folder written by
either side can be read by the other):

* ``<folder>/<name>.dat`` — the tensor's bytes, C order, native endianness, no header;
* ``<folder>/index.json`` — ``{name: {"dtype": <numpy dtype name | "bfloat16">, "shape": [...]}}``;
* bfloat16 has no numpy dtype: its bit pattern is stored as ``int16`` and the index says ``"bfloat16"``;
* 0-d tensors are stored as one element and the index keeps ``"shape": []``;
* an index entry may instead point into a safetensors file (``safetensors_file``, optional ``weight_name`` and
  ``dtype``): such weights are read straight from that file and never copied into the folder.

``OffloadedWeightsLoader`` presents "tensors that are in RAM" ∪ "tensors in the folder" ∪ "tensors in safetensors
files" as one read-only mapping; ``AlignDevicesHook`` pulls weights out of it right before a module runs.
"""

from __future__ import annotations

import json
import os
from collections.abc import Mapping
from typing import Iterable, Optional

import numpy as np
import torch

INDEX_FILE = "index.json"
_BF16 = "bfloat16"


# --------------------------------------------------------------------------- one tensor <-> one .dat file
def _storage_view(tensor: torch.Tensor):
    """(numpy array holding the bytes, dtype name for the index)."""
    if tensor.dtype == torch.bfloat16:
        return tensor.detach().cpu().contiguous().view(torch.int16).numpy(), _BF16
    arr = tensor.detach().cpu().contiguous().numpy()
    return arr, str(arr.dtype)


def _dat_path(folder, name: str) -> str:
    return os.path.join(folder, f"{name}.dat")


def offload_weight(weight: torch.Tensor, weight_name: str, offload_folder, index: Optional[dict] = None):
    """Write one tensor; when ``index`` is given, add its entry and return the index."""
    arr, dtype_name = _storage_view(weight)
    if index is not None:
        index[weight_name] = {"dtype": dtype_name, "shape": list(arr.shape)}
    flat_shape = arr.shape if arr.ndim else (1,)
    mm = np.memmap(_dat_path(offload_folder, weight_name), dtype=arr.dtype, mode="w+", shape=flat_shape)
    mm[...] = arr.reshape(flat_shape)
    mm.flush()
    del mm
    return index


def load_offloaded_weight(weight_file, weight_info: dict) -> torch.Tensor:
    """Read one ``.dat`` file back into a CPU tensor (a copy, not a view of the file)."""
    logical_shape = tuple(weight_info["shape"])
    is_bf16 = weight_info["dtype"] == _BF16
    mm = np.memmap(weight_file, dtype="int16" if is_bf16 else weight_info["dtype"], mode="r", shape=logical_shape or (1,))
    out = torch.tensor(np.asarray(mm if logical_shape else mm[0]))
    return out.view(torch.bfloat16) if is_bf16 else out


# --------------------------------------------------------------------------- folders
def save_offload_index(index: Optional[dict], offload_folder) -> None:
    """Merge ``index`` into the folder's ``index.json`` (entries already there are kept unless overwritten)."""
    if not index:
        return
    path = os.path.join(offload_folder, INDEX_FILE)
    merged = {}
    if os.path.isfile(path):
        with open(path, encoding="utf-8") as fh:
            merged = json.load(fh)
    merged.update(index)
    with open(path, "w", encoding="utf-8") as fh:
        json.dump(merged, fh, indent=2)


def offload_state_dict(save_dir, state_dict: Mapping) -> None:
    os.makedirs(save_dir, exist_ok=True)
    index: dict = {}
    for name, tensor in state_dict.items():
        offload_weight(tensor, name, save_dir, index)
    save_offload_index(index, save_dir)


def _read_index(folder) -> dict:
    with open(os.path.join(folder, INDEX_FILE), encoding="utf-8") as fh:
        return json.load(fh)


# --------------------------------------------------------------------------- mappings
class PrefixedDataset(Mapping):
    """``PrefixedDataset(d, "encoder.")["w"]`` is ``d["encoder.w"]``; iteration yields the (full) keys under the prefix."""

    def __init__(self, dataset: Mapping, prefix: str):
        self.dataset = dataset
        self.prefix = prefix

    def __getitem__(self, key):
        return self.dataset[self.prefix + key]

    def __iter__(self):
        return (k for k in self.dataset if k.startswith(self.prefix))

    def __len__(self):
        return len(self.dataset)


def _read_safetensors_entry(entry: dict, key: str, device) -> torch.Tensor:
    from safetensors import safe_open

    target = torch.device("cpu" if device is None else device)
    name = entry.get("weight_name", key)

    def fetch(dev: str):
        with safe_open(entry["safetensors_file"], framework="pt", device=dev) as fh:
            return fh.get_tensor(name)

    try:
        tensor = fetch(str(target))
    except TypeError:  # older safetensors reject some device spellings: read on CPU and move
        tensor = fetch("cpu")
    if "dtype" in entry:
        tensor = tensor.to(getattr(torch, entry["dtype"]))
    return tensor if tensor.device == target else tensor.to(target)


class OffloadedWeightsLoader(Mapping):
    """Read-only ``name -> tensor`` view over RAM-resident tensors and a disk index (see module docstring)."""

    def __init__(self, state_dict: Optional[Mapping] = None, save_folder=None, index: Optional[Mapping] = None, device=None):
        if state_dict is None and save_folder is None and index is None:
            raise ValueError("Need either a `state_dict`, a `save_folder` or an `index` containing offloaded weights.")
        self.state_dict = {} if state_dict is None else state_dict
        self.save_folder = save_folder
        if index is None:
            index = _read_index(save_folder) if save_folder is not None else {}
        self.index = index
        self.device = device
        seen = dict.fromkeys(self.state_dict)
        seen.update(dict.fromkeys(self.index))
        self.all_keys = list(seen)

    def __getitem__(self, key: str) -> torch.Tensor:
        if key in self.state_dict:
            return self.state_dict[key]
        entry = self.index[key]
        if entry.get("safetensors_file") is not None:
            return _read_safetensors_entry(entry, key, self.device)
        return load_offloaded_weight(_dat_path(self.save_folder, key), entry)

    def __iter__(self):
        return iter(self.all_keys)

    def __len__(self):
        return len(self.all_keys)


def extract_submodules_state_dict(state_dict: Mapping, submodule_names: Iterable[str]) -> dict:
    """Entries that belong to one of the named sub-modules (the module's own key or anything below ``name.``)."""
    picked = {}
    for module_name in submodule_names:
        below = module_name + "."
        for key, value in state_dict.items():
            if key == module_name or key.startswith(below):
                picked[key] = value
    return picked
