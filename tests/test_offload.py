"""Disk offload store (behaviour ports of the reference's tests/test_offload.py)."""

import json
import os

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200.utils.offload import (
    OffloadedWeightsLoader,
    PrefixedDataset,
    extract_submodules_state_dict,
    load_offloaded_weight,
    offload_state_dict,
    offload_weight,
    save_offload_index,
)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16, torch.int64, torch.int8, torch.bool])
def test_offload_weight_round_trip_per_dtype(tmp_path, dtype):
    w = (torch.randn(2, 3) * 10).to(dtype)
    index = offload_weight(w, "w", str(tmp_path), {})
    assert os.path.isfile(tmp_path / "w.dat")
    assert index == {"w": {"shape": [2, 3], "dtype": "bfloat16" if dtype == torch.bfloat16 else str(dtype).split(".")[1]}}
    back = load_offloaded_weight(str(tmp_path / "w.dat"), index["w"])
    assert back.dtype == dtype and torch.equal(back, w)


def test_scalar_tensor_keeps_empty_shape(tmp_path):
    w = torch.tensor(3.5)
    index = offload_weight(w, "s", str(tmp_path), {})
    assert index["s"]["shape"] == []
    back = load_offloaded_weight(str(tmp_path / "s.dat"), index["s"])
    assert back.shape == () and back.item() == 3.5


def test_offload_weight_without_index_returns_none(tmp_path):
    assert offload_weight(torch.ones(2), "x", str(tmp_path)) is None
    assert os.path.getsize(tmp_path / "x.dat") == 8


def test_offload_state_dict_writes_index_and_merges(tmp_path):
    model = nn.Sequential(nn.Linear(3, 4), nn.BatchNorm1d(4))
    offload_state_dict(str(tmp_path), model.state_dict())
    index = json.load(open(tmp_path / "index.json"))
    assert set(index) == set(model.state_dict()) and index["0.weight"] == {"dtype": "float32", "shape": [4, 3]}
    save_offload_index({"extra": {"dtype": "float32", "shape": [1]}}, str(tmp_path))
    merged = json.load(open(tmp_path / "index.json"))
    assert "extra" in merged and "0.weight" in merged
    save_offload_index({}, str(tmp_path))  # no-op
    assert json.load(open(tmp_path / "index.json")) == merged


def test_offloaded_weights_loader_merges_ram_and_disk(tmp_path):
    model = nn.Sequential(nn.Linear(3, 4), nn.Linear(4, 2))
    sd = model.state_dict()
    in_ram = {k: v for k, v in sd.items() if k.startswith("0.")}
    on_disk = {k: v for k, v in sd.items() if k.startswith("1.")}
    offload_state_dict(str(tmp_path), on_disk)
    loader = OffloadedWeightsLoader(state_dict=in_ram, save_folder=str(tmp_path))
    assert sorted(loader) == sorted(sd) and len(loader) == 4
    for k in sd:
        assert torch.equal(loader[k], sd[k])
    # a key present in both places is served from RAM and listed once
    offload_state_dict(str(tmp_path), sd)
    loader = OffloadedWeightsLoader(state_dict={"0.weight": torch.zeros(4, 3)}, save_folder=str(tmp_path))
    assert list(loader).count("0.weight") == 1 and loader["0.weight"].abs().sum() == 0
    with pytest.raises(ValueError):
        OffloadedWeightsLoader()


def test_offloaded_weights_loader_reads_safetensors_entries(tmp_path):
    from safetensors.torch import save_file

    t = {"a": torch.arange(6, dtype=torch.float32).view(2, 3), "b": torch.ones(2)}
    path = str(tmp_path / "w.safetensors")
    save_file(t, path)
    index = {"a": {"safetensors_file": path}, "renamed": {"safetensors_file": path, "weight_name": "b", "dtype": "float16"}}
    loader = OffloadedWeightsLoader(index=index, device="cpu")
    assert torch.equal(loader["a"], t["a"]) and loader["renamed"].dtype == torch.float16


def test_prefixed_dataset():
    data = {"x.a": 1, "x.b": 2, "y.a": 3}
    view = PrefixedDataset(data, "x.")
    assert view["a"] == 1 and view["b"] == 2 and list(view) == ["x.a", "x.b"] and len(view) == 3
    with pytest.raises(KeyError):
        view["c"]


def test_extract_submodules_state_dict_matches_whole_names_only():
    sd = {"a.1": 0, "a.10": 1, "a.2": 2, "a.1.w": 3, "b": 4}
    assert extract_submodules_state_dict(sd, ["a.1", "a.2"]) == {"a.1": 0, "a.2": 2, "a.1.w": 3}
    assert extract_submodules_state_dict(sd, ["b"]) == {"b": 4} and extract_submodules_state_dict(sd, []) == {}
