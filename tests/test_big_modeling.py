"""Meta init, device-map inference, offload hooks and stores (modelled on reference tests/test_big_modeling.py,
tests/test_modeling_utils.py, tests/test_hooks.py, tests/test_offload.py)."""

import os

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200.big_modeling import (
    cpu_offload,
    cpu_offload_with_hook,
    disk_offload,
    dispatch_model,
    init_empty_weights,
    init_on_device,
    load_checkpoint_and_dispatch,
)
from accelerate_hpc_test_b200.hooks import AlignDevicesHook, ModelHook, SequentialHook, add_hook_to_module, remove_hook_from_module
from accelerate_hpc_test_b200.utils.modeling import (
    clean_device_map,
    compute_module_sizes,
    convert_file_size_to_int,
    dtype_byte_size,
    find_tied_parameters,
    infer_auto_device_map,
    load_checkpoint_in_model,
    named_module_tensors,
    retie_parameters,
    set_module_tensor_to_device,
    shard_checkpoint,
)
from accelerate_hpc_test_b200.utils.offload import OffloadedWeightsLoader, load_offloaded_weight, offload_state_dict, offload_weight


class ModelForTest(nn.Module):
    def __init__(self):
        super().__init__()
        self.linear1 = nn.Linear(3, 4)
        self.batchnorm = nn.BatchNorm1d(4)
        self.linear2 = nn.Linear(4, 5)

    def forward(self, x):
        return self.linear2(self.batchnorm(self.linear1(x)))


class BiggerModelForTest(nn.Module):
    def __init__(self):
        super().__init__()
        self.linear1 = nn.Linear(3, 4)
        self.linear2 = nn.Linear(4, 5)
        self.batchnorm = nn.BatchNorm1d(5)
        self.linear3 = nn.Linear(5, 6)
        self.linear4 = nn.Linear(6, 5)

    def forward(self, x):
        return self.linear4(self.linear3(self.batchnorm(self.linear2(self.linear1(x)))))


def test_init_empty_weights_and_on_device():
    with init_empty_weights():
        m = nn.Linear(4, 5)
        b = nn.BatchNorm1d(4)
    assert m.weight.device == torch.device("meta") and b.running_mean.device == torch.device("cpu")
    with init_empty_weights(include_buffers=True):
        b = nn.BatchNorm1d(4)
    assert b.running_mean.device == torch.device("meta")
    with init_on_device(torch.device("cpu")):
        m = nn.Linear(2, 2)
    assert m.weight.device == torch.device("cpu")
    # the patch is undone afterwards
    assert nn.Linear(2, 2).weight.device == torch.device("cpu")


def test_set_module_tensor_to_device_and_named_tensors():
    model = ModelForTest()
    set_module_tensor_to_device(model, "linear1.weight", "meta")
    assert model.linear1.weight.device == torch.device("meta")
    with pytest.raises(ValueError):
        set_module_tensor_to_device(model, "linear1.weight", "cpu")
    set_module_tensor_to_device(model, "linear1.weight", "cpu", value=torch.randn(4, 3))
    assert model.linear1.weight.device == torch.device("cpu") and isinstance(model.linear1.weight, nn.Parameter)
    set_module_tensor_to_device(model, "batchnorm.running_mean", "meta")
    assert model.batchnorm.running_mean.device == torch.device("meta")
    names = [n for n, _ in named_module_tensors(model, recurse=True)]
    assert "linear1.weight" in names and "batchnorm.running_mean" in names
    assert "batchnorm.running_mean" not in [n for n, _ in named_module_tensors(model, include_buffers=False, recurse=True)]


def test_sizes_and_units():
    assert convert_file_size_to_int("1KB") == 1000 and convert_file_size_to_int("1KiB") == 1024 and convert_file_size_to_int("8Mb") == 10**6
    assert dtype_byte_size(torch.float32) == 4 and dtype_byte_size(torch.bfloat16) == 2 and dtype_byte_size(torch.bool) == 1 / 8
    sizes = compute_module_sizes(ModelForTest())
    assert sizes["linear1"] == (12 + 4) * 4 and sizes["linear2"] == (20 + 5) * 4
    assert sizes[""] == sizes["linear1"] + sizes["linear2"] + sizes["batchnorm"]
    assert compute_module_sizes(ModelForTest(), dtype=torch.float16)["linear1"] == (12 + 4) * 2


def test_tied_parameters():
    model = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 4))
    assert find_tied_parameters(model) == []
    model[1].weight = model[0].weight
    assert find_tied_parameters(model) == [["0.weight", "1.weight"]]
    model[1].weight = nn.Parameter(model[0].weight.detach().clone())
    retie_parameters(model, [["0.weight", "1.weight"]])
    assert model[1].weight is model[0].weight


def test_infer_auto_device_map_packs_in_order():
    model = ModelForTest()
    # sizes: linear1 64, batchnorm 72, linear2 100
    # only linear1 fits on device 0: room for the largest layer (100) is kept free there
    assert infer_auto_device_map(model, max_memory={0: 200, 1: 200}) == {"linear1": 0, "batchnorm": 1, "linear2": 1}
    dm = infer_auto_device_map(model, max_memory={0: 200, "cpu": 72, "disk": 10**6})
    assert dm["linear1"] == 0 and set(dm.values()) <= {0, "cpu", "disk"} and len(dm) >= 2
    # everything fits on one device -> a single entry after cleaning
    assert infer_auto_device_map(model, max_memory={0: 10**6}) == {"": 0}


def test_infer_auto_device_map_respects_no_split_and_tied_weights():
    model = nn.Sequential(ModelForTest(), ModelForTest(), ModelForTest())
    dm = infer_auto_device_map(model, max_memory={0: 500, 1: 300, 2: 300}, no_split_module_classes=["ModelForTest"])
    assert dm == {"0": 0, "1": 1, "2": 2}
    tied = nn.Sequential()
    tied.add_module("layer1", ModelForTest())
    tied.add_module("layer2", ModelForTest())
    tied.add_module("layer3", ModelForTest())
    tied.layer3.linear2.weight = tied.layer1.linear2.weight
    dm = infer_auto_device_map(tied, max_memory={0: 400, 1: 500})
    flat = {}
    for k, v in dm.items():
        flat[k] = v
    # the tied pair ends up on one device
    def dev(name):
        while name not in dm:
            name = name.rsplit(".", 1)[0] if "." in name else ""
        return dm[name]
    assert dev("layer1.linear2") == dev("layer3.linear2")


def test_clean_device_map():
    assert clean_device_map({"a.x": 0, "a.y": 0, "b": 1}) == {"a": 0, "b": 1}
    assert clean_device_map({"a": 0, "b": 0}) == {"": 0}


def test_hooks_protocol():
    class Pre(ModelHook):
        def pre_forward(self, module, *args, **kwargs):
            return (args[0] + 1,) + args[1:], kwargs

    class Post(ModelHook):
        def post_forward(self, module, output):
            return output + 1

    model = ModelForTest().eval()
    x = torch.randn(2, 3)
    ref = model(x)
    add_hook_to_module(model, Pre())
    assert torch.allclose(model(x), ModelForTest.forward(model, x + 1) if False else model._old_forward(x + 1))
    add_hook_to_module(model, Post(), append=True)
    assert isinstance(model._hf_hook, SequentialHook)
    assert torch.allclose(model(x), model._old_forward(x + 1) + 1)
    remove_hook_from_module(model)
    assert not hasattr(model, "_hf_hook") and torch.allclose(model(x), ref)


def test_align_devices_hook_offload_roundtrip():
    model = ModelForTest().eval()
    x = torch.randn(2, 3)
    ref = model(x)
    hook = AlignDevicesHook(execution_device="cpu", offload=True)
    add_hook_to_module(model.linear1, hook)
    assert model.linear1.weight.device == torch.device("meta")
    assert torch.allclose(model(x), ref)
    assert model.linear1.weight.device == torch.device("meta")
    remove_hook_from_module(model.linear1)
    assert model.linear1.weight.device == torch.device("cpu")


def test_cpu_and_disk_offload_equal_plain_forward(tmp_path):
    model = ModelForTest().eval()
    x = torch.randn(2, 3)
    ref = model(x)
    cpu_offload(model, execution_device=torch.device("cpu"))
    assert all(p.device == torch.device("meta") for p in model.parameters())
    assert torch.allclose(model(x), ref)
    model2 = ModelForTest().eval()
    model2.load_state_dict({k: v for k, v in ModelForTest().state_dict().items()})
    ref2 = model2(x)
    disk_offload(model2, str(tmp_path), execution_device=torch.device("cpu"))
    assert os.path.isfile(tmp_path / "index.json")
    assert torch.allclose(model2(x), ref2)


def test_cpu_offload_with_hook_chain():
    m1, m2 = nn.Linear(4, 4), nn.Linear(4, 4)
    m1, h1 = cpu_offload_with_hook(m1, execution_device=torch.device("cpu"))
    m2, h2 = cpu_offload_with_hook(m2, execution_device=torch.device("cpu"), prev_module_hook=h1)
    x = torch.randn(2, 4)
    assert m2(m1(x)).shape == (2, 4)
    h2.offload()
    h1.remove()
    assert not hasattr(m1, "_hf_hook")


def test_dispatch_model_cpu_disk(tmp_path):
    model = BiggerModelForTest().eval()
    x = torch.randn(2, 3)
    ref = model(x)
    dm = {"linear1": "cpu", "linear2": "cpu", "batchnorm": "disk", "linear3": "disk", "linear4": "cpu"}
    dispatch_model(model, dm, offload_dir=str(tmp_path))
    assert model.hf_device_map == dm
    assert torch.allclose(model(x), ref, atol=1e-6)
    with pytest.raises(RuntimeError):
        model.to("cpu")


def test_offload_store_roundtrip(tmp_path):
    for dtype in (torch.float32, torch.float16, torch.bfloat16):
        w = torch.randn(2, 3, dtype=dtype)
        index = offload_weight(w, "w", str(tmp_path), {})
        assert index["w"]["shape"] == [2, 3] and index["w"]["dtype"] == str(dtype).split(".")[1]
        back = load_offloaded_weight(str(tmp_path / "w.dat"), index["w"])
        assert torch.equal(back, w) and back.dtype == dtype
    sd = ModelForTest().state_dict()
    offload_state_dict(str(tmp_path / "sd"), sd)
    loader = OffloadedWeightsLoader(state_dict={"extra": torch.ones(1)}, save_folder=str(tmp_path / "sd"))
    assert sorted(loader) == sorted(["extra"] + list(sd.keys()))
    assert torch.allclose(loader["linear1.weight"], sd["linear1.weight"])


def test_shard_checkpoint_and_load_checkpoint_in_model(tmp_path):
    from safetensors.torch import save_file

    model = BiggerModelForTest()
    sd = model.state_dict()
    shards, index = shard_checkpoint(sd, max_shard_size=200, weights_name="model.safetensors")
    assert index is not None and len(shards) > 1 and set(index["weight_map"]) == set(sd)
    for fname, shard in shards.items():
        save_file({k: v.contiguous() for k, v in shard.items()}, str(tmp_path / fname), metadata={"format": "pt"})
    import json

    with open(tmp_path / "model.safetensors.index.json", "w") as f:
        json.dump(index, f)
    with init_empty_weights():
        fresh = BiggerModelForTest()
    dm = {"linear1": "cpu", "linear2": "cpu", "batchnorm": "cpu", "linear3": "disk", "linear4": "cpu"}
    load_checkpoint_in_model(fresh, str(tmp_path), device_map=dm, offload_folder=str(tmp_path / "off"))
    assert fresh.linear3.weight.device == torch.device("meta") and torch.equal(fresh.linear1.weight, sd["linear1.weight"])
    assert os.path.isfile(tmp_path / "off" / "linear3.weight.dat")
    with init_empty_weights():
        fresh2 = BiggerModelForTest()
    x = torch.randn(2, 3)
    model.eval()
    out = load_checkpoint_and_dispatch(fresh2, str(tmp_path), device_map=dm, offload_folder=str(tmp_path / "off2")).eval()(x)
    assert torch.allclose(out, model(x), atol=1e-6)


def test_context_parallel_hooks_force_causal_attention_without_a_mask():
    from accelerate_hpc_test_b200.big_modeling import _attach_context_parallel_hooks

    seen = {}

    class Attention(nn.Module):
        def forward(self, x, attention_mask=None, is_causal=False):
            seen.update(mask=attention_mask, causal=is_causal)
            return x

    class Layer(nn.Module):
        def __init__(self):
            super().__init__()
            self.self_attn = Attention()
            self.other = Attention()

    layer = Layer()
    assert _attach_context_parallel_hooks(layer) is None
    layer.self_attn(torch.zeros(1), attention_mask=torch.ones(1, 4))
    assert seen == {"mask": None, "causal": True}
    layer.self_attn(torch.zeros(1))                                     # no mask passed: the call is left alone
    assert seen == {"mask": None, "causal": False}
    layer.other(torch.zeros(1), attention_mask=torch.ones(1, 4))        # only `*.self_attn` modules are touched
    assert seen["mask"] is not None and seen["causal"] is False


def test_layerwise_casting_stores_small_and_computes_big():
    from accelerate_hpc_test_b200.big_modeling import attach_layerwise_casting_hooks

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.embed = nn.Embedding(10, 8)
            self.proj = nn.Linear(8, 8)
            self.norm = nn.LayerNorm(8)
            self.head = nn.Linear(8, 4)

        def forward(self, ids):
            return self.head(self.norm(self.proj(self.embed(ids))))

    torch.manual_seed(0)
    net = Net()
    reference = net(torch.arange(6))
    attach_layerwise_casting_hooks(net, storage_dtype=torch.bfloat16, compute_dtype=torch.float32, skip_modules_pattern=["head"])
    # linear layers rest in the storage dtype; the skipped head, the embedding and the norm are untouched
    assert net.proj.weight.dtype == torch.bfloat16 and net.head.weight.dtype == torch.float32
    assert net.embed.weight.dtype == torch.float32 and net.norm.weight.dtype == torch.float32
    out = net(torch.arange(6))
    assert out.dtype == torch.float32 and net.proj.weight.dtype == torch.bfloat16        # computed big, back to small afterwards
    assert torch.allclose(out, reference, atol=5e-2) and not torch.equal(out, reference)  # only the rounding of the stored weights differs
    single = Net()
    attach_layerwise_casting_hooks(single, torch.bfloat16, torch.float32, skip_modules_pattern="^head$")       # one expression as a string
    assert single.proj.weight.dtype == torch.bfloat16 and single.head.weight.dtype == torch.float32
    # whole classes can be skipped too
    other = Net()
    attach_layerwise_casting_hooks(other, torch.bfloat16, torch.float32, skip_modules_classes=(nn.Linear,))
    assert other.proj.weight.dtype == torch.float32
