"""utils/modeling.py: sizes, placement, ties, budgets, device maps, checkpoint streaming (behaviour of reference
tests/test_modeling_utils.py, on CPU / meta / disk)."""

import json
import os
from collections import OrderedDict

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200.big_modeling import cpu_offload, init_empty_weights
from accelerate_hpc_test_b200.utils.dataclasses import CustomDtype
from accelerate_hpc_test_b200.utils.modeling import (
    align_module_device,
    calculate_maximum_sizes,
    check_device_map,
    check_device_same,
    check_tied_parameters_on_same_device,
    clean_device_map,
    compute_module_sizes,
    compute_module_total_buffer_size,
    convert_file_size_to_int,
    dtype_byte_size,
    find_tied_parameters,
    get_balanced_memory,
    get_max_layer_size,
    get_max_memory,
    get_module_leaves,
    get_non_persistent_buffers,
    get_state_dict_offloaded_model,
    has_offloaded_params,
    id_tensor_storage,
    infer_auto_device_map,
    load_checkpoint_in_model,
    load_state_dict,
    named_module_tensors,
    retie_parameters,
    set_module_tensor_to_device,
    shard_checkpoint,
)


class Block(nn.Module):
    def __init__(self, width=4):
        super().__init__()
        self.linear1 = nn.Linear(3, width)
        self.batchnorm = nn.BatchNorm1d(width)
        self.linear2 = nn.Linear(width, 5)

    def forward(self, x):
        return self.linear2(self.batchnorm(self.linear1(x)))


class Nested(nn.Module):
    def __init__(self):
        super().__init__()
        self.first = Block()
        self.second = Block()
        self.head = nn.Linear(5, 2)


@pytest.mark.parametrize("text,expected", [("5GB", 5 * 10**9), ("2MiB", 2 * 2**20), ("1.5KB", 1500), ("8Gb", 10**9), ("3 kib", 3 * 1024 // 8),
                                           ("10GIB", 10 * 2**30), (123, 123)])
def test_convert_file_size_to_int(text, expected):
    assert convert_file_size_to_int(text) == expected


@pytest.mark.parametrize("bad", ["5", "GB", "5TB!", "five MB"])
def test_convert_file_size_rejects_garbage(bad):
    with pytest.raises(ValueError, match="valid format"):
        convert_file_size_to_int(bad)


def test_dtype_byte_size_covers_torch_and_custom_dtypes():
    assert dtype_byte_size(torch.float32) == 4 and dtype_byte_size(torch.bfloat16) == 2 and dtype_byte_size(torch.int64) == 8
    assert dtype_byte_size(torch.uint8) == 1 and dtype_byte_size(torch.bool) == 1 / 8
    assert dtype_byte_size(torch.float8_e4m3fn) == 1 and dtype_byte_size(torch.complex64) == 8
    assert dtype_byte_size(CustomDtype.FP8) == 1 and dtype_byte_size(CustomDtype.INT4) == 0.5 and dtype_byte_size(CustomDtype.INT2) == 0.25


def test_check_device_same_and_storage_identity():
    assert check_device_same(torch.device("cpu"), torch.device("cpu"))
    assert check_device_same(torch.device("cuda"), torch.device("cuda", 0))
    assert not check_device_same(torch.device("cuda", 1), torch.device("cuda"))
    assert not check_device_same(torch.device("cpu"), torch.device("cuda"))
    base = torch.zeros(8)
    assert id_tensor_storage(base[:4]) == id_tensor_storage(base[4:])
    assert id_tensor_storage(base) != id_tensor_storage(torch.zeros(8))
    assert id_tensor_storage(torch.empty(3, device="meta"))[0] == torch.device("meta")


def test_named_module_tensors_and_non_persistent_buffers():
    model = Block()
    model.register_buffer("scratch", torch.zeros(2), persistent=False)
    model.batchnorm.register_buffer("tmp", torch.zeros(1), persistent=False)
    own = [n for n, _ in named_module_tensors(model)]
    assert own == ["scratch"]
    every = [n for n, _ in named_module_tensors(model, recurse=True)]
    assert every[:2] == ["linear1.weight", "linear1.bias"] and "batchnorm.running_mean" in every and "scratch" in every
    saved = [n for n, _ in named_module_tensors(model, recurse=True, remove_non_persistent=True)]
    assert "scratch" not in saved and "batchnorm.running_mean" in saved
    assert [n for n, _ in named_module_tensors(model, include_buffers=False, recurse=True)] == [n for n, _ in model.named_parameters()]
    assert get_non_persistent_buffers(model) == {"scratch"}
    assert get_non_persistent_buffers(model, recurse=True, fqns=True) == {"scratch", "batchnorm.tmp"}


def test_compute_module_sizes_with_target_and_special_dtypes():
    model = Block()
    sizes = compute_module_sizes(model)
    assert sizes["linear1"] == (3 * 4 + 4) * 4 and sizes["linear1.weight"] == 48
    assert sizes["batchnorm"] == 4 * 4 * 4 + 8  # weight, bias, running mean, running var, and an int64 counter
    assert sizes[""] == sizes["linear1"] + sizes["batchnorm"] + sizes["linear2"]
    half = compute_module_sizes(model, dtype=torch.float16)
    assert half["linear1"] == sizes["linear1"] // 2 and half["batchnorm"] == 4 * 4 * 2 + 8  # the integer counter is not cast
    assert compute_module_sizes(model, dtype="torch.float16")["linear2"] == half["linear2"]
    assert compute_module_sizes(model.double(), dtype=torch.float32)["linear1"] == sizes["linear1"]
    special = compute_module_sizes(Block(), dtype=torch.float16, special_dtypes={"linear1.weight": torch.float32})
    assert special["linear1"] == 48 + 4 * 2
    int4 = compute_module_sizes(Block(), special_dtypes={"linear2.weight": CustomDtype.INT4})
    assert int4["linear2"] == 20 * 0.5 + 5 * 4
    assert compute_module_total_buffer_size(Block()) == 4 * 4 * 2 + 8


def test_max_layer_size_respects_no_split_classes():
    model = Nested()
    sizes = compute_module_sizes(model)
    top = list(model.named_children())
    size, names = get_max_layer_size(top, sizes, [])
    assert size == sizes["first.linear2"] == 100 and names == ["first.linear2", "second.linear2"]
    size, names = get_max_layer_size(top, sizes, ["Block"])
    assert size == sizes["first"] and names == ["first", "second"]
    model._no_split_modules = ["Block"]
    total, (largest, holders) = calculate_maximum_sizes(model)
    assert total == sizes[""] and largest == sizes["first"] and holders == ["first", "second"]


def test_get_max_memory_parses_orders_and_validates():
    budget = get_max_memory({"cpu": "1GB", 1: "2GiB", 0: 100, "disk": "5MB"})
    assert list(budget) == [0, 1, "cpu", "disk"] and budget[1] == 2 * 2**30 and budget["cpu"] == 10**9 and budget["disk"] == 5 * 10**6
    with pytest.raises(ValueError, match="not recognized"):
        get_max_memory({"tpu": 10})
    auto = get_max_memory()
    assert auto["cpu"] > 0


def test_get_balanced_memory_without_gpus_is_the_plain_budget():
    assert get_balanced_memory(Block(), max_memory={"cpu": "1GB"}) == {"cpu": 10**9}


def test_module_leaves():
    sizes = compute_module_sizes(Nested())
    assert get_module_leaves(sizes) == ["", "first.batchnorm.running_mean", "first.batchnorm.running_var", "first.batchnorm.num_batches_tracked",
                                        "second.batchnorm.running_mean", "second.batchnorm.running_var", "second.batchnorm.num_batches_tracked"]
    assert get_module_leaves({"a": 1, "a.b": 1, "a.b.weight": 1, "a.c": 1, "d": 1}) == ["a.c", "d"]


def test_clean_device_map_folds_subtrees_bottom_up():
    flat = {"first.linear1": 0, "first.batchnorm": 0, "first.linear2": 0, "second.linear1": 0, "second.batchnorm": 1, "second.linear2": 1, "head": 1}
    assert clean_device_map(dict(flat)) == {"first": 0, "second.linear1": 0, "second.batchnorm": 1, "second.linear2": 1, "head": 1}
    assert clean_device_map({"a.b.c": 0, "a.b.d": 0, "a.e": 0, "f": 0}) == {"": 0}
    assert clean_device_map({"a.b.c": 0, "a.b.d": 0, "a.e": 1}) == {"a.b": 0, "a.e": 1}
    assert clean_device_map({"a.b.c": "cpu"}) == {"a.b.c": "cpu"}  # a single entry is left where it is
    deep = {"a.x.p": 0, "a.x.q": 0, "a.y": 1, "b": 0}
    assert clean_device_map(dict(deep), module_name="a") == {"a.x": 0, "a.y": 1, "b": 0}


def test_check_device_map():
    model = Nested()
    check_device_map(model, {"": "cpu"})
    check_device_map(model, {"first": 0, "second": "cpu", "head": "disk"})
    with pytest.raises(ValueError, match="head.weight, head.bias"):
        check_device_map(model, {"first": 0, "second": "cpu"})
    with pytest.warns(UserWarning, match="do not match any submodules"):
        check_device_map(model, {"first": 0, "second": 0, "head": 0, "tail": 0})


def test_tied_parameters_found_checked_and_retied():
    model = Nested()
    assert find_tied_parameters(model) == []
    model.second.linear1.weight = model.first.linear1.weight
    model.second.linear1.bias = model.first.linear1.bias
    groups = find_tied_parameters(model)
    assert groups == [["first.linear1.weight", "second.linear1.weight"], ["first.linear1.bias", "second.linear1.bias"]]
    with pytest.warns(FutureWarning):
        assert groups.values() == ["second.linear1.weight", "second.linear1.bias"]
    # loading breaks the tie...
    model.second.linear1.weight = nn.Parameter(model.first.linear1.weight.detach().clone())
    assert find_tied_parameters(model) == [["first.linear1.bias", "second.linear1.bias"]]
    retie_parameters(model, groups)
    assert model.second.linear1.weight is model.first.linear1.weight
    # ...and a meta member never becomes the keeper
    with init_empty_weights():
        hollow = Nested()
    hollow.second.linear1.weight = nn.Parameter(torch.ones(4, 3))
    retie_parameters(hollow, [["first.linear1.weight", "second.linear1.weight"]])
    assert hollow.first.linear1.weight.device.type == "cpu" and hollow.first.linear1.weight is hollow.second.linear1.weight


def test_check_tied_parameters_on_same_device_warns(caplog):
    groups = [["first.linear1.weight", "second.linear1.weight"]]
    with caplog.at_level("WARNING"):
        check_tied_parameters_on_same_device(groups, {"first": 0, "second": 0, "head": "cpu"})
        assert "different devices" not in caplog.text
        check_tied_parameters_on_same_device(groups, {"first": 0, "second.linear1": "cpu", "second": 0})
        assert "different devices" in caplog.text
    with pytest.raises(ValueError, match="does not contain the module"):
        check_tied_parameters_on_same_device(groups, {"head": 0})


def test_set_module_tensor_keeps_kind_dtype_and_grad_flag():
    model = Block()
    model.linear1.weight.requires_grad_(False)
    set_module_tensor_to_device(model, "linear1.weight", "meta")
    assert model.linear1.weight.device.type == "meta" and not model.linear1.weight.requires_grad and isinstance(model.linear1.weight, nn.Parameter)
    with pytest.raises(ValueError, match="need a `value`"):
        set_module_tensor_to_device(model, "linear1.weight", "cpu")
    set_module_tensor_to_device(model, "linear1.weight", "cpu", value=torch.ones(4, 3, dtype=torch.float64))
    assert model.linear1.weight.dtype == torch.float32 and model.linear1.weight.sum() == 12  # value follows the tensor's dtype
    set_module_tensor_to_device(model, "linear1.weight", "cpu", value=torch.ones(4, 3), dtype=torch.bfloat16)
    assert model.linear1.weight.dtype == torch.bfloat16
    with pytest.raises(ValueError, match="looks incorrect"):
        set_module_tensor_to_device(model, "linear1.weight", "cpu", value=torch.ones(5, 3))
    with pytest.raises(ValueError, match="does not have a parameter or a buffer"):
        set_module_tensor_to_device(model, "linear1.nothing", "cpu")
    # buffers stay buffers; integer buffers are not cast
    set_module_tensor_to_device(model, "batchnorm.num_batches_tracked", "cpu", value=torch.tensor(7), dtype=torch.float16)
    assert model.batchnorm.num_batches_tracked.dtype == torch.int64 and "num_batches_tracked" in model.batchnorm._buffers
    set_module_tensor_to_device(model, "batchnorm.running_mean", "meta", dtype=torch.float16)
    assert model.batchnorm.running_mean.device.type == "meta"


def test_set_module_tensor_shares_tied_copies_through_the_map():
    model = Nested()
    model.second.linear1.weight = model.first.linear1.weight
    stored = model.first.linear1.weight.detach().clone()
    tied_map = {stored.data_ptr(): {}}
    set_module_tensor_to_device(model, "first.linear1.weight", "meta")
    set_module_tensor_to_device(model, "second.linear1.weight", "meta")
    set_module_tensor_to_device(model, "first.linear1.weight", "cpu", value=stored, tied_params_map=tied_map)
    materialised = tied_map[stored.data_ptr()]["cpu"]
    assert materialised is model.first.linear1.weight
    set_module_tensor_to_device(model, "second.linear1.weight", "cpu", value=stored, tied_params_map=tied_map)
    assert model.second.linear1.weight is materialised  # the second member re-uses the copy instead of making its own


def test_align_module_device_moves_and_restores_and_serves_offloaded_modules():
    layer = nn.Linear(3, 2)
    before = layer.weight
    with align_module_device(layer):
        assert layer.weight is before  # nothing asked: nothing moves
    with align_module_device(layer, torch.device("cpu")):
        assert layer.weight.device.type == "cpu"
    model = Block()
    expected = {k: v.clone() for k, v in model.state_dict().items()}
    cpu_offload(model, execution_device=torch.device("cpu"))
    assert has_offloaded_params(model.linear1) and model.linear1.weight.device.type == "meta"
    with align_module_device(model.linear1):
        assert torch.equal(model.linear1.weight, expected["linear1.weight"])
    assert model.linear1.weight.device.type == "meta"
    gathered = get_state_dict_offloaded_model(model)
    assert set(gathered) == set(expected) and all(torch.equal(gathered[k], expected[k]) for k in expected)


def test_shard_checkpoint_keeps_shared_storage_together_and_names_files():
    weights = {"a": torch.zeros(100), "b": torch.zeros(100), "c": torch.zeros(100), "note": "text entries are skipped"}
    weights["a_view"] = weights["a"][:50]
    one, index = shard_checkpoint(weights, max_shard_size="10KB", weights_name="model.safetensors")
    assert list(one) == ["model.safetensors"] and index is None
    files, index = shard_checkpoint(weights, max_shard_size=800, weights_name="model.safetensors")
    assert list(files) == ["model-00001-of-00002.safetensors", "model-00002-of-00002.safetensors"]
    assert set(files["model-00001-of-00002.safetensors"]) == {"a", "b", "a_view"} and set(files["model-00002-of-00002.safetensors"]) == {"c"}
    assert index["metadata"]["total_size"] == 1200 and index["weight_map"]["a_view"] == "model-00001-of-00002.safetensors"
    big, _ = shard_checkpoint({"huge": torch.zeros(1000), "small": torch.zeros(1)}, max_shard_size=100, weights_name="pytorch_model.bin")
    assert list(big) == ["pytorch_model-00001-of-00002.bin", "pytorch_model-00002-of-00002.bin"]  # an oversized tensor gets its own file


def test_load_state_dict_safetensors_routes_by_device_map(tmp_path):
    from safetensors.torch import save_file

    model = Nested()
    path = str(tmp_path / "model.safetensors")
    save_file(model.state_dict(), path, metadata={"format": "pt"})
    plain = load_state_dict(path)
    assert set(plain) == set(model.state_dict())
    routed = load_state_dict(path, device_map={"first": "cpu", "second": "disk", "head": "cpu"})
    assert set(routed) == set(plain) and all(t.device.type == "cpu" for t in routed.values())
    assert torch.equal(routed["second.linear2.weight"], model.second.linear2.weight)
    single = load_state_dict(path, device_map={"": "disk"})
    assert all(t.device.type == "cpu" for t in single.values())
    save_file(model.state_dict(), str(tmp_path / "tf.safetensors"), metadata={"format": "tf"})
    with pytest.raises(ValueError, match="we need a the pt format"):
        load_state_dict(str(tmp_path / "tf.safetensors"))
    save_file(model.state_dict(), str(tmp_path / "odd.safetensors"), metadata={"format": "onnx"})
    with pytest.raises(OSError, match="valid metadata"):
        load_state_dict(str(tmp_path / "odd.safetensors"))
    torch.save(model.state_dict(), tmp_path / "weights.bin")
    assert set(load_state_dict(str(tmp_path / "weights.bin"))) == set(plain)


def _save_sharded(model, folder, limit=200):
    from safetensors.torch import save_file

    files, index = shard_checkpoint(model.state_dict(), max_shard_size=limit, weights_name="model.safetensors")
    for name, tensors in files.items():
        save_file({k: v.contiguous() for k, v in tensors.items()}, os.path.join(folder, name), metadata={"format": "pt"})
    with open(os.path.join(folder, "model.safetensors.index.json"), "w") as f:
        json.dump(index, f)
    return files


def test_load_checkpoint_in_model_cpu_disk_and_parked_state_dict(tmp_path):
    torch.manual_seed(0)
    source = Nested()
    ckpt = tmp_path / "ckpt"
    ckpt.mkdir()
    assert len(_save_sharded(source, str(ckpt))) > 1
    with init_empty_weights():
        target = Nested()
    device_map = {"first": "cpu", "second": "disk", "head": "cpu"}
    with pytest.raises(ValueError, match="offload_folder"):
        load_checkpoint_in_model(target, str(ckpt), device_map=device_map)
    offload = tmp_path / "offload"
    load_checkpoint_in_model(target, str(ckpt), device_map=device_map, offload_folder=str(offload), offload_state_dict=True)
    assert torch.equal(target.first.linear1.weight, source.first.linear1.weight) and torch.equal(target.head.bias, source.head.bias)
    assert target.second.linear1.weight.device.type == "meta"
    index = json.load(open(offload / "index.json"))
    assert "second.linear1.weight" in index and "second.batchnorm.running_mean" not in index  # buffers stay unless offload_buffers
    assert os.path.exists(offload / "second.linear2.bias.dat")
    # the index file itself, and the dtype / keep-in-fp32 handling
    with init_empty_weights():
        halved = Nested()
    load_checkpoint_in_model(halved, str(ckpt / "model.safetensors.index.json"), device_map={"": "cpu"}, dtype="float16",
                             keep_in_fp32_modules=["batchnorm"])
    assert halved.first.linear1.weight.dtype == torch.float16 and halved.first.batchnorm.weight.dtype == torch.float32
    assert halved.first.batchnorm.num_batches_tracked.dtype == torch.int64


def test_load_checkpoint_in_model_without_map_and_error_paths(tmp_path):
    source = Nested()
    torch.save({**source.state_dict(), "stray.weight": torch.zeros(1)}, tmp_path / "pytorch_model.bin")
    target = Nested()
    load_checkpoint_in_model(target, str(tmp_path))  # folder holding a single weights file; the stray key is only reported
    assert all(torch.equal(a, b) for a, b in zip(target.state_dict().values(), source.state_dict().values()))
    with pytest.raises(ValueError, match="should be the path"):
        load_checkpoint_in_model(target, str(tmp_path / "nowhere"))
    empty = tmp_path / "empty"
    empty.mkdir()
    with pytest.raises(ValueError, match="is not a folder containing"):
        load_checkpoint_in_model(target, str(empty))
    for name in ("a.index.json", "b.index.json"):
        (empty / name).write_text("{}")
    with pytest.raises(ValueError, match="more than one"):
        load_checkpoint_in_model(target, str(empty))
    with init_empty_weights():
        hollow = Nested()
    with pytest.raises(ValueError, match="doesn't have any device set"):
        load_checkpoint_in_model(hollow, str(tmp_path / "pytorch_model.bin"), device_map={"first": "cpu"})


def test_load_checkpoint_reties_tied_weights(tmp_path):
    source = Nested()
    source.second.linear1.weight = source.first.linear1.weight
    torch.save(source.state_dict(), tmp_path / "pytorch_model.bin")
    with init_empty_weights():
        target = Nested()
    target.second.linear1.weight = target.first.linear1.weight
    load_checkpoint_in_model(target, str(tmp_path / "pytorch_model.bin"), device_map={"": "cpu"})
    assert target.second.linear1.weight is target.first.linear1.weight and target.first.linear1.weight.device.type == "cpu"


def test_infer_auto_device_map_spills_in_order_and_reserves_the_largest_layer():
    model = Nested()
    sizes = compute_module_sizes(model)
    everything = infer_auto_device_map(model, max_memory={"cpu": sizes[""] * 2})
    assert everything == {"": "cpu"}
    # room for `first` plus the reserve for the largest layer, not for `second`: the rest goes to disk, in definition order
    tight = infer_auto_device_map(model, max_memory={"cpu": sizes["first"] * 2 + 8}, no_split_module_classes=["Block"])
    assert tight["first"] == "cpu" and tight["second"] == "disk" and tight["head"] == "disk"


# ------------------------------------------------------------------------------------------------- ties and fallback placement
def _stack_of_linears(n, tie=False):
    import torch.nn as nn

    stack = nn.ModuleList([nn.Linear(10, 10) for _ in range(n)])
    if tie:
        for i in range(1, n):
            stack[i].weight = stack[0].weight
    return stack


def test_size_with_ties_charges_each_holder_once_and_not_the_shared_tensor():
    from accelerate_hpc_test_b200.utils.modeling import compute_module_sizes, find_tied_parameters, get_module_size_with_ties

    # twelve layers so that "1" is a string prefix of "10" and "11" without holding their tensors
    stack = _stack_of_linears(12, tie=True)
    sizes = compute_module_sizes(stack)
    group = sorted(find_tied_parameters(stack)[0])
    assert len(group) == 12
    pending = list(stack.named_children())[1:]
    outside = [p for p in group if p != "0.weight"]
    total, names, holders = get_module_size_with_ties(outside, sizes["0"], sizes, pending)
    assert total == sizes["0"] + 11 * sizes["0.bias"]          # the shared weight is paid for once
    assert names == [p.rsplit(".", 1)[0] for p in outside] and holders == [stack[int(n)] for n in names]
    # nothing tied: the size comes back unchanged
    assert get_module_size_with_ties([], 123, sizes, pending) == (123, [], [])
    # two tied tensors below one holder: the holder is charged once
    import torch.nn as nn

    pair = nn.Sequential(nn.Linear(10, 10), nn.Sequential(nn.Linear(10, 10), nn.Linear(10, 10)))
    pair[1][0].weight = pair[0].weight
    pair[1][1].weight = pair[0].weight
    sizes = compute_module_sizes(pair)
    total, names, _ = get_module_size_with_ties(["1.0.weight", "1.1.weight"], sizes["0"], sizes, [("1", pair[1])])
    assert names == ["1"] and total == sizes["0"] + 2 * sizes["1.0.bias"]


def test_fallback_allocate_finds_the_first_piece_that_fits_and_opens_its_ancestors():
    import torch.nn as nn

    from accelerate_hpc_test_b200.utils.modeling import compute_module_sizes, fallback_allocate

    model = nn.Sequential(OrderedDict(big=nn.Sequential(OrderedDict(a=nn.Linear(100, 100), b=nn.Linear(4, 4))), tail=nn.Linear(50, 50)))
    sizes = compute_module_sizes(model)
    entries = list(model.named_children())
    # nothing as large as `big` fits, its second child does; `big` is replaced by its remaining pieces in what is left
    name, piece, rest = fallback_allocate(entries, sizes, sizes["big.b"])
    assert name == "big.b" and piece is model.big.b and [n for n, _ in rest] == ["big.a", "tail"]
    # with `big` un-splittable the search moves on to `tail`
    name, piece, rest = fallback_allocate(entries, sizes, sizes["tail"], no_split_module_classes=["Sequential"])
    assert name == "tail" and [n for n, _ in rest] == ["big"]
    # nothing fits / a limit that is not a size: the list comes back untouched
    assert fallback_allocate(entries, sizes, 8) == (None, None, entries)
    assert fallback_allocate(entries, sizes, "a lot") == (None, None, entries)
    # sizes given the way `max_memory` spells them
    assert fallback_allocate(entries, sizes, "11KB", no_split_module_classes=["Sequential"])[0] == "tail"
    assert fallback_allocate(entries, sizes, "11KB")[0] == "big.b"
    # a piece tied to something outside it only fits if its partner's holder fits too
    model.tail2 = nn.Linear(4, 4)
    model.tail2.weight = model.big.b.weight
    sizes = compute_module_sizes(model)
    entries = list(model.named_children())
    ties = [["big.b.weight", "tail2.weight"]]
    assert fallback_allocate(entries, sizes, sizes["big.b"], tied_parameters=ties)[0] is None
    name, _, rest = fallback_allocate(entries, sizes, sizes["big.b"] + sizes["tail2.bias"], tied_parameters=ties)
    assert name == "big.b" and [n for n, _ in rest] == ["big.a", "tail", "tail2"]


def test_device_map_with_fallback_fills_the_first_device_instead_of_leaving_it_empty():
    import torch.nn as nn

    from accelerate_hpc_test_b200.utils.modeling import compute_module_sizes, infer_auto_device_map

    model = nn.Sequential(OrderedDict(wide=nn.Linear(200, 200), narrow=nn.Linear(10, 10), other=nn.Linear(200, 200)))
    sizes = compute_module_sizes(model)
    # device 0 must keep room for the largest layer, so `wide` itself cannot go there
    budget = {0: sizes["wide"] + sizes["narrow"] + 10, 1: 10 * sizes[""]}
    plain = infer_auto_device_map(model, max_memory=budget)
    assert 0 not in plain.values()
    filled = infer_auto_device_map(model, max_memory=budget, fallback_allocation=True)
    assert filled == {"narrow": 0, "wide": 1, "other": 1}
    # tied pieces stay together under fallback as well
    model.twin = nn.Linear(10, 10)
    model.twin.weight = model.narrow.weight
    filled = infer_auto_device_map(model, max_memory={0: budget[0] + sizes["narrow.bias"], 1: budget[1]}, fallback_allocation=True)
    assert filled["narrow"] == filled["twin"] == 0 and filled["wide"] == 1
