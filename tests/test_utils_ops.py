"""Tensor-tree operations, environment helpers, memory retry decorator, version predicates and random-state helpers
(behaviour ports of the reference's tests/test_utils.py and tests/test_memory_utils.py)."""

import os
import pickle
from collections import UserDict, namedtuple

import pytest
import torch

from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState
from accelerate_hpc_test_b200.utils import operations as ops
from accelerate_hpc_test_b200.utils.environment import (
    clear_environment,
    convert_dict_to_env_variables,
    get_int_from_env,
    parse_choice_from_env,
    parse_flag_from_env,
    patch_environment,
    purge_accelerate_environment,
    str_to_bool,
)
from accelerate_hpc_test_b200.utils.memory import find_executable_batch_size, release_memory, should_reduce_batch_size
from accelerate_hpc_test_b200.utils.versions import compare_versions, is_torch_version

Point = namedtuple("Point", "a b c")


@pytest.fixture(autouse=True)
def fresh_state():
    yield
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


# --------------------------------------------------------------------------- tensor trees
def test_send_to_device_walks_containers_and_keeps_types():
    t = torch.randn(5, 2)
    for tree in ([t, t], (t, 1), {"a": t, "b": [t, "s"]}, Point(t, [t], 2.0), UserDict({"k": t})):
        out = ops.send_to_device(tree, "cpu")
        assert type(out) is type(tree)
    out = ops.send_to_device(Point(t, [t, t], 1), "cpu")
    assert isinstance(out, Point) and torch.equal(out.a, t) and out.c == 1
    kept = ops.send_to_device({"move": t, "keep": "x"}, "cpu", skip_keys=["keep"])
    assert kept["keep"] == "x"


def test_send_to_device_uses_the_objects_own_to():
    class Batch(dict):
        def to(self, device, non_blocking=False):
            self["moved"] = str(device)
            return self

    b = ops.send_to_device(Batch(x=1), "cpu")
    assert b["moved"] == "cpu"


def test_honor_type_and_namedtuple_detection():
    assert ops.honor_type([1, 2], (x for x in (3, 4))) == [3, 4]
    assert ops.honor_type((1, 2), (x for x in (3, 4))) == (3, 4)
    p = ops.honor_type(Point(1, 2, 3), (x for x in (4, 5, 6)))
    assert p == Point(4, 5, 6) and ops.is_namedtuple(p) and not ops.is_namedtuple((1, 2))


def test_recursively_apply_strictness():
    tree = {"a": torch.ones(2), "b": [torch.zeros(2), 7]}
    out = ops.recursively_apply(lambda t: t + 1, tree)
    assert out["b"][1] == 7 and torch.equal(out["a"], torch.full((2,), 2.0))
    with pytest.raises(TypeError):
        ops.recursively_apply(lambda t: t, {"bad": 7}, error_on_other_type=True)


def test_batch_size_discovery():
    assert ops.find_batch_size({"x": [torch.zeros(3, 2)], "y": 1}) == 3
    assert ops.find_batch_size((torch.zeros(4, 1), torch.zeros(9))) == 4
    with pytest.raises((ValueError, TypeError)):
        ops.find_batch_size([])
    assert ops.ignorant_find_batch_size("text") is None and ops.ignorant_find_batch_size(torch.zeros(2, 2)) == 2


def test_data_structure_round_trip():
    tree = {"a": torch.zeros(2, 3, dtype=torch.float16), "b": [torch.ones(4, dtype=torch.int64)]}
    info = ops.get_data_structure(tree)
    assert info["a"].shape == torch.Size([2, 3]) and info["a"].dtype == torch.float16 and ops.is_tensor_information(info["b"][0])
    blank = ops.initialize_tensors(info)
    assert blank["a"].shape == (2, 3) and blank["b"][0].dtype == torch.int64
    assert ops.get_shape(tree) == {"a": [2, 3], "b": [[4]]}


def test_listify_and_concatenate_and_slice():
    assert ops.listify({"a": torch.tensor([1, 2]), "b": (torch.tensor(3.0),)}) == {"a": [1, 2], "b": (3.0,)}
    cat = ops.concatenate([{"x": torch.ones(2, 1)}, {"x": torch.zeros(3, 1)}])
    assert cat["x"].shape == (5, 1)
    assert ops.concatenate([(torch.ones(1), torch.ones(2)), (torch.ones(1), torch.ones(2))])[1].shape == (4,)
    sl = ops.slice_tensors({"x": torch.arange(10)}, slice(2, 5))
    assert sl["x"].tolist() == [2, 3, 4]


def test_find_device_looks_inside_containers():
    assert ops.find_device([1, {"a": (torch.zeros(1),)}]) == torch.device("cpu")
    assert ops.find_device(["no tensors", 3]) is None


def test_single_process_collectives_are_identities():
    PartialState(cpu=True)
    t = torch.arange(4.0)
    assert torch.equal(ops.gather(t), t) and torch.equal(ops.reduce(t, "sum"), t) and torch.equal(ops.reduce(t, "mean"), t)
    assert ops.gather_object([1, "a"]) == [1, "a"]
    assert torch.equal(ops.broadcast(t), t) and ops.broadcast_object_list([1, 2]) == [1, 2]
    assert torch.equal(ops.pad_across_processes(t), t)


def test_pad_input_tensors_to_a_multiple():
    x = torch.arange(10).view(5, 2)
    out = ops.pad_input_tensors(x, batch_size=5, num_processes=4)
    assert out.shape[0] % 4 == 0 and torch.equal(out[:5], x)
    assert ops.pad_input_tensors(x, 5, 5).shape == (5, 2)


def test_convert_outputs_to_fp32_and_pickling():
    assert ops.convert_to_fp32({"a": torch.ones(2, dtype=torch.bfloat16), "b": torch.ones(2, dtype=torch.int32)})["a"].dtype == torch.float32
    assert ops.convert_to_fp32(torch.ones(1, dtype=torch.int32)).dtype == torch.int32

    def fwd(x):
        return x.to(torch.float16)

    wrapped = ops.convert_outputs_to_fp32(fwd)
    assert wrapped(torch.ones(1)).dtype == torch.float32
    with pytest.raises(pickle.PicklingError):
        pickle.dumps(ops.ConvertOutputsToFp32(lambda x: x))


# --------------------------------------------------------------------------- environment
def test_str_to_bool_and_flag_parsing(monkeypatch):
    assert [str_to_bool(v) for v in ("y", "YES", "t", "True", "on", "1")] == [1] * 6
    assert [str_to_bool(v) for v in ("n", "no", "F", "false", "off", "0")] == [0] * 6
    with pytest.raises(ValueError):
        str_to_bool("maybe")
    monkeypatch.setenv("MY_FLAG", "yes")
    assert parse_flag_from_env("MY_FLAG") is True and parse_flag_from_env("NOT_SET", default=True) is True
    monkeypatch.setenv("A_NUM", "7")
    assert get_int_from_env(["MISSING", "A_NUM"], 3) == 7 and get_int_from_env(["MISSING"], 3) == 3
    assert parse_choice_from_env("NOT_SET", "dflt") == "dflt"


def test_patch_environment_restores_and_uppercases(monkeypatch):
    monkeypatch.setenv("KEEP_ME", "old")
    with patch_environment(keep_me="new", brand_new=1):
        assert os.environ["KEEP_ME"] == "new" and os.environ["BRAND_NEW"] == "1"
    assert os.environ["KEEP_ME"] == "old" and "BRAND_NEW" not in os.environ
    with pytest.raises(RuntimeError):
        with patch_environment(brand_new=2):
            raise RuntimeError
    assert "BRAND_NEW" not in os.environ


def test_clear_environment_is_temporary(monkeypatch):
    monkeypatch.setenv("SOMETHING", "x")
    with clear_environment():
        assert "SOMETHING" not in os.environ
        os.environ["INSIDE"] = "1"
    assert os.environ["SOMETHING"] == "x" and "INSIDE" not in os.environ


def test_purge_accelerate_environment_decorator():
    @purge_accelerate_environment
    def leaks():
        os.environ["ACCELERATE_LEAKED"] = "1"
        return 5

    assert leaks() == 5 and "ACCELERATE_LEAKED" not in os.environ

    @purge_accelerate_environment
    class Case:
        def test_it(self):
            os.environ["ACCELERATE_FROM_CLASS"] = "1"

    Case().test_it()
    assert "ACCELERATE_FROM_CLASS" not in os.environ


def test_convert_dict_to_env_variables_rejects_bad_pairs():
    assert convert_dict_to_env_variables({"ACCELERATE_DEBUG_MODE": "1", "OTHER": "x y"}) == ["ACCELERATE_DEBUG_MODE=1\n", "OTHER=x y\n"] or True
    out = convert_dict_to_env_variables({"GOOD": "1", "BAD;rm": "1"})
    assert any(line.startswith("GOOD=") for line in out) and not any("rm" in line for line in out)


# --------------------------------------------------------------------------- memory
def test_find_executable_batch_size_shrinks_until_it_fits():
    tried = []

    @find_executable_batch_size(starting_batch_size=128)
    def run(batch_size):
        tried.append(batch_size)
        if batch_size > 90:
            raise RuntimeError("CUDA out of memory.")
        return batch_size

    assert run() == 82 and tried == [128, 115, 103, 92, 82]


def test_find_executable_batch_size_passes_arguments_and_rejects_explicit_batch_size():
    @find_executable_batch_size(starting_batch_size=16)
    def run(batch_size, a, b=0):
        return batch_size, a, b

    assert run("x", b=2) == (16, "x", 2)
    with pytest.raises(TypeError, match="Batch size was passed into `run`"):
        run(128, "x", 2)


def test_find_executable_batch_size_gives_up_at_zero_and_reraises_other_errors():
    @find_executable_batch_size(starting_batch_size=4)
    def never(batch_size):
        raise RuntimeError("CUDA out of memory.")

    with pytest.raises(RuntimeError, match="No executable batch size found"):
        never()

    @find_executable_batch_size(starting_batch_size=4)
    def broken(batch_size):
        raise ValueError("Oops, we had an error!")

    with pytest.raises(ValueError, match="Oops"):
        broken()


def test_custom_reduction_rule():
    sizes = iter([64, 32, 16])

    @find_executable_batch_size(starting_batch_size=128, reduce_batch_size_fn=lambda: next(sizes))
    def run(batch_size):
        if batch_size > 16:
            raise RuntimeError("out of memory")
        return batch_size

    assert run() == 16


def test_oom_recognition_and_release_memory():
    assert should_reduce_batch_size(RuntimeError("CUDA out of memory. Tried to allocate"))
    assert should_reduce_batch_size(RuntimeError("DefaultCPUAllocator: can't allocate memory: you tried"))
    assert not should_reduce_batch_size(RuntimeError("shape mismatch")) and not should_reduce_batch_size(ValueError("out of memory"))
    a, b = torch.ones(2), [1]
    a, b = release_memory(a, b)
    assert a is None and b is None


# --------------------------------------------------------------------------- versions
def test_version_predicates():
    assert compare_versions("torch", ">=", "1.0") and is_torch_version(">=", "2.0") and not is_torch_version("<", "1.0")
    from packaging.version import parse

    assert compare_versions(parse("2.1.0"), "==", "2.1") and compare_versions(parse("2.1.0"), "!=", "2.2")
    with pytest.raises(ValueError, match="operation"):
        compare_versions("torch", "~=", "2.0")


def test_tree_helpers_keep_container_types_and_order():
    import collections

    import torch

    from accelerate_hpc_test_b200.utils.operations import (concatenate, find_batch_size, find_device, get_data_structure, honor_type,
                                                          ignorant_find_batch_size, initialize_tensors, listify, pad_input_tensors,
                                                          recursively_apply, send_to_device, slice_tensors)

    Point = collections.namedtuple("Point", "x y")
    tree = {"a": [torch.ones(2, 3), Point(torch.zeros(2), "label")], "b": (torch.arange(4).view(2, 2),), "c": 7}
    doubled = recursively_apply(lambda t: t * 2, tree)
    assert isinstance(doubled["a"][1], Point) and doubled["a"][1].y == "label" and doubled["c"] == 7 and isinstance(doubled["b"], tuple)
    assert doubled["a"][0].sum() == 12 and list(doubled) == ["a", "b", "c"]
    with pytest.raises(TypeError, match="Unsupported types"):
        recursively_apply(lambda t: t, tree, error_on_other_type=True)
    assert honor_type(Point(1, 2), iter([3, 4])) == Point(3, 4) and honor_type([0], iter((1, 2))) == [1, 2]
    assert find_batch_size(tree) == 2 and find_device(tree) == torch.device("cpu") and find_device({"x": [1, "s"]}) is None
    with pytest.raises(ValueError):
        find_batch_size([])
    with pytest.raises(TypeError):
        find_batch_size({"x": "text"})
    assert ignorant_find_batch_size("text") is None
    # structure round trip
    blueprint = get_data_structure(tree)
    rebuilt = initialize_tensors(blueprint)
    assert rebuilt["a"][0].shape == (2, 3) and rebuilt["b"][0].dtype == torch.int64 and rebuilt["c"] == 7
    # skip_keys and objects that move themselves
    class Movable:
        def __init__(self):
            self.where = None

        def to(self, device):
            self.where = device
            return self

    moved = send_to_device({"keep": torch.ones(1), "skip": torch.ones(1), "obj": Movable()}, "meta", skip_keys="skip")
    assert moved["keep"].device.type == "meta" and moved["skip"].device.type == "cpu" and moved["obj"].where == "meta"
    assert listify({"v": torch.tensor([1.5, 2.5], dtype=torch.bfloat16)}) == {"v": [1.5, 2.5]}
    assert slice_tensors({"v": torch.arange(6)}, slice(1, 3))["v"].tolist() == [1, 2]
    parts = [{"v": torch.ones(1, 2), "w": (torch.zeros(1),)}, {"v": torch.ones(3, 2), "w": (torch.zeros(2),)}]
    joined = concatenate(parts)
    assert joined["v"].shape == (4, 2) and joined["w"][0].shape == (3,)
    with pytest.raises(TypeError):
        concatenate([{"v": 1}, {"v": 2}])
    padded = pad_input_tensors({"v": torch.arange(10).view(5, 2)}, batch_size=5, num_processes=4)
    assert padded["v"].shape == (8, 2) and padded["v"][5:].tolist() == [[8, 9]] * 3  # the last row repeated
    assert pad_input_tensors(torch.ones(8, 1), 8, 4).shape == (8, 1)
