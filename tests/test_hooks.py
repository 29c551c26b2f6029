"""Module hooks (behaviour ports of the reference's tests/test_hooks.py): protocol, installation, device alignment, weight
streaming from a weights map, tied weights, block-wise device maps, whole-model CPU offload, storage-dtype casting."""

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200.hooks import (
    AlignDevicesHook,
    CpuOffload,
    LayerwiseCastingHook,
    ModelHook,
    SequentialHook,
    UserCpuOffloadHook,
    add_hook_to_module,
    attach_align_device_hook,
    attach_align_device_hook_on_blocks,
    attach_execution_device_hook,
    remove_hook_from_module,
    remove_hook_from_submodules,
)
from accelerate_hpc_test_b200.state import PartialState

META = torch.device("meta")


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.linear1 = nn.Linear(3, 4)
        self.batchnorm = nn.BatchNorm1d(4)
        self.linear2 = nn.Linear(4, 5)

    def forward(self, x):
        return self.linear2(self.batchnorm(self.linear1(x)))


class PreAdd(ModelHook):
    def pre_forward(self, module, *args, **kwargs):
        return (args[0] + 1,) + args[1:], kwargs


class PostAdd(ModelHook):
    def post_forward(self, module, output):
        return output + 1


class PostDetach(ModelHook):
    def post_forward(self, module, output):
        return output.detach()


class Tagger(ModelHook):
    def __init__(self, log, tag):
        self.log, self.tag = log, tag

    def init_hook(self, module):
        self.log.append(("init", self.tag))
        return module

    def pre_forward(self, module, *args, **kwargs):
        self.log.append(("pre", self.tag))
        return args, kwargs

    def post_forward(self, module, output):
        self.log.append(("post", self.tag))
        return output

    def detach_hook(self, module):
        self.log.append(("detach", self.tag))
        return module


@pytest.fixture
def x():
    torch.manual_seed(0)
    return torch.randn(2, 3)


# --------------------------------------------------------------------------- protocol + installation
def test_add_and_remove_keep_forward_signature_and_name(x):
    m = Net()
    original = m.forward
    hook = ModelHook()
    add_hook_to_module(m, hook)
    assert m._hf_hook is hook and hasattr(m, "_old_forward")
    assert m.forward.__name__ == "forward" and m.forward is not original
    assert list(__import__("inspect").signature(m.forward).parameters) == ["x"]
    remove_hook_from_module(m)
    assert not hasattr(m, "_hf_hook") and not hasattr(m, "_old_forward")
    assert m.forward.__func__ is Net.forward


def test_pre_forward_changes_inputs(x):
    m = Net().eval()
    expected = m(x + 1)
    add_hook_to_module(m, PreAdd())
    assert torch.allclose(m(x), expected, atol=1e-6)
    # adding a second hook without append REPLACES the first
    add_hook_to_module(m, PreAdd())
    assert torch.allclose(m(x), expected, atol=1e-6)
    # with append they chain: +2
    add_hook_to_module(m, PreAdd(), append=True)
    assert isinstance(m._hf_hook, SequentialHook)
    remove_hook_from_module(m)
    assert torch.allclose(m(x + 2), m(x + 2))


def test_post_forward_changes_outputs(x):
    m = Net().eval()
    base = m(x)
    add_hook_to_module(m, PostAdd())
    assert torch.allclose(m(x), base + 1, atol=1e-6)
    add_hook_to_module(m, SequentialHook(PostAdd(), PostAdd()))
    assert torch.allclose(m(x), base + 2, atol=1e-6)


def test_no_grad_flag_runs_forward_without_graph(x):
    m = Net()
    assert m(x).requires_grad
    h = ModelHook()
    h.no_grad = True
    add_hook_to_module(m, h)
    assert not m(x).requires_grad
    add_hook_to_module(m, PostDetach())
    assert not m(x).requires_grad


def test_sequential_hook_order_and_detach(x):
    log = []
    m = Net().eval()
    add_hook_to_module(m, Tagger(log, "a"))
    add_hook_to_module(m, Tagger(log, "b"), append=True)
    m(x)
    remove_hook_from_module(m)
    # the first hook is detached when the chain is rebuilt, and re-initialised inside the chain
    assert log[0] == ("init", "a") and ("pre", "a") in log and log.index(("pre", "a")) < log.index(("pre", "b"))
    assert log.index(("post", "a")) < log.index(("post", "b"))
    assert log[-2:] == [("detach", "a"), ("detach", "b")]


def test_remove_hook_recurse_and_submodules(x):
    m = Net()
    for sub in (m, m.linear1, m.linear2):
        add_hook_to_module(sub, ModelHook())
    remove_hook_from_module(m, recurse=True)
    assert not any(hasattr(s, "_hf_hook") for s in m.modules())
    for sub in (m.linear1, m.batchnorm):
        add_hook_to_module(sub, ModelHook())
    remove_hook_from_submodules(m)
    assert not any(hasattr(s, "_hf_hook") for s in m.modules())


# --------------------------------------------------------------------------- AlignDevicesHook
def test_align_devices_as_simple_placement(x):
    m = Net().eval()
    base = m(x)
    for sub in (m.linear1, m.batchnorm, m.linear2):
        add_hook_to_module(sub, AlignDevicesHook(execution_device="cpu"))
    assert torch.allclose(m(x), base)
    assert all(p.device.type == "cpu" for p in m.parameters())
    assert "execution_device=cpu" in repr(m.linear1._hf_hook)


def test_align_devices_offload_parks_weights_on_meta_between_calls(x):
    m = Net().eval()
    base = m(x)
    for sub in (m.linear1, m.batchnorm, m.linear2):
        add_hook_to_module(sub, AlignDevicesHook(execution_device="cpu", offload=True))
    assert m.linear1.weight.device == META and m.batchnorm.weight.device == META
    assert m.batchnorm.running_mean.device.type == "cpu"  # buffers stay unless offload_buffers
    assert torch.allclose(m(x), base, atol=1e-6)
    assert m.linear2.weight.device == META
    remove_hook_from_submodules(m)
    assert all(p.device.type == "cpu" for p in m.parameters())
    assert torch.allclose(m(x), base, atol=1e-6)


def test_align_devices_offload_buffers_too(x):
    m = Net().eval()
    base = m(x)
    for sub in (m.linear1, m.batchnorm, m.linear2):
        add_hook_to_module(sub, AlignDevicesHook(execution_device="cpu", offload=True, offload_buffers=True))
    assert m.batchnorm.running_mean.device == META
    assert torch.allclose(m(x), base, atol=1e-6)
    assert m.batchnorm.running_var.device == META
    remove_hook_from_submodules(m)
    assert m.batchnorm.running_mean.device.type == "cpu"


def test_align_devices_reads_from_given_weights_map(x):
    m = Net().eval()
    base = m(x)
    stored = {k: v.clone() for k, v in m.linear1.state_dict().items()}
    with torch.no_grad():
        m.linear1.weight.zero_()  # the module's own values must not matter any more
    add_hook_to_module(m.linear1, AlignDevicesHook(execution_device="cpu", offload=True, weights_map=stored))
    assert torch.allclose(m(x), base, atol=1e-6)


def test_io_same_device_and_skip_keys(x):
    class TakesKw(nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = nn.Linear(3, 3)

        def forward(self, x, extra=None):
            return self.lin(x), extra

    m = TakesKw()
    add_hook_to_module(m, AlignDevicesHook(execution_device="cpu", io_same_device=True, skip_keys="extra"))
    marker = object()
    out, extra = m(x, extra=marker)
    assert extra is marker and out.device == x.device and m._hf_hook.input_device == x.device


def test_tied_weights_are_materialised_once_and_released():
    class Tied(nn.Module):
        def __init__(self):
            super().__init__()
            self.a = nn.Linear(4, 4, bias=False)
            self.b = nn.Linear(4, 4, bias=False)
            self.b.weight = self.a.weight

        def forward(self, x):
            return self.b(self.a(x))

    m = Tied().eval()
    x = torch.randn(2, 4)
    base = m(x)
    shared = {m.a.weight.data_ptr(): {}}
    weights = {"a.weight": m.a.weight.detach().clone()}
    weights["b.weight"] = weights["a.weight"]
    shared = {weights["a.weight"].data_ptr(): {}}
    from accelerate_hpc_test_b200.utils.offload import PrefixedDataset

    for name in ("a", "b"):
        add_hook_to_module(getattr(m, name), AlignDevicesHook(execution_device="cpu", offload=True,
                                                              weights_map=PrefixedDataset(weights, f"{name}."), tied_params_map=shared))
    assert torch.allclose(m(x), base, atol=1e-6)
    assert m.a.weight.device == META and m.b.weight.device == META
    assert all(len(v) == 0 for v in shared.values())  # nothing left alive after the forward


def test_attach_align_device_hook_hooks_only_tensor_owners(x):
    m = Net().eval()
    base = m(x)
    attach_align_device_hook(m, execution_device="cpu", offload=True)
    assert not hasattr(m, "_hf_hook") and all(hasattr(s, "_hf_hook") for s in (m.linear1, m.batchnorm, m.linear2))
    assert torch.allclose(m(x), base, atol=1e-6)
    remove_hook_from_submodules(m)
    assert all(p.device.type == "cpu" for p in m.parameters())


def test_attach_align_device_hook_preload_module_classes(x):
    class Block(nn.Module):
        def __init__(self):
            super().__init__()
            self.inner = nn.Linear(3, 3)

        def forward(self, x):
            return self.inner(x)

    class Outer(nn.Module):
        def __init__(self):
            super().__init__()
            self.block = Block()
            self.head = nn.Linear(3, 2)

        def forward(self, x):
            return self.head(self.block(x))

    m = Outer().eval()
    base = m(x)
    attach_align_device_hook(m, execution_device="cpu", offload=True, weights_map={k: v.clone() for k, v in m.state_dict().items()},
                             preload_module_classes=["Block"])
    assert hasattr(m.block, "_hf_hook") and m.block._hf_hook.place_submodules and not hasattr(m.block.inner, "_hf_hook")
    assert torch.allclose(m(x), base, atol=1e-6)


def test_attach_on_blocks_with_device_map(x):
    m = Net().eval()
    base = m(x)
    weights = {k: v.clone() for k, v in m.state_dict().items()}
    execution = {"linear1": "cpu", "batchnorm": "cpu", "linear2": "cpu", "": "cpu"}
    offload = {"linear1": True, "batchnorm": False, "linear2": True}
    attach_align_device_hook_on_blocks(m, execution_device=execution, offload=offload, weights_map=weights)
    assert m.linear1.weight.device == META and m.batchnorm.weight.device.type == "cpu" and m.linear2.weight.device == META
    assert m._hf_hook.io_same_device
    assert torch.allclose(m(x), base, atol=1e-6)


def test_attach_on_blocks_uniform_arguments(x):
    m = Net().eval()
    base = m(x)
    attach_align_device_hook_on_blocks(m, execution_device="cpu", offload=False)
    assert m._hf_hook.place_submodules and m._hf_hook.io_same_device
    assert torch.allclose(m(x), base)
    m2 = Net().eval()
    base2 = m2(x)
    attach_align_device_hook_on_blocks(m2, execution_device="cpu", offload=True)
    assert m2.linear1.weight.device == META and torch.allclose(m2(x), base2, atol=1e-6)


def test_attach_execution_device_hook_skips_stateless_modules():
    m = nn.Sequential(nn.Linear(2, 2), nn.ReLU(), nn.Linear(2, 2))
    attach_execution_device_hook(m, "cpu")
    assert hasattr(m, "_hf_hook") and hasattr(m[0], "_hf_hook") and not hasattr(m[1], "_hf_hook")


# --------------------------------------------------------------------------- CpuOffload / casting
def test_cpu_offload_chain_evicts_previous_model():
    PartialState(cpu=True)
    a, b = nn.Linear(2, 2), nn.Linear(2, 2)
    ha = CpuOffload(execution_device=torch.device("cpu"))
    add_hook_to_module(a, ha)
    handle_a = UserCpuOffloadHook(a, ha)
    calls = []
    handle_a.offload = lambda: calls.append("evict")  # record instead of moving (everything is on CPU here)
    hb = CpuOffload(execution_device=torch.device("cpu"), prev_module_hook=handle_a)
    add_hook_to_module(b, hb)
    b(a(torch.randn(1, 2)))
    assert calls == []  # previous model already on CPU: nothing to evict
    UserCpuOffloadHook(b, hb).remove()
    assert not hasattr(b, "_hf_hook")
    PartialState._reset_state()


def test_layerwise_casting_hook_round_trip():
    lin = nn.Linear(4, 4)
    add_hook_to_module(lin, LayerwiseCastingHook(torch.float16, torch.float32, non_blocking=False))
    assert lin.weight.dtype == torch.float16
    seen = []
    lin.register_forward_pre_hook(lambda mod, args: seen.append(mod.weight.dtype))
    out = lin(torch.randn(2, 4))
    assert out.dtype == torch.float32 and lin.weight.dtype == torch.float16
