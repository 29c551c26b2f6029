"""AcceleratedScheduler / AcceleratedOptimizer semantics (behaviour ports of the reference's tests/test_scheduler.py and
tests/test_optimizer.py)."""

import pickle

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.optimizer import AcceleratedOptimizer, move_to_device
from accelerate_hpc_test_b200.scheduler import AcceleratedScheduler
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState
from accelerate_hpc_test_b200.utils import GradientAccumulationPlugin


@pytest.fixture(autouse=True)
def fresh_state():
    yield
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


def trio(accelerator, lr=1.0, **sched_kw):
    model = nn.Linear(2, 2)
    opt = torch.optim.SGD(model.parameters(), lr=lr)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda n: 1 - n / 10)
    return accelerator.prepare(model, opt, sched)


def one_step(acc, model, opt, sched):
    with acc.accumulate(model):
        acc.backward(model(torch.ones(1, 2)).sum())
        opt.step()
        sched.step()
        opt.zero_grad()


# --------------------------------------------------------------------------- scheduler
def test_scheduler_follows_real_optimizer_steps_only():
    acc = Accelerator(cpu=True, gradient_accumulation_steps=2)
    model, opt, sched = trio(acc)
    lrs = []
    for _ in range(4):
        one_step(acc, model, opt, sched)
        lrs.append(sched.get_last_lr()[0])
    assert lrs == pytest.approx([1.0, 0.9, 0.9, 0.8])  # advances every second call


def test_scheduler_adjust_false_does_not_touch_step_count():
    acc = Accelerator(cpu=True, gradient_accumulation_plugin=GradientAccumulationPlugin(num_steps=2, adjust_scheduler=False))
    model, opt, sched = trio(acc)
    before = sched.scheduler._step_count
    one_step(acc, model, opt, sched)  # accumulating micro-step
    assert sched.scheduler._step_count == before
    acc2_state = acc.gradient_state
    assert acc2_state.adjust_scheduler is False


def test_scheduler_bumps_internal_counter_when_gradient_state_asks_for_it(monkeypatch):
    acc = Accelerator(cpu=True, gradient_accumulation_steps=2)
    model, opt, sched = trio(acc)
    before = sched.scheduler._step_count
    one_step(acc, model, opt, sched)
    # as in the reference, the plugin's default (True) is not among the "changed" kwargs, so the state reports False
    assert acc.gradient_state.adjust_scheduler is False and sched.scheduler._step_count == before
    monkeypatch.setitem(acc.gradient_state.plugin_kwargs, "adjust_scheduler", True)
    acc.gradient_state._set_sync_gradients(False)
    sched.step()
    assert sched.scheduler._step_count == before + 1 and sched.get_last_lr()[0] == 1.0


def test_scheduler_ticks_once_per_process_unless_batches_are_split(monkeypatch):
    acc = Accelerator(cpu=True)
    model, opt, sched = trio(acc)
    monkeypatch.setitem(PartialState._shared_state, "num_processes", 3)
    one_step(acc, model, opt, sched)
    assert sched.get_last_lr()[0] == pytest.approx(0.7)
    sched.split_batches = True
    one_step(acc, model, opt, sched)
    assert sched.get_last_lr()[0] == pytest.approx(0.6)


def test_scheduler_never_steps_past_total_steps(monkeypatch):
    acc = Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    sched = torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=0.1, total_steps=4)
    model, opt, sched = acc.prepare(model, opt, sched)
    monkeypatch.setitem(PartialState._shared_state, "num_processes", 8)
    one_step(acc, model, opt, sched)  # would be 8 ticks: must stop at the horizon instead of raising
    assert sched.scheduler._step_count <= 5


def test_scheduler_holds_when_optimizer_step_was_skipped():
    acc = Accelerator(cpu=True)
    model, opt, sched = trio(acc)
    opt._is_overflow = True
    sched.step()
    assert sched.get_last_lr()[0] == 1.0
    opt._is_overflow = False
    sched.step()
    assert sched.get_last_lr()[0] == pytest.approx(0.9)


def test_scheduler_manual_mode_and_forwarding_and_pickle():
    acc = Accelerator(cpu=True, step_scheduler_with_optimizer=False, gradient_accumulation_steps=4)
    model, opt, sched = trio(acc)
    sched.step()
    sched.step()
    assert sched.get_last_lr()[0] == pytest.approx(0.8)  # passes straight through, accumulation ignored
    state = sched.state_dict()
    sched.load_state_dict(state)
    assert isinstance(sched, AcceleratedScheduler) and not hasattr(sched, "definitely_not_there")
    with pytest.raises(AttributeError):
        sched.nope
    clone = pickle.loads(pickle.dumps(sched.scheduler.state_dict()))
    assert clone["_step_count"] == state["_step_count"]


# --------------------------------------------------------------------------- optimizer
def test_optimizer_is_a_noop_while_accumulating():
    acc = Accelerator(cpu=True, gradient_accumulation_steps=2)
    model, opt, sched = trio(acc)
    w0 = model.weight.detach().clone()
    with acc.accumulate(model):
        acc.backward(model(torch.ones(1, 2)).sum())
        opt.step()
        opt.zero_grad()
    assert torch.equal(model.weight, w0) and model.weight.grad is not None  # neither stepped nor cleared
    with acc.accumulate(model):
        acc.backward(model(torch.ones(1, 2)).sum())
        opt.step()
        opt.zero_grad()
    assert not torch.equal(model.weight, w0) and model.weight.grad is None


def test_optimizer_exposes_wrapped_state_and_groups():
    acc = Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    inner = torch.optim.AdamW(model.parameters(), lr=0.5)
    opt = acc.prepare_optimizer(inner)
    assert isinstance(opt, AcceleratedOptimizer) and isinstance(opt, torch.optim.Optimizer)
    assert opt.param_groups is inner.param_groups and opt.state is inner.state and opt.defaults is inner.defaults
    opt.param_groups = inner.param_groups
    opt.add_param_group({"params": [nn.Parameter(torch.zeros(1))]})
    assert len(inner.param_groups) == 2
    opt.train()
    opt.eval()  # no-ops for optimizers without modes


def test_optimizer_state_follows_the_device_on_prepare():
    acc = Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    inner = torch.optim.Adam(model.parameters())
    model(torch.ones(1, 2)).sum().backward()
    inner.step()
    opt = AcceleratedOptimizer(inner, device_placement=True)
    assert all(v["exp_avg"].device.type == "cpu" for v in opt.state.values())
    nested = move_to_device({"a": [torch.ones(1), (torch.zeros(1), "s")], "n": 3}, "cpu")
    assert isinstance(nested["a"][1], tuple) and nested["n"] == 3


class FakeScaler:
    """GradScaler stand-in: calls optimizer.step() only when told that the gradients are finite."""

    def __init__(self):
        self.finite = True
        self.updates = 0

    def step(self, optimizer, closure=None):
        if self.finite:
            optimizer.step()

    def update(self):
        self.updates += 1


def test_optimizer_detects_steps_skipped_by_the_scaler():
    Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    inner = torch.optim.SGD(model.parameters(), lr=0.1)
    scaler = FakeScaler()
    opt = AcceleratedOptimizer(inner, device_placement=False, scaler=scaler)
    model(torch.ones(1, 2)).sum().backward()
    opt.step()
    assert opt.step_was_skipped is False and scaler.updates == 1
    scaler.finite = False
    w = model.weight.detach().clone()
    opt.step()
    assert opt.step_was_skipped is True and torch.equal(model.weight, w)
    assert inner.step.__func__ is torch.optim.SGD.step or callable(inner.step)  # original step restored
    scaler.finite = True
    opt.step()
    assert opt.step_was_skipped is False


def test_optimizer_pickles_without_transient_fields():
    Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    opt = AcceleratedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1), device_placement=False, scaler=FakeScaler())
    opt.attach_engine(object())
    clone = pickle.loads(pickle.dumps(opt))
    assert clone._engine is None and clone._optimizer_patched_step_method is not None and clone.param_groups[0]["lr"] == 0.1


def test_optimizer_routes_to_an_attached_engine():
    Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    inner = torch.optim.SGD(model.parameters(), lr=0.1)
    opt = AcceleratedOptimizer(inner, device_placement=False)

    class Engine:
        def __init__(self):
            self.calls = []

        def optimizer_step(self, optimizer, scaler):
            self.calls.append("step")
            return True

        def zero_grad(self):
            self.calls.append("zero")

        def optimizer_state_dict(self, optimizer):
            return {"engine": True}

        def load_optimizer_state_dict(self, optimizer, sd):
            self.calls.append(("load", sd))

    e = Engine()
    opt.attach_engine(e)
    opt.step()
    opt.zero_grad()
    opt.load_state_dict({"x": 1})
    assert e.calls == ["step", "zero", ("load", {"x": 1})] and opt.state_dict() == {"engine": True} and opt.step_was_skipped is True


def test_zero_grad_set_to_none_argument():
    Accelerator(cpu=True)
    model = nn.Linear(2, 2)
    opt = AcceleratedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1), device_placement=False)
    model(torch.ones(1, 2)).sum().backward()
    opt.zero_grad(set_to_none=False)
    assert model.weight.grad is not None and model.weight.grad.abs().sum() == 0
    opt.zero_grad()
    assert model.weight.grad is None


def test_fused_adamw_switches_to_updates_behind_the_reduction_only_after_two_clean_steps():
    """The engine-driven fusion (DDP: update a bucket right behind its all-reduce) is armed by probing: two optimizer steps in a row
    that each followed exactly one reduction of every parameter.  A loop that accumulates by hand never gets there."""
    import torch

    from accelerate_hpc_test_b200.optim import FusedAdamW

    def fresh():
        torch.manual_seed(0)
        p = torch.nn.Parameter(torch.randn(16))
        opt = FusedAdamW([p], lr=0.1, weight_decay=0.0)
        assert opt.enable_update_after_reduce(stream=None, mode="auto") and opt._bwd_mode == "probing"
        assert not opt.enable_update_after_reduce(stream=None)  # armed once
        return p, opt

    def backward(p, opt, reductions=1):
        p.grad = torch.ones_like(p)
        done = [opt.update_reduced(p, p.grad) for _ in range(reductions)]
        if all(done):
            p.grad = None  # spent, as the engine does
        return all(done)

    # one backward per step: probing, probing -> active from the third step on
    p, opt = fresh()
    reference = torch.nn.Parameter(p.detach().clone())
    ref_opt = FusedAdamW([reference], lr=0.1, weight_decay=0.0)
    modes = []
    for _ in range(4):
        updated = backward(p, opt)
        opt.step()
        modes.append((opt._bwd_mode, updated))
        reference.grad = torch.ones_like(reference)
        ref_opt.step()
    assert modes == [("probing", False), ("active", False), ("active", True), ("active", True)]
    assert torch.allclose(p, reference, atol=1e-6)  # the same update, only earlier
    # step() at micro-batch 0, then after every second backward (the `step % accumulation == 0` idiom): never armed
    p, opt = fresh()
    pattern = [1, 2, 2, 2]
    for reductions in pattern:
        assert backward(p, opt, reductions) is False
        opt.step()
    assert opt._bwd_mode == "probing" and opt._probe_clean_steps == 0
    # asking for clipping during probing switches the fusion off for good
    p, opt = fresh()
    backward(p, opt)
    opt.set_clip_coef(torch.ones(1))
    opt.step()
    assert opt._bwd_mode == "off"
    # once active, a second reduction before step() is an error rather than a silent double update
    p, opt = fresh()
    for _ in range(2):
        backward(p, opt)
        opt.step()
    assert backward(p, opt) is True
    with pytest.raises(RuntimeError, match="reduced twice"):
        opt.update_reduced(p, torch.ones_like(p))
