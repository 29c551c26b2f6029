"""save_state / load_state (behaviour ports of the reference's tests/test_state_checkpointing.py:98-444)."""

import os
import random

import numpy as np
import pytest
import torch
import torch.nn as nn
from torch.utils.data import DataLoader, TensorDataset

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.utils import ProjectConfiguration, set_seed


@pytest.fixture(autouse=True)
def fresh_state():
    yield
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


class Line(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Parameter(torch.randn(1))
        self.b = nn.Parameter(torch.randn(1))

    def forward(self, x):
        return x * self.a + self.b


def loaders(n=64, bs=16):
    x = torch.randn(n)
    y = 3 * x + 2 + 0.1 * torch.randn(n)
    return DataLoader(TensorDataset(x, y), batch_size=bs, shuffle=True)


def train(acc, model, opt, dl, epochs=1, sched=None):
    for _ in range(epochs):
        for x, y in dl:
            opt.zero_grad()
            acc.backward(nn.functional.mse_loss(model(x), y))
            opt.step()
            if sched is not None:
                sched.step()
    return random.random()


def ab(model):
    m = getattr(model, "module", model)
    return m.a.item(), m.b.item()


def test_total_limit_keeps_only_newest_checkpoints(tmp_path):
    set_seed(42)
    acc = Accelerator(cpu=True, project_config=ProjectConfiguration(project_dir=str(tmp_path), automatic_checkpoint_naming=True, total_limit=1))
    model = Line()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    model, opt, dl = acc.prepare(model, opt, loaders())
    acc.save_state()
    acc.save_state()
    assert os.listdir(tmp_path / "checkpoints") == ["checkpoint_1"]


def test_resume_reproduces_parameters_optimizer_and_rng(tmp_path):
    set_seed(42)
    acc = Accelerator(cpu=True)
    model = Line()
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    model, opt, dl = acc.prepare(model, opt, loaders())
    train(acc, model, opt, dl)
    first = str(tmp_path / "initial")
    acc.save_state(first)
    a1, b1 = ab(model)
    import copy

    opt_state1 = copy.deepcopy(opt.state_dict())
    train(acc, model, opt, dl)
    assert ab(model) != (a1, b1)
    acc.load_state(first)
    assert ab(model) == (a1, b1)
    s = opt.state_dict()
    assert all(torch.equal(s["state"][k]["exp_avg"], opt_state1["state"][k]["exp_avg"]) for k in s["state"])
    # rng: the draw after loading equals the draw made right after saving
    acc.save_state(first)
    drawn = (random.random(), np.random.rand(), torch.rand(1).item())
    acc.load_state(first)
    assert (random.random(), np.random.rand(), torch.rand(1).item()) == drawn


def test_file_layout_and_safe_serialization_switch(tmp_path):
    acc = Accelerator(cpu=True)
    model = Line()
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    sched = torch.optim.lr_scheduler.StepLR(opt, 1)
    model, opt, dl, sched = acc.prepare(model, opt, loaders(), sched)
    acc.save_state(str(tmp_path / "safe"))
    assert sorted(os.listdir(tmp_path / "safe")) == ["model.safetensors", "optimizer.bin", "random_states_0.pkl", "scheduler.bin"]
    acc.save_state(str(tmp_path / "unsafe"), safe_serialization=False)
    assert "pytorch_model.bin" in os.listdir(tmp_path / "unsafe") and "model.safetensors" not in os.listdir(tmp_path / "unsafe")
    acc.load_state(str(tmp_path / "unsafe"))


def test_automatic_naming_counts_up_and_load_picks_latest(tmp_path):
    set_seed(1)
    acc = Accelerator(cpu=True, project_config=ProjectConfiguration(project_dir=str(tmp_path), automatic_checkpoint_naming=True))
    model = Line()
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    model, opt, dl = acc.prepare(model, opt, loaders())
    acc.save_state()
    train(acc, model, opt, dl)
    acc.save_state()
    latest = ab(model)
    train(acc, model, opt, dl)
    acc.load_state()
    assert ab(model) == latest and sorted(os.listdir(tmp_path / "checkpoints")) == ["checkpoint_0", "checkpoint_1"]
    assert acc.project_configuration.iteration == 2
    with pytest.raises(ValueError):
        acc.save_state()  # checkpoint_2 does not exist yet -> fine; saving again into an existing one is refused
        acc.project_configuration.iteration = 1
        acc.save_state()


def test_load_state_from_missing_directory_fails(tmp_path):
    acc = Accelerator(cpu=True)
    with pytest.raises(ValueError):
        acc.load_state(str(tmp_path / "nope"))
    with pytest.raises(ValueError):
        acc.load_state()  # no input_dir and no automatic naming


def test_scheduler_state_is_restored(tmp_path):
    set_seed(3)
    acc = Accelerator(cpu=True)
    model = Line()
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.9)
    model, opt, dl, sched = acc.prepare(model, opt, loaders(), sched)
    train(acc, model, opt, dl, sched=sched)
    acc.save_state(str(tmp_path / "s"))
    saved = sched.state_dict()
    train(acc, model, opt, dl, sched=sched)
    assert sched.state_dict()["last_epoch"] != saved["last_epoch"]
    acc.load_state(str(tmp_path / "s"))
    assert sched.state_dict()["last_epoch"] == saved["last_epoch"] and sched.get_last_lr() == pytest.approx(saved["_last_lr"])


def test_custom_objects_round_trip_and_count_check(tmp_path):
    class Counter:
        def __init__(self):
            self.n = 0

        def state_dict(self):
            return {"n": self.n}

        def load_state_dict(self, sd):
            self.n = sd["n"]

    acc = Accelerator(cpu=True)
    c = Counter()
    acc.register_for_checkpointing(c)
    c.n = 5
    acc.save_state(str(tmp_path / "c"))
    assert "custom_checkpoint_0.pkl" in os.listdir(tmp_path / "c")
    c.n = 9
    acc.load_state(str(tmp_path / "c"))
    assert c.n == 5
    with pytest.raises(ValueError, match="state_dict"):
        acc.register_for_checkpointing(object())
    acc.register_for_checkpointing(Counter())
    with pytest.raises(RuntimeError):
        acc.load_state(str(tmp_path / "c"))  # two registered objects, one file


def test_pre_hooks_see_models_and_can_be_removed(tmp_path):
    acc = Accelerator(cpu=True)
    model = acc.prepare(Line())
    calls = []
    h1 = acc.register_save_state_pre_hook(lambda models, weights, out: calls.append(("save", len(models), len(weights), out)))
    h2 = acc.register_load_state_pre_hook(lambda models, inp: calls.append(("load", len(models), inp)))
    d = str(tmp_path / "h")
    acc.save_state(d)
    acc.load_state(d)
    assert calls == [("save", 1, 1, d), ("load", 1, d)]
    h1.remove()
    h2.remove()
    acc.save_state(d)
    assert len(calls) == 2


def test_map_location_and_strict_kwargs(tmp_path):
    acc = Accelerator(cpu=True)
    model = acc.prepare(Line())
    d = str(tmp_path / "m")
    acc.save_state(d)
    acc.load_state(d, map_location="cpu")
    with pytest.raises(TypeError):
        acc.load_state(d, map_location="gpu?")


def test_mid_epoch_resume_with_skip_first_batches(tmp_path):
    set_seed(5)
    acc = Accelerator(cpu=True)
    data = TensorDataset(torch.arange(40).float())
    dl = acc.prepare(DataLoader(data, batch_size=4))
    seen = [b[0].tolist() for b in dl]
    resumed = acc.skip_first_batches(dl, 6)
    rest = [b[0].tolist() for b in resumed]
    assert rest == seen[6:] and len(rest) == 4
    assert [b[0].tolist() for b in dl] == seen  # the original loader is untouched
