"""Tracker contract and the package-backed integrations (behaviour ports of the reference's tests/test_tracking.py).

TensorBoard runs for real (the package is installed); the other back-ends are exercised against small stand-in modules
injected into ``sys.modules`` that record the calls the adapter makes."""

import json
import os
import sys
import types

import pytest
import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200 import tracking
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.tracking import (
    LOGGER_TYPE_TO_CLASS,
    ClearMLTracker,
    GeneralTracker,
    JSONLTracker,
    TensorBoardTracker,
    filter_trackers,
    get_available_trackers,
    on_main_process,
)
from accelerate_hpc_test_b200.utils.dataclasses import LoggerType


@pytest.fixture(autouse=True)
def fresh_state():
    yield
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


class Recorder:
    """Stand-in run object: remembers every call as (name, args, kwargs)."""

    def __init__(self, log):
        self._log = log
        self.config = types.SimpleNamespace(update=lambda *a, **k: log.append(("config.update", a, k)))

    def __getattr__(self, name):
        def call(*a, **k):
            self._log.append((name, a, k))
            return self

        return call


def fake_module(monkeypatch, name, **attrs):
    mod = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(mod, k, v)
    monkeypatch.setitem(sys.modules, name, mod)
    return mod


# --------------------------------------------------------------------------- contract
def test_registry_names_match_class_names():
    for key, cls in LOGGER_TYPE_TO_CLASS.items():
        assert cls.name == key and issubclass(cls, GeneralTracker)
    assert LoggerType.JSONL in get_available_trackers()


def test_tracker_without_required_attributes_is_rejected():
    class NoName(GeneralTracker):
        requires_logging_directory = False

        @property
        def tracker(self):
            return None

    with pytest.raises(NotImplementedError, match="`name`"):
        NoName()

    class Nothing(GeneralTracker):
        pass

    with pytest.raises(NotImplementedError) as err:
        Nothing()
    assert all(word in str(err.value) for word in ("`name`", "`requires_logging_directory`", "`tracker`"))
    GeneralTracker(_blank=True)  # the placeholder handed out on non-main processes


def test_on_main_process_respects_main_process_only():
    Accelerator(cpu=True)
    calls = []

    class T:
        main_process_only = True

        @on_main_process
        def hit(self, x):
            calls.append(x)
            return x

    assert T().hit(3) == 3 and calls == [3]


def test_filter_trackers_rules(tmp_path):
    Accelerator(cpu=True)
    assert filter_trackers(None) == []
    assert filter_trackers("jsonl", str(tmp_path)) == [LoggerType.JSONL]
    assert filter_trackers(["jsonl", "jsonl"], str(tmp_path)) == [LoggerType.JSONL]
    with pytest.raises(ValueError, match="requires a `logging_dir`"):
        filter_trackers("jsonl")
    with pytest.raises(ValueError, match="Unsupported logging capability"):
        filter_trackers("carrier_pigeon")
    mine = JSONLTracker("r", str(tmp_path))
    everything = filter_trackers(["all", mine], str(tmp_path))
    assert everything[0] is mine and LoggerType.JSONL in everything
    # a known integration whose package is missing is skipped, not an error
    assert filter_trackers(["aim"], str(tmp_path)) in ([], [LoggerType.AIM])


# --------------------------------------------------------------------------- JSONL + TensorBoard (real)
def test_jsonl_tracker_appends_one_object_per_call(tmp_path):
    Accelerator(cpu=True)
    t = JSONLTracker("run", str(tmp_path))
    t.store_init_configuration({"lr": 1e-3, "w": torch.tensor(2.0)})
    t.log({"loss": torch.tensor(0.5), "vec": torch.tensor([1, 2])}, step=1)
    t.log({"loss": 0.25})
    t.finish()
    rows = [json.loads(x) for x in open(tmp_path / "run" / "metrics.jsonl")]
    assert rows[0]["step"] == 1 and rows[0]["loss"] == 0.5 and rows[0]["vec"] == [1, 2] and rows[1]["step"] is None
    assert json.load(open(tmp_path / "run" / "hparams.json")) == {"lr": 1e-3, "w": 2.0}
    assert t.tracker is None


def test_tensorboard_writes_event_files_and_hparams(tmp_path):
    acc = Accelerator(cpu=True, log_with="tensorboard", project_dir=str(tmp_path))
    acc.init_trackers("tb_run", config={"num_iterations": 12, "learning_rate": 1e-2, "some_boolean": False, "some_string": "x"})
    acc.log({"total_loss": 0.1, "iteration": 1, "my_text": "some_value", "group": {"a": 1.0, "b": 2.0}}, step=0)
    tb = acc.get_tracker("tensorboard")
    assert isinstance(tb, TensorBoardTracker) and tb.tracker is not None
    tb.log_images({"imgs": torch.rand(2, 3, 4, 4)}, step=0)
    acc.end_training()
    run_dir = tmp_path / "tb_run"
    assert any(f.startswith("events.out.tfevents") for f in os.listdir(run_dir))
    stamped = [d for d in os.listdir(run_dir) if (run_dir / d / "hparams.yml").is_file()]
    assert len(stamped) == 1


def test_tensorboard_needs_a_logging_dir():
    with pytest.raises(ValueError, match="logging_dir"):
        Accelerator(cpu=True, log_with="tensorboard")


# --------------------------------------------------------------------------- package-backed adapters against stand-ins
def test_wandb_adapter_calls(monkeypatch):
    Accelerator(cpu=True)
    seen = []
    fake_module(monkeypatch, "wandb", init=lambda **k: (seen.append(("init", (), k)), Recorder(seen))[1],
                Image=lambda x: ("img", x), Table=lambda **k: ("table", k))
    t = tracking.WandBTracker("proj", entity="me")
    t.start()
    t.store_init_configuration({"lr": 1})
    t.log({"loss": 1.0}, step=5)
    t.log_images({"pics": [1, 2]}, step=5)
    t.log_table("tab", columns=["a"], data=[[1]])
    t.finish()
    names = [c[0] for c in seen]
    assert names == ["init", "config.update", "log", "log", "log", "finish"]
    assert seen[0][2] == {"project": "proj", "entity": "me"} and seen[1][2] == {"allow_val_change": True}
    assert seen[2][1] == ({"loss": 1.0},) and seen[2][2] == {"step": 5}
    assert seen[3][1][0] == {"pics": [("img", 1), ("img", 2)]}


def test_mlflow_adapter_filters_and_batches(monkeypatch):
    Accelerator(cpu=True)
    seen = []
    validation = types.SimpleNamespace(MAX_PARAM_VAL_LENGTH=10)
    fake_module(monkeypatch, "mlflow", utils=types.SimpleNamespace(validation=validation),
                search_experiments=lambda filter_string: [], create_experiment=lambda **k: "exp1",
                start_run=lambda **k: (seen.append(("start_run", k)), "run")[1], log_params=lambda d: seen.append(("log_params", d)),
                log_metrics=lambda d, step=None: seen.append(("log_metrics", d, step)), end_run=lambda: seen.append(("end_run",)))
    t = tracking.MLflowTracker(experiment_name="e", run_name="r")
    t.start()
    params = {f"p{i}": i for i in range(150)}
    params["too_long"] = "x" * 50
    t.store_init_configuration(params)
    t.log({"loss": 1.5, "note": "text", "acc": torch.tensor(0.5)}, step=2)
    t.finish()
    batches = [c[1] for c in seen if c[0] == "log_params"]
    assert [len(b) for b in batches] == [100, 50] and all("too_long" not in b for b in batches)
    assert ("log_metrics", {"loss": 1.5, "acc": 0.5}, 2) in seen and seen[-1] == ("end_run",)
    assert seen[0][1]["experiment_id"] == "exp1" and seen[0][1]["run_name"] == "r"


def test_clearml_title_series_split_and_single_values(monkeypatch):
    Accelerator(cpu=True)
    seen = []
    sink = Recorder(seen)
    task = types.SimpleNamespace(get_logger=lambda: sink, connect_configuration=lambda v: seen.append(("connect", v)),
                                 close=lambda: seen.append(("close",)))
    Task = types.SimpleNamespace(current_task=lambda: None, init=lambda **k: (seen.append(("init", k)), task)[1])
    fake_module(monkeypatch, "clearml", Task=Task)
    t = ClearMLTracker("proj")
    t.start()
    t.store_init_configuration({"a": 1})
    t.log({"eval_loss": 0.5, "loss": 0.7, "skip": "text"}, step=3)
    t.log({"final_acc": 0.9})
    t.finish()
    assert ClearMLTracker._title_series("test_f1") == ("f1", "test") and ClearMLTracker._title_series("f1") == ("f1", "train")
    scalars = [c for c in seen if c[0] == "report_scalar"]
    assert {(c[2]["title"], c[2]["series"]) for c in scalars} == {("loss", "eval"), ("loss", "train")}
    assert any(c[0] == "report_single_value" and c[2]["name"] == "final_acc" for c in seen)
    assert seen[0][1]["project_name"] == "proj" and seen[-1] == ("close",)


def test_dvclive_and_comet_and_aim_adapters(monkeypatch):
    Accelerator(cpu=True)
    seen = []
    live = Recorder(seen)
    t = tracking.DVCLiveTracker(live=live)
    t.start()
    t.store_init_configuration({"lr": 1})
    t.log({"loss": 1.0, "bad": [1, 2]}, step=7)
    t.finish()
    assert live.step == 7 and [c[0] for c in seen] == ["log_params", "log_metric", "next_step", "end"]

    seen.clear()
    fake_module(monkeypatch, "comet_ml", start=lambda **k: (seen.append(("start", k)), Recorder(seen))[1])
    c = tracking.CometMLTracker("proj")
    c.start()
    c.log({"loss": 1.0, "tag": "x", "grp": {"a": 1}}, step=4)
    c.finish()
    assert [x[0] for x in seen] == ["start", "set_step", "log_metric", "log_other", "log_metrics", "end"]

    seen.clear()

    class Run(dict):
        def __init__(self, repo=None, **k):
            seen.append(("Run", repo))

        def track(self, value, name=None, step=None, **k):
            seen.append(("track", name, value, step))

        def close(self):
            seen.append(("close",))

    fake_module(monkeypatch, "aim", Run=Run, Image=lambda img, caption="", **k: ("aimimg", img, caption))
    a = tracking.AimTracker("r", logging_dir="/tmp/aimrepo")
    a.start()
    a.store_init_configuration({"lr": 2})
    a.log({"loss": 0.3}, step=1)
    a.log_images({"pic": ("IMG", "cap")}, step=1)
    a.finish()
    assert a.tracker["hparams"] == {"lr": 2} and a.tracker.name == "r"
    assert ("track", "loss", 0.3, 1) in seen and ("track", "pic", ("aimimg", "IMG", "cap"), 1) in seen and seen[-1] == ("close",)


def test_accelerator_get_tracker_unwrap_and_blank(tmp_path):
    acc = Accelerator(cpu=True, log_with="jsonl", project_dir=str(tmp_path))
    acc.init_trackers("r")
    assert isinstance(acc.get_tracker("jsonl"), JSONLTracker)
    assert acc.get_tracker("jsonl", unwrap=True) is acc.get_tracker("jsonl").tracker
    with pytest.raises(ValueError):
        acc.get_tracker("wandb")
    acc.end_training()


def test_trackio_and_swanlab_adapters(monkeypatch):
    Accelerator(cpu=True)
    seen = []
    fake_module(monkeypatch, "trackio", init=lambda **k: (seen.append(("init", (), k)), Recorder(seen))[1])
    t = tracking.TrackioTracker("proj", space_id="me/space")
    t.start()
    t.store_init_configuration({"lr": 1e-3})
    t.log({"loss": 0.5}, step=3)                     # trackio keeps its own step counter: the step is not forwarded
    t.finish()
    assert [c[0] for c in seen] == ["init", "config.update", "log", "finish"]
    assert seen[0][2] == {"project": "proj", "space_id": "me/space"} and seen[2][1] == ({"loss": 0.5},) and seen[2][2] == {}
    assert t.tracker is not None and tracking.TrackioTracker.name == "trackio" and not tracking.TrackioTracker.requires_logging_directory

    seen.clear()
    class Config(dict):
        def update(self, values, **k):
            super().update(values)
            seen.append(("config.update", (values,), k))

    swan_config = Config()
    fake_module(monkeypatch, "swanlab", init=lambda **k: (seen.append(("init", (), k)), Recorder(seen))[1], config=swan_config,
                Image=lambda x: ("img", x))
    s = tracking.SwanLabTracker("proj", mode="offline")
    s.start()
    assert swan_config["FRAMEWORK"] == "accelerate" and seen[0][2] == {"project": "proj", "mode": "offline"}
    s.store_init_configuration({"epochs": 3})
    assert swan_config["epochs"] == 3 and seen[1][2] == {"allow_val_change": True}
    s.log({"acc": 0.9}, step=7)
    s.log_images({"samples": ["a", "b"]}, step=7)
    s.finish()
    assert seen[2] == ("log", ({"acc": 0.9},), {"step": 7})
    assert seen[3][1][0] == {"samples": [("img", "a"), ("img", "b")]} and seen[3][2] == {"step": 7}
    assert seen[-1][0] == "finish"
