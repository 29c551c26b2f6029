"""Rank-aware logging and the kwargs-handler / plugin dataclasses (behaviour ports of the reference's tests/test_logging.py and
tests/test_kwargs_handlers.py)."""

import logging
import os

import pytest
import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.logging import MultiProcessAdapter, get_logger
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState
from accelerate_hpc_test_b200.utils import (
    AutocastKwargs,
    DataLoaderConfiguration,
    DistributedDataParallelKwargs,
    GradientAccumulationPlugin,
    GradScalerKwargs,
    InitProcessGroupKwargs,
    ProfileKwargs,
    ProjectConfiguration,
    TorchDynamoPlugin,
)
from accelerate_hpc_test_b200.utils.dataclasses import KwargsHandler
from accelerate_hpc_test_b200.utils.environment import patch_environment


@pytest.fixture(autouse=True)
def fresh_state():
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    yield
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


# --------------------------------------------------------------------------- logging
def test_logger_needs_an_initialised_state():
    log = get_logger("needs_state")
    with pytest.raises(RuntimeError, match="initialize the accelerate state"):
        log.info("too early")


def test_main_process_only_and_in_order(caplog):
    PartialState(cpu=True)
    log = get_logger("acc_test", log_level="INFO")
    assert isinstance(log, MultiProcessAdapter)
    with caplog.at_level(logging.INFO, logger="acc_test"):
        log.info("a")
        log.info("b", main_process_only=False)
        log.info("c", in_order=True)
        log.debug("hidden")
    assert [r.message for r in caplog.records] == ["a", "b", "c"]
    assert caplog.records[0].funcName == "test_main_process_only_and_in_order"  # stacklevel points at the caller


def test_non_main_process_is_silent(caplog, monkeypatch):
    PartialState(cpu=True)
    monkeypatch.setitem(PartialState._shared_state, "process_index", 1)
    log = get_logger("acc_test_rank1", log_level="INFO")
    with caplog.at_level(logging.INFO, logger="acc_test_rank1"):
        log.info("only on main")
        log.info("everyone", main_process_only=False)
    assert [r.message for r in caplog.records] == ["everyone"]


def test_warning_once(caplog):
    PartialState(cpu=True)
    log = get_logger("acc_once", log_level="WARNING")
    with caplog.at_level(logging.WARNING, logger="acc_once"):
        for _ in range(3):
            log.warning_once("same text")
        log.warning_once("other text")
    assert [r.message for r in caplog.records] == ["same text", "other text"]


def test_log_level_from_environment():
    PartialState(cpu=True)
    with patch_environment(accelerate_log_level="error"):
        log = get_logger("acc_env_level")
    assert log.logger.level == logging.ERROR
    assert get_logger("acc_explicit", log_level="debug").logger.level == logging.DEBUG


# --------------------------------------------------------------------------- kwargs handlers
def test_to_kwargs_reports_only_changed_fields():
    from dataclasses import dataclass

    @dataclass
    class Mock(KwargsHandler):
        a: int = 0
        b: bool = False
        c: float = 3.0

    assert Mock().to_kwargs() == {}
    assert Mock(a=2).to_kwargs() == {"a": 2}
    assert Mock(a=2, b=True).to_kwargs() == {"a": 2, "b": True}
    assert Mock(a=2, c=2.25).to_kwargs() == {"a": 2, "c": 2.25}


def test_grad_scaler_kwargs_reach_the_scaler():
    scaler_handler = GradScalerKwargs(init_scale=1024, growth_factor=2)
    assert scaler_handler.to_kwargs() == {"init_scale": 1024}  # growth_factor=2 is the default
    acc = Accelerator(cpu=True, mixed_precision="fp16", kwargs_handlers=[scaler_handler])
    assert acc.scaler is not None and acc.scaler._init_scale == 1024.0 and acc.scaler._growth_interval == 2000


def test_duplicate_handlers_are_rejected():
    with pytest.raises(ValueError):
        Accelerator(cpu=True, kwargs_handlers=[GradScalerKwargs(), GradScalerKwargs()])


def test_autocast_kwargs_disable_autocast():
    acc = Accelerator(cpu=True, mixed_precision="bf16", kwargs_handlers=[AutocastKwargs(enabled=False)])
    with acc.autocast():
        assert not torch.is_autocast_enabled("cpu")
    acc2_kwargs = AutocastKwargs(enabled=True, cache_enabled=False).to_kwargs()
    assert acc2_kwargs == {"cache_enabled": False}


def test_ddp_and_init_process_group_kwargs_defaults():
    ddp = DistributedDataParallelKwargs(bucket_cap_mb=15, find_unused_parameters=True)
    kw = ddp.to_kwargs()
    assert kw["bucket_cap_mb"] == 15 and kw["find_unused_parameters"] is True and "gradient_as_bucket_view" not in kw
    assert "comm_hook" not in DistributedDataParallelKwargs().to_kwargs()
    ipg = InitProcessGroupKwargs()
    assert ipg.backend in (None, "nccl") and ipg.timeout is not None


def test_profile_kwargs_build_a_profiler(tmp_path):
    kw = ProfileKwargs(activities=["cpu"], record_shapes=True, output_trace_dir=str(tmp_path))
    prof = kw.build()
    with prof:
        torch.ones(4).sum()
    assert prof.key_averages() is not None
    acc = Accelerator(cpu=True, kwargs_handlers=[kw])
    with acc.profile() as p:
        torch.ones(4).sum()
    assert any(f.startswith("profile_0") for f in os.listdir(tmp_path))
    assert p is not None


def test_gradient_accumulation_plugin_and_env(monkeypatch):
    acc = Accelerator(cpu=True, gradient_accumulation_plugin=GradientAccumulationPlugin(num_steps=3, sync_with_dataloader=False))
    assert acc.gradient_accumulation_steps == 3 and acc.gradient_state.sync_with_dataloader is False
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    with patch_environment(accelerate_gradient_accumulation_steps="5"):
        assert Accelerator(cpu=True).gradient_accumulation_steps == 5
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    with pytest.raises(ValueError):
        Accelerator(cpu=True, gradient_accumulation_steps=2, gradient_accumulation_plugin=GradientAccumulationPlugin(num_steps=3))


def test_project_configuration_directories(tmp_path):
    cfg = ProjectConfiguration(project_dir=str(tmp_path))
    assert cfg.logging_dir == str(tmp_path)
    cfg.set_directories(str(tmp_path / "other"))
    assert cfg.project_dir == str(tmp_path / "other") and cfg.logging_dir == str(tmp_path)  # an existing logging_dir is kept
    cfg2 = ProjectConfiguration(project_dir="p", logging_dir="l", automatic_checkpoint_naming=True, total_limit=2)
    assert cfg2.logging_dir == "l" and cfg2.iteration == 0


def test_dataloader_configuration_defaults():
    cfg = DataLoaderConfiguration()
    assert cfg.split_batches is False and cfg.dispatch_batches is None and cfg.even_batches is True and cfg.use_seedable_sampler is False
    acc = Accelerator(cpu=True, dataloader_config=DataLoaderConfiguration(split_batches=True, even_batches=False))
    assert acc.split_batches is True and acc.even_batches is False


def test_dynamo_plugin_reads_environment():
    with patch_environment(accelerate_dynamo_backend="inductor", accelerate_dynamo_mode="max-autotune", accelerate_dynamo_use_fullgraph="true",
                           accelerate_dynamo_use_dynamic="false", accelerate_dynamo_use_regional_compilation="true"):
        plug = TorchDynamoPlugin()
    kw = plug.to_kwargs()
    assert str(plug.backend).lower().endswith("inductor") and plug.mode == "max-autotune" and plug.fullgraph is True
    assert plug.use_regional_compilation is True and kw.get("fullgraph") is True
