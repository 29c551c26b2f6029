"""Accelerator / state behaviour on CPU (modelled on reference tests/test_accelerator.py,
tests/test_state_checkpointing.py, tests/test_scheduler.py, tests/test_optimizer.py, tests/test_tracking.py)."""

import json
import os
import pickle

import pytest
import torch
from torch.utils.data import DataLoader, TensorDataset

from accelerate_hpc_test_b200 import Accelerator, AcceleratorState, PartialState
from accelerate_hpc_test_b200.data_loader import DataLoaderShard
from accelerate_hpc_test_b200.optimizer import AcceleratedOptimizer
from accelerate_hpc_test_b200.scheduler import AcceleratedScheduler
from accelerate_hpc_test_b200.test_utils.training import RegressionDataset, RegressionModel
from accelerate_hpc_test_b200.utils import (
    DistributedType,
    GradientAccumulationPlugin,
    ProjectConfiguration,
    patch_environment,
    set_seed,
)
from accelerate_hpc_test_b200.utils.dataclasses import DistributedDataParallelKwargs, FullyShardedDataParallelPlugin, GradScalerKwargs


def components():
    model = torch.nn.Linear(2, 4)
    optimizer = torch.optim.AdamW(model.parameters(), lr=1.0)
    scheduler = torch.optim.lr_scheduler.OneCycleLR(optimizer, max_lr=0.01, steps_per_epoch=2, epochs=1)
    train_dl = DataLoader(TensorDataset(torch.tensor([1, 2, 3])))
    valid_dl = DataLoader(TensorDataset(torch.tensor([4, 5, 6])))
    return model, optimizer, scheduler, train_dl, valid_dl


def test_state_is_a_shared_singleton_and_resets():
    a = Accelerator(cpu=True)
    assert PartialState().device.type == "cpu" and AcceleratorState().num_processes == 1
    assert a.distributed_type == DistributedType.NO and a.is_main_process and a.is_last_process
    with pytest.raises(ValueError):
        Accelerator(mixed_precision="bf16")  # state already fixed with "no"
    AcceleratorState._reset_state(True)
    with pytest.raises(AttributeError, match="has no attribute `distributed_type`"):
        a.state.distributed_type
    b = Accelerator(cpu=True, mixed_precision="bf16")
    assert b.mixed_precision == "bf16"


def test_env_configures_mixed_precision_and_cpu():
    with patch_environment(accelerate_mixed_precision="bf16", accelerate_use_cpu="true"):
        acc = Accelerator()
        assert acc.mixed_precision == "bf16" and acc.device.type == "cpu"


def test_prepare_returns_wrapped_objects_in_order_and_is_idempotent():
    acc = Accelerator(cpu=True)
    model, optimizer, scheduler, train_dl, valid_dl = components()
    pm, po, ps, pt, pv = acc.prepare(model, optimizer, scheduler, train_dl, valid_dl)
    assert isinstance(po, AcceleratedOptimizer) and isinstance(ps, AcceleratedScheduler)
    assert type(pt) is DataLoaderShard and isinstance(pt, DataLoader)
    for obj in (pm, po, ps, pt, pv):
        assert getattr(obj, "_is_accelerate_prepared", False)
    assert len(acc._models) == 1 and len(acc._optimizers) == 1 and len(acc._schedulers) == 1 and len(acc._dataloaders) == 2
    again = acc.prepare(pm, po, ps, pt, pv)
    assert all(x is y for x, y in zip(again, (pm, po, ps, pt, pv)))
    assert len(acc._models) == 1 and len(acc._dataloaders) == 2
    acc.free_memory()
    assert len(acc._models) == 0 and len(acc._optimizers) == 0 and len(acc._dataloaders) == 0


def test_kwargs_handler_only_reports_non_defaults():
    assert DistributedDataParallelKwargs(bucket_cap_mb=15, find_unused_parameters=True).to_kwargs() == {"bucket_cap_mb": 15, "find_unused_parameters": True}
    assert GradScalerKwargs(init_scale=1024, growth_factor=2).to_kwargs() == {"init_scale": 1024}
    with pytest.raises(TypeError):
        Accelerator(cpu=True, kwargs_handlers=[object()])


def test_gradient_accumulation_makes_optimizer_and_scheduler_noops():
    acc = Accelerator(cpu=True, gradient_accumulation_steps=2)
    model = RegressionModel()
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0 / (s + 1))
    dl = DataLoader(RegressionDataset(length=16), batch_size=4)
    model, opt, sched, dl = acc.prepare(model, opt, sched, dl)
    a_hist, lr_hist, sync_hist = [], [], []
    for batch in dl:
        with acc.accumulate(model):
            loss = torch.nn.functional.mse_loss(model(batch["x"]), batch["y"])
            acc.backward(loss)
            opt.step()
            sched.step()
            opt.zero_grad()
        a_hist.append(model.a.item())
        lr_hist.append(sched.get_last_lr()[0])
        sync_hist.append(acc.sync_gradients)
    assert sync_hist == [False, True, False, True]
    assert a_hist[0] == 0.0 and a_hist[1] != 0.0 and a_hist[2] == a_hist[1] and a_hist[3] != a_hist[2]
    assert lr_hist[0] == lr_hist[1] * 2 or lr_hist[0] > lr_hist[1]


def test_accumulate_syncs_at_end_of_dataloader():
    acc = Accelerator(cpu=True, gradient_accumulation_plugin=GradientAccumulationPlugin(num_steps=3))
    dl = acc.prepare(DataLoader(list(range(8)), batch_size=2))  # 4 batches: sync at 3 and at the end (4)
    model = acc.prepare(RegressionModel())
    syncs = []
    for _ in dl:
        with acc.accumulate(model):
            syncs.append(acc.sync_gradients)
    assert syncs == [False, False, True, True]


def test_backward_divides_by_accumulation_steps():
    acc = Accelerator(cpu=True, gradient_accumulation_steps=4)
    p = torch.nn.Parameter(torch.tensor(1.0))
    acc.backward(p * 8.0)
    assert p.grad.item() == 2.0


def test_clip_grad_norm_and_unwrap():
    acc = Accelerator(cpu=True, mixed_precision="bf16")
    model = acc.prepare(torch.nn.Linear(4, 4))
    out = model(torch.randn(2, 4))
    assert out.dtype == torch.float32  # bf16 autocast output is converted back
    acc.backward(out.sum() * 100)
    norm = acc.clip_grad_norm_(model.parameters(), 1.0)
    assert norm > 1.0
    total = torch.sqrt(sum((p.grad.float() ** 2).sum() for p in model.parameters()))
    assert abs(total.item() - 1.0) < 1e-3
    inner = acc.unwrap_model(model, keep_fp32_wrapper=False)
    assert inner(torch.randn(2, 4)).dtype == torch.float32 and not hasattr(inner, "_original_forward")
    pickle.dumps(inner)


def test_gather_for_metrics_drops_duplicated_tail():
    acc = Accelerator(cpu=True)
    dl = acc.prepare(DataLoader(list(range(10)), batch_size=4))
    seen = []
    for batch in dl:
        seen += acc.gather_for_metrics(batch).tolist()
    assert seen == list(range(10))
    assert acc.gather_for_metrics(["a", "b"], use_gather_object=True) == ["a", "b"]


def test_save_and_load_state_restores_everything(tmp_path):
    set_seed(0)
    acc = Accelerator(cpu=True)
    model = RegressionModel()
    opt = torch.optim.Adam(model.parameters(), lr=0.1)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.9)
    dl = DataLoader(RegressionDataset(length=32, seed=1), batch_size=8, shuffle=True)
    model, opt, sched, dl = acc.prepare(model, opt, sched, dl)

    class Counter:
        def __init__(self):
            self.n = 0

        def state_dict(self):
            return {"n": self.n}

        def load_state_dict(self, sd):
            self.n = sd["n"]

    counter = Counter()
    acc.register_for_checkpointing(counter)
    with pytest.raises(ValueError):
        acc.register_for_checkpointing(object())

    def epoch():
        for batch in dl:
            opt.zero_grad()
            acc.backward(torch.nn.functional.mse_loss(model(batch["x"]), batch["y"]))
            opt.step()
        sched.step()
        counter.n += 1
        return model.a.item(), model.b.item(), torch.rand(1).item()

    epoch()
    hook_calls = []
    acc.register_save_state_pre_hook(lambda models, weights, d: hook_calls.append(("save", len(models), len(weights))))
    h = acc.register_load_state_pre_hook(lambda models, d: hook_calls.append(("load", len(models))))
    ckpt = str(tmp_path / "ckpt")
    acc.save_state(ckpt)
    for name in ("model.safetensors", "optimizer.bin", "scheduler.bin", "random_states_0.pkl", "custom_checkpoint_0.pkl"):
        assert os.path.isfile(os.path.join(ckpt, name)), name
    first = epoch()
    acc.load_state(ckpt)
    assert counter.n == 1 and hook_calls == [("save", 1, 1), ("load", 1)]
    h.remove()
    second = epoch()
    assert first == second  # same parameters, same optimizer / scheduler state, same RNG stream


def test_automatic_checkpoint_naming_and_total_limit(tmp_path):
    acc = Accelerator(cpu=True, project_config=ProjectConfiguration(project_dir=str(tmp_path), automatic_checkpoint_naming=True, total_limit=2))
    model = acc.prepare(RegressionModel())
    for _ in range(3):
        acc.save_state()
    kept = sorted(os.listdir(tmp_path / "checkpoints"))
    assert kept == ["checkpoint_1", "checkpoint_2"] and acc.save_iteration == 3
    with torch.no_grad():
        model.a.fill_(5.0)
    acc.load_state()  # latest
    assert model.a.item() == 0.0
    with pytest.raises(ValueError):
        acc.project_configuration.iteration = 2
        acc.save_state()


def test_save_model_shards_and_index(tmp_path):
    acc = Accelerator(cpu=True)
    model = torch.nn.Sequential(torch.nn.Linear(64, 64), torch.nn.Linear(64, 64))
    acc.save_model(model, str(tmp_path / "one"))
    assert os.listdir(tmp_path / "one") == ["model.safetensors"]
    acc.save_model(model, str(tmp_path / "many"), max_shard_size="17KB")
    files = sorted(os.listdir(tmp_path / "many"))
    assert "model.safetensors.index.json" in files and len([f for f in files if f.endswith(".safetensors")]) == 2
    index = json.load(open(tmp_path / "many" / "model.safetensors.index.json"))
    assert set(index["weight_map"]) == set(model.state_dict())
    acc.save_model(model, str(tmp_path / "bin"), safe_serialization=False)
    assert os.listdir(tmp_path / "bin") == ["pytorch_model.bin"]


def test_scheduler_steps_num_processes_times_unless_split(monkeypatch):
    acc = Accelerator(cpu=True)
    model, optimizer, _, _, _ = components()
    sched = torch.optim.lr_scheduler.LambdaLR(optimizer, lambda s: 1.0)
    model, optimizer, sched = acc.prepare(model, optimizer, sched)
    monkeypatch.setitem(PartialState._shared_state, "num_processes", 4)
    optimizer.step()
    sched.step()
    assert sched.scheduler._step_count == 1 + 4
    sched.split_batches = True
    sched.step()
    assert sched.scheduler._step_count == 1 + 4 + 1


def test_jsonl_tracker_and_custom_tracker(tmp_path):
    from accelerate_hpc_test_b200.tracking import GeneralTracker

    class Mine(GeneralTracker):
        name = "mine"
        requires_logging_directory = False

        def __init__(self):
            super().__init__()
            self.logged = []

        @property
        def tracker(self):
            return self.logged

        def log(self, values, step=None, **kw):
            self.logged.append((step, values))

    mine = Mine()
    acc = Accelerator(cpu=True, log_with=["jsonl", mine], project_dir=str(tmp_path))
    acc.init_trackers("run", config={"lr": 0.1})
    acc.log({"loss": torch.tensor(1.5), "tokens_per_s": 10.0}, step=3)
    assert acc.get_tracker("mine", unwrap=True) == [(3, {"loss": torch.tensor(1.5), "tokens_per_s": 10.0})]
    acc.end_training()
    lines = open(tmp_path / "run" / "metrics.jsonl").read().strip().splitlines()
    rec = json.loads(lines[0])
    assert rec["step"] == 3 and rec["loss"] == 1.5 and json.load(open(tmp_path / "run" / "hparams.json")) == {"lr": 0.1}
    with pytest.raises(ValueError):
        Accelerator(cpu=True, log_with="tensorboard")  # needs a logging dir
    class Broken(GeneralTracker):
        pass

    with pytest.raises(NotImplementedError):
        Broken()


def test_split_between_processes_single_process_is_identity():
    state = PartialState()
    with state.split_between_processes([1, 2, 3]) as x:
        assert x == [1, 2, 3]
    calls = []
    state.on_main_process(lambda: calls.append(1))()
    state.on_process(lambda: calls.append(2), process_index=0)()
    assert calls == [1, 2]


def test_fsdp_plugin_env_parsing_and_mapping():
    with patch_environment(fsdp_version="2", fsdp_auto_wrap_policy="TRANSFORMER_BASED_WRAP", fsdp_transformer_cls_to_wrap="LlamaDecoderLayer,Foo",
                           fsdp_state_dict_type="FULL_STATE_DICT", fsdp_reshard_after_forward="false", fsdp_activation_checkpointing="true"):
        p = FullyShardedDataParallelPlugin()
    assert p.fsdp_version == 2 and p.auto_wrap_policy == "TRANSFORMER_BASED_WRAP" and p.transformer_cls_names_to_wrap == ["LlamaDecoderLayer", "Foo"]
    assert p.state_dict_type == "FULL_STATE_DICT" and p.reshard_after_forward is False and p.activation_checkpointing is True
    old = FullyShardedDataParallelPlugin(fsdp_version=1, sharding_strategy="SHARD_GRAD_OP")
    assert old.effective_reshard_after_forward is False
    assert FullyShardedDataParallelPlugin(fsdp_version=1, sharding_strategy=1).sharding_strategy == "FULL_SHARD"
    with pytest.raises(ValueError):
        FullyShardedDataParallelPlugin(fsdp_version=2, reshard_after_forward="FULL_SHARD")
    p.set_mixed_precision("bf16")
    assert p.mixed_precision_policy.param_dtype == torch.bfloat16
    with pytest.raises(ValueError):
        p.set_mixed_precision({"bogus": 1})


def test_parallelism_config_validation():
    from accelerate_hpc_test_b200 import ParallelismConfig

    pc = ParallelismConfig(dp_shard_size=4, tp_size=2)
    assert pc.total_size == 8 and pc._get_mesh() == (("dp_shard", "tp"), (4, 2)) and pc.fsdp_dim_names == ["dp_shard_cp"]
    with pytest.raises(ValueError, match="mutually exclusive"):
        ParallelismConfig(cp_size=2, sp_size=2)
    with pytest.raises(ValueError):
        ParallelismConfig(dp_replicate_size=2, tp_size=2)
    with patch_environment(parallelism_config_dp_shard_size="2", parallelism_config_cp_size="2"):
        assert ParallelismConfig().dp_shard_cp_dim_names == ["dp_shard", "cp"]


def test_optimizer_remap_after_sharding_reports_what_it_cannot_map():
    import types

    from accelerate_hpc_test_b200.accelerator import _remap_after_sharding

    a, b, c = (torch.nn.Parameter(torch.zeros(1)) for _ in range(3))
    new_a, new_b = (torch.nn.Parameter(torch.ones(1)) for _ in range(2))
    model = types.SimpleNamespace(config=types.SimpleNamespace(tie_word_embeddings=False))
    assert _remap_after_sharding({"x": a, "y": b}, {"x": new_a, "y": new_b}, model) == {id(a): new_a, id(b): new_b}
    with pytest.raises(KeyError, match="Parameters missing after FSDP2 wrapping") as raised:
        _remap_after_sharding({"x": a, "y": b, "some_other.weight": c}, {"x": new_a, "y": new_b}, model)
    assert "some_other.weight" in str(raised.value)
    model.config.tie_word_embeddings = True
    with pytest.raises(ValueError, match="FSDP2 mapping failed") as raised:
        _remap_after_sharding({"model.embed_tokens.weight": a, "lm_head.weight": c}, {"model.embed_tokens.weight": new_a}, model)
    assert "tied embeddings" in str(raised.value) and "lm_head.weight" in str(raised.value) and "tie_word_embeddings = False" in str(raised.value)
