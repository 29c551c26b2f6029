"""DeepSpeed-style configurations on the native engines: where the plugin's values come from (fields, file, environment), how ``auto``
entries are filled, which combinations are refused, and what ``prepare`` resolves from the objects it is handed.  Behaviour of reference
tests/deepspeed/test_deepspeed.py (plugin / autofill / assertion tests), without DeepSpeed itself."""

import json

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200 import debug_launcher
from accelerate_hpc_test_b200.utils import DeepSpeedPlugin, patch_environment
from accelerate_hpc_test_b200.utils.deepspeed import HfDeepSpeedConfig

ZERO3_AUTO = {
    "bf16": {"enabled": "auto"},
    "fp16": {"enabled": "auto"},
    "optimizer": {"type": "AdamW", "params": {"lr": "auto", "weight_decay": "auto", "betas": [0.9, 0.95], "eps": 1e-8}},
    "scheduler": {"type": "WarmupDecayLR", "params": {"warmup_min_lr": "auto", "warmup_max_lr": "auto", "warmup_num_steps": "auto",
                                                      "total_num_steps": "auto"}},
    "zero_optimization": {"stage": 3, "overlap_comm": True, "reduce_bucket_size": "auto", "stage3_prefetch_bucket_size": "auto",
                          "stage3_param_persistence_threshold": "auto", "stage3_gather_16bit_weights_on_model_save": "auto",
                          "offload_optimizer": {"device": "none"}, "offload_param": {"device": "cpu"}},
    "gradient_accumulation_steps": "auto", "gradient_clipping": "auto", "train_batch_size": "auto", "train_micro_batch_size_per_gpu": "auto",
}


def test_fields_and_environment_write_the_config():
    plugin = DeepSpeedPlugin(zero_stage=1, gradient_accumulation_steps=2, gradient_clipping=0.5, offload_optimizer_device="cpu")
    config = plugin.deepspeed_config
    assert config["zero_optimization"]["stage"] == 1 and config["gradient_accumulation_steps"] == 2 and config["gradient_clipping"] == 0.5
    assert config["zero_optimization"]["offload_optimizer"] == {"device": "cpu", "nvme_path": None}
    assert config["train_batch_size"] == config["train_micro_batch_size_per_gpu"] == "auto" and config["steps_per_print"] == float("inf")
    assert isinstance(plugin.hf_ds_config, HfDeepSpeedConfig) and plugin.hf_ds_config.config is config
    with patch_environment(accelerate_deepspeed_zero_stage="3", accelerate_gradient_accumulation_steps="4", accelerate_gradient_clipping="1.5",
                           accelerate_deepspeed_offload_param_device="nvme", accelerate_deepspeed_offload_param_nvme_path="/mnt/nvme",
                           accelerate_deepspeed_zero3_save_16bit_model="true"):
        plugin = DeepSpeedPlugin()
    zero = plugin.deepspeed_config["zero_optimization"]
    assert (plugin.zero_stage, plugin.gradient_accumulation_steps, plugin.gradient_clipping) == (3, 4, 1.5) and plugin.zero3_init_flag is True
    assert zero["offload_param"] == {"device": "nvme", "nvme_path": "/mnt/nvme"} and zero["stage3_gather_16bit_weights_on_model_save"] is True
    # the init flag only means something at stage 3
    with pytest.warns(UserWarning, match="only applicable for ZeRO Stage 3"):
        assert DeepSpeedPlugin(zero_stage=2, zero3_init_flag=True).zero3_init_flag is False


def test_a_config_file_rules_and_the_fields_follow_it(tmp_path):
    path = tmp_path / "zero3.json"
    path.write_text(json.dumps(ZERO3_AUTO))
    plugin = DeepSpeedPlugin(hf_ds_config=str(path), gradient_accumulation_steps=8)
    assert plugin.zero_stage == 3 and plugin.offload_param_device == "cpu" and plugin.offload_optimizer_device == "none"
    assert plugin.deepspeed_config["gradient_accumulation_steps"] == 8 and plugin.is_auto("gradient_clipping")      # nothing stated yet
    assert plugin.is_auto("train_batch_size") and not plugin.is_auto("zero_optimization.stage") and not plugin.is_auto("no.such.key")
    assert plugin.get_value("optimizer.params.betas") == [0.9, 0.95] and plugin.get_value("missing.key", 7) == 7
    assert plugin.to_fsdp_plugin().reshard_after_forward is True and plugin.zero3_init_flag is True
    # the same through the environment variable the launcher sets
    with patch_environment(accelerate_deepspeed_config_file=str(path)):
        assert DeepSpeedPlugin().zero_stage == 3
    # a dict is copied, an HfDeepSpeedConfig is used as it is; ZeRO settings are mandatory; accumulation defaults to 1
    given = {"zero_optimization": {"stage": 2}}
    plugin = DeepSpeedPlugin(hf_ds_config=given)
    assert plugin.deepspeed_config is not given and plugin.deepspeed_config["gradient_accumulation_steps"] == 1 and plugin.zero_stage == 2
    ready = HfDeepSpeedConfig({"zero_optimization": {"stage": 1}})
    assert DeepSpeedPlugin(hf_ds_config=ready).hf_ds_config is ready
    with pytest.raises(ValueError, match="specify the ZeRO optimization config"):
        DeepSpeedPlugin(hf_ds_config={"bf16": {"enabled": True}})
    # settings that the accelerate config file also carried are refused rather than silently ignored
    with patch_environment(accelerate_config_ds_fields="deepspeed_config_file,zero_stage"):
        with pytest.raises(ValueError, match="following accelerate config variables will be ignored"):
            DeepSpeedPlugin(hf_ds_config=str(path))
    with patch_environment(accelerate_config_ds_fields="deepspeed_config_file,zero3_init_flag"):
        DeepSpeedPlugin(hf_ds_config=str(path))


def test_auto_entries_are_filled_and_stated_ones_compared():
    plugin = DeepSpeedPlugin(hf_ds_config=json.loads(json.dumps(ZERO3_AUTO)))
    with pytest.raises(ValueError, match="`optimizer.params.lr` not found in kwargs"):
        plugin.fill_match("optimizer.params.lr")
    plugin.fill_match("optimizer.params.lr", **{"optimizer.params.lr": 3e-4})
    assert plugin.get_value("optimizer.params.lr") == 3e-4
    plugin.fill_match("absent.path", **{"absent.path": 1})                     # unknown paths are ignored
    found = []
    plugin.fill_match("optimizer.params.lr", found, **{"optimizer.params.lr": 1e-3})
    assert found == ["- ds optimizer.params.lr=0.0003 vs arg optimizer.params.lr=0.001"]
    plugin.fill_match("optimizer.params.lr", found := [], must_match=False, **{"optimizer.params.lr": 1e-3})
    assert found == []
    known = {"optimizer.params.weight_decay": 0.1, "scheduler.params.warmup_min_lr": 0, "scheduler.params.warmup_max_lr": 3e-4,
             "scheduler.params.warmup_num_steps": 10, "scheduler.params.total_num_steps": 100, "zero_optimization.reduce_bucket_size": 16,
             "zero_optimization.stage3_prefetch_bucket_size": 14, "zero_optimization.stage3_param_persistence_threshold": 40,
             "zero_optimization.stage3_gather_16bit_weights_on_model_save": False, "gradient_clipping": 1.0, "train_batch_size": 16,
             "train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 4, "bf16.enabled": True, "fp16.enabled": False}
    plugin.deepspeed_config_process(must_match=False, **known)
    flat = json.dumps(plugin.deepspeed_config, default=str)
    assert '"auto"' not in flat and plugin.get_value("train_batch_size") == 16
    with pytest.raises(ValueError, match="mismatch kwargs") as raised:
        plugin.deepspeed_config_process(**{**known, "train_batch_size": 32, "optimizer.params.lr": 1.0})
    assert "ds train_batch_size=16 vs arg train_batch_size=32" in str(raised.value) and "optimizer.params.lr" in str(raised.value)


def test_precision_sections():
    plugin = DeepSpeedPlugin(hf_ds_config=json.loads(json.dumps(ZERO3_AUTO)))
    assert plugin.stated_precision() is None
    plugin.set_mixed_precision("bf16")
    assert plugin.deepspeed_config["bf16"]["enabled"] is True and plugin.deepspeed_config["fp16"]["enabled"] is False
    assert plugin.stated_precision() == "bf16"
    bare = DeepSpeedPlugin(zero_stage=2)
    bare.set_mixed_precision("fp16")
    assert bare.deepspeed_config["fp16"] == {"enabled": True, "auto_cast": True} and bare.deepspeed_config["bf16"] == {"enabled": False}
    none = DeepSpeedPlugin(zero_stage=2)
    none.set_mixed_precision("no")
    assert none.deepspeed_config["fp16"] == {"enabled": False} and none.deepspeed_config["bf16"] == {"enabled": False}
    stated = DeepSpeedPlugin(hf_ds_config={"zero_optimization": {"stage": 2}, "fp16": {"enabled": True}})
    with pytest.raises(ValueError, match="cannot be set to `bf16` when `fp16` is set"):
        stated.set_mixed_precision("bf16")
    stated.set_mixed_precision("fp16")


def test_selection_init_flag_context_and_moe_leaves():
    plugin = DeepSpeedPlugin(hf_ds_config=json.loads(json.dumps(ZERO3_AUTO)), transformer_moe_cls_names="Experts")
    assert plugin.selected is False
    with pytest.raises(ValueError, match="must be enabled manually"):
        plugin.select()
    with pytest.raises(NotImplementedError):
        plugin.selected = True
    plugin.select(_from_accelerator_state=True)
    assert plugin.selected and plugin.dschf.get_value("train_micro_batch_size_per_gpu") == 1 and "train_batch_size" not in plugin.dschf.config
    assert plugin.is_zero3_init_enabled()
    with plugin.zero3_init_context_manager(enable=False):
        assert not plugin.is_zero3_init_enabled()
    assert plugin.is_zero3_init_enabled()

    class Experts(nn.Module):
        def __init__(self):
            super().__init__()
            self.experts = nn.ModuleList(nn.Linear(4, 4) for _ in range(3))

    model = nn.Sequential(nn.Linear(4, 4), Experts(), Experts())
    assert plugin.set_moe_leaf_modules(model) == [Experts]
    assert [getattr(m, "_z3_leaf", False) for m in model] == [False, True, True]
    plugin.transformer_moe_cls_names = "Experts,Router"
    with pytest.raises(Exception, match="Could not find a transformer layer class called 'Router'"):
        plugin.set_moe_leaf_modules(model)
    with pytest.raises(NotImplementedError, match="MS-AMP is not supported for ZeRO Stage 3"):
        DeepSpeedPlugin(zero_stage=3, enable_msamp=True)
    assert DeepSpeedPlugin(zero_stage=2, enable_msamp=True, msamp_opt_level="O2").deepspeed_config["msamp"] == {"enabled": True, "opt_level": "O2"}


# ------------------------------------------------------------------------------------------------- prepare-level resolution (2 ranks)
class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.config = type("Config", (), {"hidden_size": 16})()
        self.body = nn.Sequential(nn.Linear(16, 16), nn.Tanh(), nn.Linear(16, 1))

    def forward(self, x):
        return self.body(x)


def _loader(batch_size):
    from torch.utils.data import DataLoader, TensorDataset

    g = torch.Generator().manual_seed(0)
    return DataLoader(TensorDataset(torch.randn(64, 16, generator=g), torch.randn(64, 1, generator=g)), batch_size=batch_size)


def _prepare_resolves():
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.state import AcceleratorState
    from accelerate_hpc_test_b200.utils import DummyOptim, DummyScheduler

    config = json.loads(json.dumps(ZERO3_AUTO))
    config["gradient_accumulation_steps"] = 2
    config["gradient_clipping"] = 0.05
    config["zero_optimization"]["offload_param"] = {"device": "none"}
    plugin = DeepSpeedPlugin(hf_ds_config=config)
    acc = Accelerator(deepspeed_plugin=plugin, gradient_accumulation_steps=1)
    assert acc.distributed_type.value == "DEEPSPEED" and acc.deepspeed_plugin is plugin and acc.mixed_precision == "no"
    model = _Net()
    opt = DummyOptim(model.parameters(), lr=1e-2, weight_decay=0.01)
    sched = DummyScheduler(opt, total_num_steps=40, warmup_num_steps=4)
    model, opt, train, evaluate, sched = acc.prepare(model, opt, _loader(8), _loader(4), sched)
    resolved = plugin.deepspeed_config
    assert '"auto"' not in json.dumps(resolved, default=str), resolved
    assert resolved["train_micro_batch_size_per_gpu"] == 4 and resolved["train_batch_size"] == 4 * 2 * 2     # smallest loader x steps x ranks
    assert acc.gradient_accumulation_steps == 2                                                                 # the configuration's value won
    zero = resolved["zero_optimization"]
    assert (zero["reduce_bucket_size"], zero["stage3_prefetch_bucket_size"], zero["stage3_param_persistence_threshold"]) == (256, 230, 160)
    assert resolved["optimizer"]["params"]["lr"] == 1e-2 and resolved["optimizer"]["params"]["weight_decay"] == 0.01
    assert resolved["scheduler"]["params"] == {"warmup_min_lr": 0, "warmup_max_lr": 1e-2, "warmup_num_steps": 4, "total_num_steps": 20}
    inner = opt.optimizer
    assert type(inner).__name__ == "FusedAdamW" and tuple(inner.param_groups[0]["betas"]) == (0.9, 0.95)
    # the configuration's clipping is applied at step boundaries: the update after the synchronising backward used clipped gradients
    snapshot = lambda: {k: v.detach().clone() for k, v in acc.get_state_dict(model).items()}  # noqa: E731 - gathered from the shards
    before = snapshot()
    for index, (x, y) in enumerate(train):
        with acc.accumulate(model):
            loss = nn.functional.mse_loss(model(x * 50), y)
            acc.backward(loss)
            norm = acc.clip_grad_norm_(model.parameters(), 1000.0)          # reports, does not clip again
            opt.step()
            sched.step()
            opt.zero_grad()
        if index == 3:      # two optimizer steps: the first one runs at the warm-up's learning rate of zero
            break
    assert norm is not None and float(norm) > 0.05
    after = snapshot()
    assert any(not torch.equal(before[k], after[k]) for k in before)

    # what is refused
    AcceleratorState._reset_state(reset_partial_state=False)
    acc = Accelerator(deepspeed_plugin=DeepSpeedPlugin(hf_ds_config=json.loads(json.dumps(ZERO3_AUTO))))
    net = _Net()
    with pytest.raises(ValueError, match="optimizer in the config file and in the code at the same time"):
        acc.prepare(net, torch.optim.AdamW(net.parameters()), _loader(4))
    with pytest.raises(ValueError, match="requires you to pass at least one of training or evaluation dataloaders"):
        acc.prepare(net, DummyOptim(net.parameters()))
    AcceleratorState._reset_state(reset_partial_state=False)
    acc = Accelerator(deepspeed_plugin=DeepSpeedPlugin(zero_stage=2))
    net = _Net()
    with pytest.raises(ValueError, match="cannot create a `DummyOptim` without specifying an optimizer in the config file"):
        acc.prepare(net, DummyOptim(net.parameters()), _loader(4))
    real = torch.optim.SGD(net.parameters(), lr=0.1)
    with pytest.raises(ValueError, match="Either specify a scheduler in the config file"):
        acc.prepare(net, real, _loader(4), DummyScheduler(real, total_num_steps=10))
    AcceleratorState._reset_state(reset_partial_state=False)
    acc = Accelerator(deepspeed_plugin=DeepSpeedPlugin(zero_stage=2))
    net = _Net()
    net, real, loader = acc.prepare(net, torch.optim.SGD(net.parameters(), lr=0.1), _loader(4))
    assert acc.deepspeed_plugin.deepspeed_config["zero_allow_untested_optimizer"] is True
    assert acc.deepspeed_plugin.deepspeed_config["gradient_clipping"] == 1.0          # `auto` clipping means 1.0, as under DeepSpeed


def _moe_blocks_stay_whole():
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    class Experts(nn.Module):
        def __init__(self):
            super().__init__()
            self.experts = nn.ModuleList(nn.Linear(8, 8) for _ in range(3))

        def forward(self, x):
            return self.experts[int(x.sum() > 0)](x)           # which expert runs depends on the data

    model = nn.Sequential(nn.Linear(8, 8), Experts(), nn.Linear(8, 8))
    plugin = DeepSpeedPlugin(zero_stage=3, transformer_moe_cls_names="Experts")
    assert plugin.set_moe_leaf_modules(model) == [Experts]
    acc = Accelerator(fsdp_plugin=FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                                 transformer_cls_names_to_wrap=["Linear"]))
    engine = acc.prepare(model)
    units = [u.name for u in engine._units]
    assert "1" in units and not any(name.startswith("1.experts") for name in units), units
    assert {"0", "2"} <= set(units)
    x = torch.randn(4, 8) + (1.0 if acc.process_index else -1.0)      # the two ranks take different experts
    engine(x).sum().backward()


def test_marked_moe_blocks_are_single_units_two_processes():
    debug_launcher(_moe_blocks_stay_whole, num_processes=2)


def test_prepare_resolves_the_configuration_two_processes():
    debug_launcher(_prepare_resolves, num_processes=2)
