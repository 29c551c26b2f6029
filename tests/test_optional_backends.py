"""Adapters for optional libraries (DeepSpeed config objects, bitsandbytes plumbing, torchao / TransformerEngine guards): the parts that
run without those libraries installed."""

import base64
import json

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200.utils import ao, bnb, transformer_engine
from accelerate_hpc_test_b200.utils.deepspeed import DummyOptim, DummyScheduler, HfDeepSpeedConfig, map_pytorch_optim_to_deepspeed
from accelerate_hpc_test_b200.utils.imports import is_torchao_available, is_transformer_engine_available

ZERO3 = {"zero_optimization": {"stage": 3, "offload_param": {"device": "cpu"}, "overlap_comm": True, "contiguous_gradients": False},
         "bf16": {"enabled": True}, "train_batch_size": "auto"}


def test_deepspeed_config_sources(tmp_path):
    from_dict = HfDeepSpeedConfig(ZERO3)
    assert from_dict.config == ZERO3 and from_dict.config is not ZERO3
    path = tmp_path / "ds.json"
    path.write_text(json.dumps(ZERO3))
    assert HfDeepSpeedConfig(str(path)).config == ZERO3
    encoded = base64.urlsafe_b64encode(json.dumps(ZERO3).encode()).decode()
    assert HfDeepSpeedConfig(encoded).config == ZERO3
    with pytest.raises(ValueError, match="Expected a string path"):
        HfDeepSpeedConfig("definitely-not-a-config")


def test_deepspeed_config_paths_and_flags():
    cfg = HfDeepSpeedConfig(ZERO3)
    assert cfg.get_value("zero_optimization.stage") == 3 and cfg.get_value("zero_optimization.nothing", 7) == 7
    assert cfg.get_value("no.such.path") is None and cfg.get_value("train_batch_size") == "auto"
    assert cfg.is_zero3() and not cfg.is_zero2() and cfg.is_offload()
    assert cfg.is_true("zero_optimization.overlap_comm") and not cfg.is_false("zero_optimization.overlap_comm")
    assert cfg.is_false("zero_optimization.contiguous_gradients") and not cfg.is_true("zero_optimization.contiguous_gradients")
    assert not cfg.is_true("zero_optimization.absent") and not cfg.is_false("zero_optimization.absent")
    holder, leaf = cfg.find_config_node("zero_optimization.offload_param.device")
    assert holder == {"device": "cpu"} and leaf == "device"
    cfg.del_config_sub_tree("zero_optimization.offload_param")
    assert "offload_param" not in cfg.config["zero_optimization"]
    cfg.del_config_sub_tree("zero_optimization.offload_param")  # absent: fine unless it must exist
    with pytest.raises(ValueError, match="Can't find"):
        cfg.del_config_sub_tree("zero_optimization.offload_param", must_exist=True)
    cfg.set_stage_and_offload()
    assert not cfg.is_offload()
    stage2 = HfDeepSpeedConfig({"zero_optimization": {"stage": 2, "offload_optimizer": {"device": "nvme"}}})
    assert stage2.is_zero2() and stage2.is_offload()
    assert not HfDeepSpeedConfig({"zero_optimization": {"stage": 1, "offload_optimizer": {"device": "cpu"}}}).is_offload()


def test_dummy_optimizer_and_scheduler_materialize():
    layer = nn.Linear(4, 4)
    placeholder = DummyOptim(layer.parameters(), lr=0.5, weight_decay=0.1, betas=(0.8, 0.9), unknown=1)
    opt = placeholder.materialize()
    group = opt.param_groups[0]
    assert group["lr"] == 0.5 and group["weight_decay"] == 0.1 and tuple(group["betas"]) == (0.8, 0.9)
    sched = DummyScheduler(opt, total_num_steps=10, warmup_num_steps=2).materialize(opt)
    rates = []
    for _ in range(10):
        rates.append(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()
    assert rates[0] == 0.0 and rates[1] == 0.25 and rates[2] == 0.5 and rates[-1] < rates[3]
    custom = DummyScheduler(opt, lr_scheduler_callable=lambda o: torch.optim.lr_scheduler.ConstantLR(o, factor=1.0)).materialize(opt)
    assert isinstance(custom, torch.optim.lr_scheduler.ConstantLR)
    mapped = map_pytorch_optim_to_deepspeed(torch.optim.AdamW(layer.parameters(), lr=0.25, weight_decay=0.0))
    assert type(mapped).__name__ == "FusedAdamW" and mapped.param_groups[0]["lr"] == 0.25


class Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.embed = nn.Embedding(8, 16)
        self.body = nn.Sequential(nn.Linear(16, 32), nn.Linear(32, 16), nn.Linear(16, 10))
        self.lm_head = nn.Linear(16, 8, bias=False)


def test_bnb_exclusion_rules():
    model = Tiny()
    assert bnb.get_keys_to_not_convert(model) == ["lm_head"]
    model.lm_head.weight = model.embed.weight
    assert bnb.get_keys_to_not_convert(model) == ["embed", "lm_head"]
    assert bnb._excluded("model.lm_head.proj", ["lm_head"]) and bnb._excluded("lm_head", ["lm_head"])
    assert not bnb._excluded("model.lm_headless", ["lm_head"])
    assert not bnb.has_4bit_bnb_layers(model)
    cfg = bnb.BnbQuantizationConfig(load_in_8bit=True)
    with pytest.raises(ImportError, match="not compatible with 8bit"):
        bnb.load_and_quantize_model(model, cfg)


def test_ao_filters():
    model = Tiny()
    assert ao.find_first_last_linear_layers(model) == ("body.0", "lm_head")
    assert ao.find_first_last_linear_layers(nn.ReLU()) == (None, None)
    assert ao.filter_linear_layers(model.body[1], "body.1", ["body.0", "lm_head"])
    assert not ao.filter_linear_layers(model.body[0], "body.0", ["body.0", "lm_head"])
    assert not ao.filter_linear_layers(model.body[2], "body.2", [])  # 10 output features: off the fp8 tile
    assert ao.filter_linear_layers(nn.ReLU(), "act", [])
    assert not ao.has_ao_layers(model) or is_torchao_available()
    if not is_torchao_available():
        with pytest.raises(ImportError, match="torchao is not installed"):
            ao.convert_model_to_fp8_ao(model)


def test_transformer_engine_guards():
    model = Tiny()
    assert not transformer_engine.has_transformer_engine_layers(model) or is_transformer_engine_available()
    if not is_transformer_engine_available():
        with pytest.raises(ImportError, match="transformer_engine"):
            transformer_engine.convert_model(model)
        with pytest.raises(ImportError, match="transformer_engine"):
            transformer_engine.apply_fp8_autowrap(model, None)


def test_prepare_turns_deepspeed_placeholders_into_real_objects():
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.scheduler import AcceleratedScheduler

    acc = Accelerator(cpu=True)
    model = nn.Linear(4, 4)
    opt = DummyOptim(model.parameters(), lr=0.25, weight_decay=0.0)
    sched = DummyScheduler(opt, total_num_steps=8, warmup_num_steps=2)
    model, opt, sched = acc.prepare(model, opt, sched)
    assert type(opt.optimizer).__name__ == "FusedAdamW" and isinstance(sched, AcceleratedScheduler)
    seen = []
    for _ in range(4):
        loss = model(torch.ones(2, 4)).sum()
        acc.backward(loss)
        seen.append(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()
        opt.zero_grad()
    assert seen[0] == 0.0 and seen[1] == 0.125 and seen[2] == 0.25 and seen[3] < 0.25  # warm-up then linear decay
    with pytest.raises(ValueError, match="needs its optimizer"):
        acc.prepare(DummyScheduler(DummyOptim(model.parameters())))


def test_quantised_device_maps_and_8bit_offload(tmp_path, monkeypatch):
    import numpy as np

    from accelerate_hpc_test_b200.utils import modeling

    model = Tiny()
    cfg = bnb.BnbQuantizationConfig(load_in_4bit=True, skip_modules=["lm_head"], keep_in_fp32_modules=[])
    # a dict is checked: 4-bit layers may not be offloaded unless they are left un-quantised
    assert bnb.get_quantized_model_device_map(model, cfg, {"embed": 0, "body": 0, "lm_head": "cpu"}) == {"embed": 0, "body": 0, "lm_head": "cpu"}
    with pytest.raises(ValueError, match="dispatched on the CPU or the disk"):
        bnb.get_quantized_model_device_map(model, cfg, {"embed": 0, "body": "disk", "lm_head": 0})
    eight = bnb.BnbQuantizationConfig(load_in_8bit=True, skip_modules=["lm_head"], keep_in_fp32_modules=[])
    assert bnb.get_quantized_model_device_map(model, eight, {"embed": 0, "body": "cpu", "lm_head": 0})["body"] == "cpu"
    with pytest.raises(ValueError, match="please choose 'auto', 'balanced', 'balanced_low_0' or 'sequential'"):
        bnb.get_quantized_model_device_map(model, eight, "everywhere")
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="A GPU is needed for quantization"):
            bnb.get_quantized_model_device_map(model, eight)
    assert bnb.get_parameter_device(model.body) == torch.device("cpu")

    # already-quantised checkpoint weights (int8 data + SCB statistics) go to the offload folder as they are, the model keeps a meta tensor
    index = {}
    weight = torch.randint(-127, 127, (32, 16), dtype=torch.int8)
    scb = torch.rand(32)
    bnb.quantize_and_offload_8bit(model, weight, "body.0.weight", torch.int8, str(tmp_path), index, scb)
    assert set(index) == {"body.0.weight", "body.0.SCB"} and model.body[0].weight.device.type == "meta"
    stored = np.memmap(tmp_path / "body.0.weight.dat", dtype=np.int8, mode="r", shape=(32, 16))
    assert np.array_equal(np.asarray(stored), weight.numpy())
    # a float weight is quantised by a pass through GPU 0 first: the placement helper is what does it
    seen = []

    def fake_place(module, name, device, value=None, dtype=None, **kwargs):
        seen.append((name, device))
        if device == 0:
            owner, leaf = modeling._owner_of(module, name)
            quantised = torch.nn.Parameter(torch.zeros(value.shape, dtype=torch.int8), requires_grad=False)
            quantised.SCB = torch.ones(value.shape[0])
            owner._parameters[leaf] = quantised

    monkeypatch.setattr(modeling, "set_module_tensor_to_device", fake_place)
    index = {}
    bnb.quantize_and_offload_8bit(model, torch.randn(16, 32), "body.1.weight", torch.float16, str(tmp_path), index, None)
    assert seen == [("body.1.weight", 0), ("body.1.weight", "meta")] and set(index) == {"body.1.weight", "body.1.SCB"}


def test_transformer_engine_adapter_through_a_stand_in(monkeypatch):
    import contextlib
    import enum
    import sys
    import types

    entered = []

    class TELinear(nn.Linear):
        def __init__(self, i, o, bias=True, params_dtype=torch.float32):
            super().__init__(i, o, bias=bias, dtype=params_dtype)

    class TELayerNorm(nn.LayerNorm):
        def __init__(self, width, eps=1e-5, params_dtype=torch.float32):
            super().__init__(width, eps=eps, dtype=params_dtype)

    class TETransformerLayer(nn.Module):
        pass

    @contextlib.contextmanager
    def fp8_autocast(enabled, fp8_recipe):
        entered.append((enabled, fp8_recipe))
        yield

    class Format(enum.Enum):
        E4M3 = 1
        HYBRID = 2

    pytorch = types.ModuleType("transformer_engine.pytorch")
    pytorch.Linear, pytorch.LayerNorm, pytorch.TransformerLayer, pytorch.fp8_autocast = TELinear, TELayerNorm, TETransformerLayer, fp8_autocast
    recipe = types.ModuleType("transformer_engine.common.recipe")
    recipe.Format = Format
    recipe.DelayedScaling = lambda **k: ("delayed", k)
    recipe.MXFP8BlockScaling = lambda **k: ("mxfp8", k)
    common = types.ModuleType("transformer_engine.common")
    common.recipe = recipe
    root = types.ModuleType("transformer_engine")
    root.pytorch, root.common = pytorch, common
    for name, module in {"transformer_engine": root, "transformer_engine.pytorch": pytorch, "transformer_engine.common": common,
                         "transformer_engine.common.recipe": recipe}.items():
        monkeypatch.setitem(sys.modules, name, module)
    monkeypatch.setattr(transformer_engine, "is_transformer_engine_available", lambda: True)

    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(32, 64), nn.LayerNorm(64), nn.Sequential(nn.Linear(64, 16, bias=False)), nn.Linear(16, 10))
    x = torch.randn(3, 32)
    before = model(x)
    transformer_engine.convert_model(model)
    assert isinstance(model[0], TELinear) and isinstance(model[1], TELayerNorm) and isinstance(model[2][0], TELinear)
    assert model[2][0].bias is None and type(model[3]) is nn.Linear          # 10 outputs: off the fp8 tile, left as it is
    assert torch.equal(model(x), before) and transformer_engine.has_transformer_engine_layers(model)     # weights came along
    transformer_engine.convert_model(model, to_transformer_engine=False)
    assert type(model[0]) is nn.Linear and type(model[1]) is nn.LayerNorm and torch.equal(model(x), before)
    assert not transformer_engine.has_transformer_engine_layers(model)

    from accelerate_hpc_test_b200.utils import TERecipeKwargs

    wrapped = transformer_engine.apply_fp8_autowrap(model, TERecipeKwargs(fp8_format="E4M3", amax_history_len=16, use_autocast_during_eval=False))
    wrapped.train()
    wrapped(x)
    wrapped.eval()
    wrapped(x)
    assert [flag for flag, _ in entered] == [True, False]                    # on while training, off in eval unless asked
    kind, options = entered[0][1]
    assert kind == "delayed" and options["fp8_format"] is Format.E4M3 and options["amax_history_len"] == 16     # only what differs from TE's defaults
    entered.clear()
    block = transformer_engine.apply_fp8_autowrap(nn.Linear(16, 16), TERecipeKwargs(fp8_format="E4M3", use_mxfp8_block_scaling=True,
                                                                                    use_autocast_during_eval=True))
    block.eval()
    block(torch.randn(2, 16))
    assert entered[0][0] is True and entered[0][1][0] == "mxfp8" and entered[0][1][1]["fp8_format"] is Format.E4M3


def test_torchao_adapter_through_a_stand_in(monkeypatch):
    import sys
    import types

    class Float8Linear(nn.Linear):
        pass

    converted = {}

    def convert_to_float8_training(model, module_filter_fn=None, config=None):
        converted["config"] = config
        converted["names"] = []
        for name, module in list(model.named_modules()):
            if isinstance(module, nn.Linear) and module_filter_fn(module, name):
                converted["names"].append(name)
                owner = model.get_submodule(name.rpartition(".")[0]) if "." in name else model
                setattr(owner, name.rpartition(".")[2], Float8Linear(module.in_features, module.out_features, bias=module.bias is not None))

    float8 = types.ModuleType("torchao.float8")
    float8.convert_to_float8_training = convert_to_float8_training
    float8_linear = types.ModuleType("torchao.float8.float8_linear")
    float8_linear.Float8Linear = Float8Linear
    float8.float8_linear = float8_linear
    root = types.ModuleType("torchao")
    root.float8 = float8
    for name, module in {"torchao": root, "torchao.float8": float8, "torchao.float8.float8_linear": float8_linear}.items():
        monkeypatch.setitem(sys.modules, name, module)
    monkeypatch.setattr(ao, "is_torchao_available", lambda: True)

    model = nn.Sequential(nn.Linear(16, 32), nn.Linear(32, 32), nn.Linear(32, 24), nn.Linear(32, 32), nn.Linear(32, 16))
    ao.convert_model_to_fp8_ao(model, config="recipe")
    # the first and the last linear stay precise, so does the one whose width is off the fp8 tile
    assert converted == {"config": "recipe", "names": ["1", "3"]} and ao.has_ao_layers(model)
    assert type(model[0]) is nn.Linear and type(model[2]) is nn.Linear and type(model[4]) is nn.Linear
    custom = nn.Sequential(nn.Linear(16, 16), nn.Linear(16, 16))
    ao.convert_model_to_fp8_ao(custom, module_filter_func=lambda module, fqn: fqn == "0")
    assert converted["names"] == ["0"]
    with pytest.raises(NotImplementedError, match="resolved by `convert_model_to_fp8_ao`"):
        ao.filter_first_and_last_linear_layers(custom[0], "0")
