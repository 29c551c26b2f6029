"""``FullyShardedDataParallelPlugin``: every field from its argument or from the launcher's ``FSDP_*`` environment, FSDP1 spellings
mapped onto FSDP2 meaning, wrap policies resolved against a model, mixed-precision policies, ignored modules by regular expression.
Behaviour of reference tests/fsdp/test_fsdp.py:68-472 (the plugin tests), on CPU."""

import pytest
import torch
import torch.nn as nn

from accelerate_hpc_test_b200 import debug_launcher
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin, patch_environment
from accelerate_hpc_test_b200.utils.constants import FSDP_AUTO_WRAP_POLICY, FSDP_BACKWARD_PREFETCH, FSDP_SHARDING_STRATEGY, FSDP_STATE_DICT_TYPE
from accelerate_hpc_test_b200.utils.dataclasses import MixedPrecisionPolicy


class Block(nn.Module):
    def __init__(self):
        super().__init__()
        self.q_proj, self.o_proj = nn.Linear(8, 8), nn.Linear(8, 8)

    def forward(self, x):
        return self.o_proj(torch.tanh(self.q_proj(x)))


class Net(nn.Module):
    _no_split_modules = ["Block"]

    def __init__(self):
        super().__init__()
        self.embed = nn.Linear(8, 8)
        self.layers = nn.ModuleList(Block() for _ in range(2))

    def forward(self, x):
        x = self.embed(x)
        for layer in self.layers:
            x = layer(x)
        return x


def test_sharding_strategy_by_name_number_and_environment():
    for index, name in enumerate(FSDP_SHARDING_STRATEGY, start=1):
        with patch_environment(fsdp_version="1", fsdp_sharding_strategy=name):
            assert FullyShardedDataParallelPlugin().sharding_strategy == name
        with patch_environment(fsdp_version="1", fsdp_sharding_strategy=str(index)):
            assert FullyShardedDataParallelPlugin().sharding_strategy == name
        assert FullyShardedDataParallelPlugin(fsdp_version=1, sharding_strategy=index).sharding_strategy == name
    with pytest.raises(ValueError, match="sharding_strategy must be one of"):
        FullyShardedDataParallelPlugin(fsdp_version=1, sharding_strategy="SHARD_EVERYTHING")
    # FSDP2: the strategy's meaning survives as reshard_after_forward
    expected = {"FULL_SHARD": True, "HYBRID_SHARD": True, "SHARD_GRAD_OP": False, "_HYBRID_SHARD_ZERO2": False}
    for name, reshard in expected.items():
        if name in FSDP_SHARDING_STRATEGY:
            assert FullyShardedDataParallelPlugin(fsdp_version=2, sharding_strategy=name).reshard_after_forward is reshard
    with patch_environment(fsdp_version="2", fsdp_reshard_after_forward="false"):
        plugin = FullyShardedDataParallelPlugin()
        assert plugin.fsdp_version == 2 and plugin.reshard_after_forward is False and plugin.state_dict_type == "SHARDED_STATE_DICT"
    assert FullyShardedDataParallelPlugin(fsdp_version=2, reshard_after_forward="True").reshard_after_forward is True
    with pytest.raises(ValueError, match="must be a bool with fsdp_version=2"):
        FullyShardedDataParallelPlugin(fsdp_version=2, reshard_after_forward="FULL_SHARD")
    with pytest.raises(ValueError, match="sharding-strategy name with fsdp_version=1"):
        FullyShardedDataParallelPlugin(fsdp_version=1, reshard_after_forward=True)
    with pytest.raises(ValueError, match="fsdp_version must be 1 or 2"):
        FullyShardedDataParallelPlugin(fsdp_version=3)


def test_backward_prefetch_and_the_fsdp1_only_switches():
    for index, name in enumerate(FSDP_BACKWARD_PREFETCH, start=1):
        wanted = None if name == "NO_PREFETCH" else name
        with patch_environment(fsdp_version="1", fsdp_backward_prefetch=name):
            assert FullyShardedDataParallelPlugin().backward_prefetch == wanted
        assert FullyShardedDataParallelPlugin(fsdp_version=1, backward_prefetch=str(index)).backward_prefetch == wanted
    assert FullyShardedDataParallelPlugin(fsdp_version=2, backward_prefetch="BACKWARD_PRE").backward_prefetch is None     # implicit in FSDP2
    with patch_environment(fsdp_version="1", fsdp_use_orig_params="true", fsdp_sync_module_states="true", fsdp_forward_prefetch="true",
                           fsdp_offload_params="true", fsdp_activation_checkpointing="true"):
        plugin = FullyShardedDataParallelPlugin()
    assert plugin.use_orig_params and plugin.sync_module_states and plugin.forward_prefetch and plugin.cpu_offload and plugin.activation_checkpointing
    with pytest.raises(ValueError, match="forward_prefetch is an FSDP1 option"):
        FullyShardedDataParallelPlugin(fsdp_version=2, forward_prefetch=True)
    assert FullyShardedDataParallelPlugin(fsdp_version=2).cpu_offload is False


def test_state_dict_type():
    for index, name in enumerate(FSDP_STATE_DICT_TYPE, start=1):
        with patch_environment(fsdp_version="1", fsdp_state_dict_type=name):
            assert FullyShardedDataParallelPlugin().state_dict_type == name
        assert FullyShardedDataParallelPlugin(fsdp_version=1, state_dict_type=str(index)).state_dict_type == name
    assert FullyShardedDataParallelPlugin(fsdp_version=1).state_dict_type == "FULL_STATE_DICT"
    assert FullyShardedDataParallelPlugin(fsdp_version=2).state_dict_type == "SHARDED_STATE_DICT"
    with pytest.raises(ValueError, match="LOCAL_STATE_DICT does not exist with fsdp_version=2"):
        FullyShardedDataParallelPlugin(fsdp_version=2, state_dict_type="LOCAL_STATE_DICT")
    with pytest.raises(ValueError, match="unknown state_dict_type"):
        FullyShardedDataParallelPlugin(fsdp_version=2, state_dict_type="HALF_STATE_DICT")
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2)
    plugin.set_state_dict_type("full_state_dict")
    assert plugin.state_dict_type == "FULL_STATE_DICT"


def test_auto_wrap_policies_resolve_against_the_model():
    model = Net()
    assert set(FSDP_AUTO_WRAP_POLICY) == {"TRANSFORMER_BASED_WRAP", "SIZE_BASED_WRAP", "NO_WRAP"}
    with patch_environment(fsdp_version="2", fsdp_auto_wrap_policy="TRANSFORMER_BASED_WRAP", fsdp_transformer_cls_to_wrap="Block"):
        plugin = FullyShardedDataParallelPlugin()
    assert plugin.transformer_cls_names_to_wrap == ["Block"]
    plugin.set_auto_wrap_policy(model)
    assert plugin.should_wrap(model.layers[0]) and not plugin.should_wrap(model.embed) and not plugin.should_wrap(model)
    # without names the model's own `_no_split_modules` are used; several names come comma separated
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap")
    plugin.set_auto_wrap_policy(model)
    assert plugin.should_wrap(model.layers[1])
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap", transformer_cls_names_to_wrap="Block, Linear")
    plugin.set_auto_wrap_policy(model)
    assert plugin.should_wrap(model.embed) and plugin.should_wrap(model.layers[0])
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap", transformer_cls_names_to_wrap=["Decoder"])
    with pytest.raises(ValueError, match="Could not find the transformer layer class Decoder"):
        plugin.set_auto_wrap_policy(model)
    # by size: a block has 144 parameters, the embedding 72
    with patch_environment(fsdp_version="2", fsdp_auto_wrap_policy="SIZE_BASED_WRAP", fsdp_min_num_params="100"):
        plugin = FullyShardedDataParallelPlugin()
    plugin.set_auto_wrap_policy(model)
    assert plugin.should_wrap(model.layers[0]) and not plugin.should_wrap(model.embed)
    with pytest.raises(ValueError, match="min_num_params must be > 0"):
        FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="size_based_wrap", min_num_params=0)
    # none, a callable, an unknown name
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="no_wrap")
    plugin.set_auto_wrap_policy(model)
    assert not plugin.should_wrap(model.layers[0])
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy=lambda module, *args, **kwargs: isinstance(module, nn.Linear))
    plugin.set_auto_wrap_policy(model)
    assert plugin.should_wrap(model.embed) and not plugin.should_wrap(model.layers[0])
    with pytest.raises(ValueError, match="auto_wrap_policy must be one of"):
        FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="wrap_whatever")


def test_mixed_precision_policies():
    for text, dtype in (("fp16", torch.float16), ("bf16", torch.bfloat16), ("fp8", torch.bfloat16)):
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2)
        plugin.set_mixed_precision(text)
        policy = plugin.mixed_precision_policy
        assert isinstance(policy, MixedPrecisionPolicy) and (policy.param_dtype, policy.reduce_dtype, policy.output_dtype) == (dtype,) * 3
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2)
    plugin.set_mixed_precision("no")
    assert plugin.mixed_precision_policy is None
    with pytest.raises(ValueError, match="Unknown mixed precision"):
        plugin.set_mixed_precision("int3")
    # a policy that is already there stays unless told to override
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, mixed_precision_policy={"param_dtype": "bf16", "reduce_dtype": torch.float32})
    assert plugin.mixed_precision_policy.param_dtype == torch.bfloat16 and plugin.mixed_precision_policy.reduce_dtype == torch.float32
    plugin.set_mixed_precision("fp16")
    assert plugin.mixed_precision_policy.param_dtype == torch.bfloat16
    plugin.set_mixed_precision("fp16", override=True)
    assert plugin.mixed_precision_policy.param_dtype == torch.float16
    with pytest.raises(ValueError, match="Invalid mixed precision policy keys"):
        FullyShardedDataParallelPlugin(fsdp_version=2, mixed_precision_policy={"params": "bf16"})
    # torch's own policy objects are read field by field
    from torch.distributed.fsdp import MixedPrecisionPolicy as TorchPolicy

    plugin = FullyShardedDataParallelPlugin(fsdp_version=2)
    plugin.set_mixed_precision(TorchPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.float32))
    assert plugin.mixed_precision_policy.reduce_dtype == torch.float32
    plugin.validate_mixed_precision_policy()
    plugin.mixed_precision_policy = "bf16"
    with pytest.raises(ValueError, match="must be a dict, a precision string or a MixedPrecisionPolicy"):
        plugin.validate_mixed_precision_policy()


def test_ram_efficient_loading_is_shared_through_the_environment():
    import os

    with patch_environment(fsdp_version="2", fsdp_cpu_ram_efficient_loading="true"):
        assert FullyShardedDataParallelPlugin().cpu_ram_efficient_loading is True
    with patch_environment(fsdp_version="2"):
        # an explicit argument is exported, so that model-loading code outside the Accelerator sees the same answer
        FullyShardedDataParallelPlugin(cpu_ram_efficient_loading=True)
        assert os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] == "True"
        FullyShardedDataParallelPlugin(cpu_ram_efficient_loading=False)
        assert os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] == "False"
    os.environ.pop("FSDP_CPU_RAM_EFFICIENT_LOADING", None)


def _ignored_by_regex():
    from accelerate_hpc_test_b200 import Accelerator

    with patch_environment(fsdp_version="2", fsdp_auto_wrap_policy="TRANSFORMER_BASED_WRAP", fsdp_transformer_cls_to_wrap="Block",
                           fsdp_ignored_modules=r".*\.q_proj$", accelerate_use_fsdp="true"):
        acc = Accelerator()
        assert acc.state.fsdp_plugin.ignored_modules == r".*\.q_proj$"
        torch.manual_seed(0)
        model = Net()
        ignored = {model.layers[0].q_proj.weight, model.layers[0].q_proj.bias, model.layers[1].q_proj.weight, model.layers[1].q_proj.bias}
        opt = torch.optim.SGD(model.parameters(), lr=0.1)
        engine, opt = acc.prepare(model, opt)
    assert set(engine._ignored_params) == ignored
    sharded_names = {s.name for u in engine._units for s in u.segs}
    assert not any("q_proj" in name for name in sharded_names) and "layers.0.o_proj.weight" in sharded_names
    # the ignored parameters stay whole, identical on every rank, and still receive gradients
    x = torch.randn(4, 8, generator=torch.Generator().manual_seed(acc.process_index))
    acc.backward(engine(x).sum())
    assert model.layers[0].q_proj.weight.shape == (8, 8) and model.layers[0].q_proj.weight.grad is not None


def test_ignored_modules_by_regular_expression_two_processes():
    debug_launcher(_ignored_by_regex, num_processes=2)


def test_cpu_offload_as_bool_or_torch_policy():
    from torch.distributed.fsdp import CPUOffload, CPUOffloadPolicy, OffloadPolicy

    off = FullyShardedDataParallelPlugin(fsdp_version=2, cpu_offload=False)
    assert off.offloads_parameters is False
    off.set_cpu_offload()
    assert type(off.cpu_offload) is OffloadPolicy and off.offloads_parameters is False     # the policy object of "no offload" is not an offload
    off.validate_cpu_offload()
    on = FullyShardedDataParallelPlugin(fsdp_version=2, cpu_offload=True)
    on.set_cpu_offload()
    assert isinstance(on.cpu_offload, CPUOffloadPolicy) and on.offloads_parameters is True
    on.set_cpu_offload()                                                                    # already a policy: left alone
    first = FullyShardedDataParallelPlugin(fsdp_version=1, cpu_offload=True)
    first.set_cpu_offload()
    assert isinstance(first.cpu_offload, CPUOffload) and first.offloads_parameters is True
    first.validate_cpu_offload()
    wrong = FullyShardedDataParallelPlugin(fsdp_version=2, cpu_offload=True)
    with pytest.raises(ValueError, match="must be an instance of `torch.distributed.fsdp.OffloadPolicy` in FSDP2"):
        wrong.validate_cpu_offload()
    wrong = FullyShardedDataParallelPlugin(fsdp_version=1, cpu_offload=False)
    with pytest.raises(ValueError, match="must be an instance of `torch.distributed.fsdp.CPUOffload` in FSDP1"):
        wrong.validate_cpu_offload()
