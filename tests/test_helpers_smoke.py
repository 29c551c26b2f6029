"""Public helpers that no other test names: each is called once with inputs small enough to check the answer by eye."""

import os

import pytest
import torch
import torch.nn as nn


def test_other_helpers():
    from accelerate_hpc_test_b200.utils import other

    assert other.merge_dicts({"a": {"x": 1}, "b": 2}, {"a": {"y": 3}, "b": 0, "c": 4}) == {"a": {"y": 3, "x": 1}, "b": 2, "c": 4}
    assert other.convert_bytes(512) == "512 bytes" and other.convert_bytes(1536) == "1.5 KB" and other.convert_bytes(3 * 1024**3) == "3.0 GB"
    model = nn.Sequential(nn.Linear(2, 3), nn.ModuleList([nn.Linear(3, 3), nn.Linear(3, 3)]))
    assert other.recursive_getattr(model, "1.0.weight") is model[1][0].weight
    assert other.get_pretty_name(other.merge_dicts) == "merge_dicts" and other.get_pretty_name(model) == "Sequential"
    order = [name for name, _ in other.get_module_children_bottom_up(model, return_fqns=True)]
    assert order == ["0", "1.0", "1.1", "1", ""] and other.get_module_children_bottom_up(model)[-1] is model
    assert other.is_repeated_blocks(model[1]) and not other.is_repeated_blocks(model) and other.has_repeated_blocks(model)
    assert not other.is_compiled_module(model) and not other.has_compiled_regions(model)
    compiled = torch.compile(nn.Linear(2, 2), backend="eager")
    assert other.is_compiled_module(compiled)
    regional = other.compile_regions(model, backend="eager")
    assert other.has_compiled_regions(regional) and regional._orig_mod is model
    assert other.extract_model_from_parallel(regional, keep_torch_compile=False) is model
    other.compile_regions_deepspeed(nn.Sequential(nn.ModuleList([nn.Linear(2, 2), nn.Linear(2, 2)])), backend="eager")
    # aliased tensors: one name per storage survives
    shared = torch.ones(4)
    cleaned = other.clean_state_dict_for_safetensors({"a": shared, "b": shared, "c": torch.zeros(2, 2).t()})
    assert set(cleaned) == {"a", "c"} and cleaned["c"].is_contiguous()
    os.environ["ACCELERATE_MIXED_PRECISION"] = "bf16"
    try:
        with other.clear_environment_for_defaults():
            assert "ACCELERATE_MIXED_PRECISION" not in os.environ
        assert os.environ["ACCELERATE_MIXED_PRECISION"] == "bf16"
    finally:
        del os.environ["ACCELERATE_MIXED_PRECISION"]
    other.check_os_kernel()
    port = other.get_free_port()
    assert isinstance(port, int) and not other.is_port_in_use(port)
    assert not other.model_has_dtensor(model)


def test_import_probes_answer_with_booleans():
    from accelerate_hpc_test_b200.utils import imports

    probes = [getattr(imports, name) for name in dir(imports) if name.startswith("is_") and name.endswith("_available")]
    assert len(probes) > 50
    for probe in probes:
        assert isinstance(probe(), bool), probe.__name__
    for gone in ("is_torch_xla_available", "is_xpu_available", "is_npu_available", "is_mlu_available", "is_musa_available",
                 "is_sdaa_available", "is_hpu_available", "is_mps_available", "is_sagemaker_available"):
        assert getattr(imports, gone)() is False, gone                  # back ends this build does not have
    assert imports.is_torch_distributed_available() and imports.is_safetensors_available() and imports.is_pytest_available()
    assert imports.is_notebook() is False and imports.is_habana_gaudi1() is False

    @imports.torchao_required
    def needs_torchao():
        return "ran"

    if imports.is_torchao_available():
        assert needs_torchao() == "ran"
    else:
        with pytest.raises(ImportError):
            needs_torchao()


def test_environment_helpers():
    from accelerate_hpc_test_b200.utils import environment

    info = environment.get_cpu_distributed_information()
    assert isinstance(info, environment.CPUInformation) and info.world_size >= 1 and info.rank >= 0
    assert environment.get_current_device_type()[0] in ("cuda", "cpu")
    assert isinstance(environment.check_cuda_p2p_ib_support(), bool) and isinstance(environment.check_cuda_fp8_capability(), bool)
    try:
        names, count = environment.get_gpu_info()                        # asks nvidia-smi
        assert len(names) == count and isinstance(environment.get_driver_version(), str)
    except FileNotFoundError:
        pass                                                             # no driver tools on this machine: the p2p check above still answered
    environment.set_numa_affinity(0)                                   # no NUMA tools in the sandbox: must not raise


def test_fsdp_utility_functions(monkeypatch):
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin, fsdp_utils

    monkeypatch.delenv("FSDP_CPU_RAM_EFFICIENT_LOADING", raising=False)
    fsdp_utils.enable_fsdp_ram_efficient_loading()
    assert os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] == "True"
    fsdp_utils.disable_fsdp_ram_efficient_loading()
    assert os.environ["FSDP_CPU_RAM_EFFICIENT_LOADING"] == "False"
    names = fsdp_utils.fsdp2_canonicalize_names({"_orig_mod.layers.0._checkpoint_wrapped_module.weight": 1, "layers.1._orig_mod.bias": 2})
    assert names == {"layers.0.weight": 1, "layers.1.bias": 2}
    old, new = nn.Parameter(torch.zeros(2)), nn.Parameter(torch.ones(2))
    placeholder = torch.empty(1)
    placeholder.data_ptr = old.data_ptr()                               # how `prepare` marks a parameter it will swap
    optimizer = torch.optim.SGD([placeholder], lr=0.1)
    fsdp_utils.fsdp2_switch_optimizer_parameters(optimizer, {old.data_ptr(): new})
    assert optimizer.param_groups[0]["params"][0] is new
    with pytest.raises(KeyError, match="couldn't be switched to its sharded version"):
        fsdp_utils.fsdp2_switch_optimizer_parameters(torch.optim.SGD([placeholder], lr=0.1), {})
    model = nn.Sequential(nn.Linear(4, 4), nn.ReLU())
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap", transformer_cls_names_to_wrap=["Linear"])
    policy = fsdp_utils.fsdp2_prepare_auto_wrap_policy(plugin, model)
    assert policy(model[0]) and not policy(model[1])
    assert isinstance(fsdp_utils.get_fsdp2_grad_scaler(), torch.amp.GradScaler)

    class Tied(nn.Module):
        _tied_weights_keys = ["head.weight"]

        def __init__(self):
            super().__init__()
            self.embed, self.head = nn.Linear(3, 3, bias=False), nn.Linear(3, 3, bias=False)
            self.head.weight = self.embed.weight

    tied = Tied()

    def fresh(module):                                                   # a param_init_fn that breaks the tie, as materialising does
        for name, p in list(module.named_parameters(recurse=False)):
            setattr(module, name, nn.Parameter(torch.zeros_like(p)))
        return module

    init = fsdp_utils.ensure_weights_retied(fresh, tied, torch.device("cpu"))
    init(tied.embed)
    init(tied.head)
    assert tied.head.weight is tied.embed.weight
    assert fsdp_utils.ensure_weights_retied(fresh, nn.Linear(2, 2), torch.device("cpu")) is fresh


def test_operation_decorators_and_modeling_helpers():
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.utils import modeling, operations

    Accelerator(cpu=True)

    @operations.chained_operation
    def outer():
        raise operations.DistributedOperationException("inner failure")

    with pytest.raises(operations.DistributedOperationException, match="Error found while calling") as raised:
        outer()
    assert "inner failure" in str(raised.value.__cause__)

    @operations.verify_operation
    def double(tensor):
        return tensor * 2

    assert torch.equal(double(torch.ones(2)), torch.full((2,), 2.0))         # one process: passes straight through
    assert operations.is_torch_tensor(torch.ones(1)) and not operations.is_torch_tensor([1])
    with modeling.get_mixed_precision_context_manager(native_amp=False):
        pass
    assert modeling.get_grad_scaler() is not None
    assert modeling.check_tied_parameters_in_config(nn.Linear(2, 2)) is False
    assert modeling.is_peft_model(nn.Linear(2, 2)) is False


def test_deepspeed_wrappers_and_megatron_mapping():
    from accelerate_hpc_test_b200.utils import MegatronLMPlugin
    from accelerate_hpc_test_b200.utils.deepspeed import DeepSpeedEngineWrapper, DeepSpeedOptimizerWrapper, DeepSpeedSchedulerWrapper
    from accelerate_hpc_test_b200.utils.megatron_lm import megatron_plugin_to_parallelism_config

    layer = nn.Linear(2, 2)
    inner = torch.optim.SGD(layer.parameters(), lr=0.5)
    wrapped = DeepSpeedOptimizerWrapper(inner)
    before = layer.weight.detach().clone()
    layer(torch.ones(1, 2)).sum().backward()
    wrapped.step()                                                           # the optimizer is real here: the update happens
    wrapped.zero_grad()
    assert not torch.equal(layer.weight, before) and wrapped.step_was_skipped is False
    inner.overflow = True
    assert wrapped.step_was_skipped is True
    schedule = torch.optim.lr_scheduler.ConstantLR(inner, factor=0.5, total_iters=1)
    DeepSpeedSchedulerWrapper(schedule, wrapped).step()
    assert inner.param_groups[0]["lr"] == 0.5

    class Engine:                                                            # DeepSpeed's contract: the engine differentiates and steps
        def __init__(self):
            self.calls = []

        def set_gradient_accumulation_boundary(self, is_boundary):
            self.calls.append(("boundary", is_boundary))

        def backward(self, loss, **kwargs):
            self.calls.append("backward")

        def step(self):
            self.calls.append("step")

        def get_global_grad_norm(self):
            return torch.tensor(1.5)

    engine = Engine()
    adapter = DeepSpeedEngineWrapper(engine)
    adapter.backward(torch.tensor(1.0), sync_gradients=False)
    adapter.backward(torch.tensor(1.0))
    assert engine.calls == [("boundary", False), "backward", ("boundary", True), "backward", "step"] and adapter.get_global_grad_norm() == 1.5
    plain = DeepSpeedEngineWrapper(layer)                                    # no engine contract: the loss is differentiated natively
    layer.zero_grad()
    plain.backward(layer(torch.ones(1, 2)).sum())
    assert layer.weight.grad is not None and plain.get_global_grad_norm() is None
    pc = megatron_plugin_to_parallelism_config(MegatronLMPlugin(tp_degree=2, context_parallel_size=1))
    assert pc.tp_size == 2 and pc.cp_size == 1
    with pytest.raises(NotImplementedError, match="pp_degree must be 1"):
        megatron_plugin_to_parallelism_config(MegatronLMPlugin(pp_degree=2))


def test_estimate_helpers():
    from accelerate_hpc_test_b200.commands import estimate

    gb = 2**30
    assert estimate.estimate_training_usage(4 * gb, "float32") == {"model": 4 * gb, "optimizer": 8 * gb, "gradients": 4 * gb, "step": 16 * gb}
    mixed = estimate.estimate_training_usage(4 * gb, "bfloat16")
    assert mixed == {"model": 4 * gb, "optimizer": 8 * gb, "gradients": 6 * gb, "step": 8 * gb}
    assert set(estimate.estimate_training_usage(gb, "int8").values()) == {-1}
    table = estimate.create_ascii_table(["dtype", "Total Size"], [["float32", "1.0 GB"], ["int4", "128.0 MB"]], "Memory Usage for loading `m`")
    lines = table.splitlines()
    assert lines[0].startswith("┌") and lines[-1].startswith("└") and len({len(line) for line in lines}) == 1       # a rectangle
    assert "Memory Usage for loading `m`" in lines[1] and "float32" in table and "128.0 MB" in table
    empty = estimate.create_empty_model("llama3-1b")
    assert next(empty.parameters()).device.type == "meta"
    with pytest.raises(OSError, match="neither a built-in config"):
        estimate.create_empty_model("somebody/some-model")


def test_offloaded_weights_come_back(tmp_path):
    from accelerate_hpc_test_b200.big_modeling import cpu_offload
    from accelerate_hpc_test_b200.utils import modeling
    from accelerate_hpc_test_b200.utils.offload import offload_weight, save_offload_index

    torch.manual_seed(0)
    source = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 2))
    index = {}
    for name, tensor in source.state_dict().items():
        offload_weight(tensor, name, str(tmp_path), index=index)
    save_offload_index(index, str(tmp_path))
    target = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 2))
    modeling.load_offloaded_weights(target, index, str(tmp_path))
    assert all(torch.equal(a, b) for a, b in zip(source.state_dict().values(), target.state_dict().values()))
    # a model whose weights sit behind offload hooks: one sub-module's entries are filled in
    hooked = cpu_offload(nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 2)), execution_device="cpu")
    wanted = {"1.weight": None, "1.bias": None, "0.weight": None}
    filled = modeling.get_state_dict_from_offload(hooked[1], "1.weight", wanted)
    assert filled["1.weight"].shape == (2, 4) and filled["1.bias"].shape == (2,) and filled["0.weight"] is None
    whole = modeling.get_state_dict_offloaded_model(hooked)
    assert set(whole) == {"0.weight", "0.bias", "1.weight", "1.bias"} and all(t.device.type == "cpu" for t in whole.values())


def test_rng_synchronisation_single_process_and_state_descriptor():
    import threading

    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.state import ThreadLocalSharedDict
    from accelerate_hpc_test_b200.utils import RNGType, synchronize_rng_state, synchronize_rng_states

    Accelerator(cpu=True)
    torch.manual_seed(5)
    expected = torch.rand(2)
    torch.manual_seed(5)
    synchronize_rng_state(RNGType.TORCH)                                # one process: the state is "broadcast" to itself
    synchronize_rng_states(["torch"])
    assert torch.equal(torch.rand(2), expected)
    generator = torch.Generator().manual_seed(9)
    synchronize_rng_state(RNGType.GENERATOR, generator=generator)
    assert torch.equal(torch.rand(2, generator=generator), torch.rand(2, generator=torch.Generator().manual_seed(9)))

    class Holder:
        shared = ThreadLocalSharedDict()

    Holder.shared["main"] = 1
    seen = {}

    def other_thread():
        seen["keys"] = list(Holder().shared)                            # a fresh thread starts with its own, empty dict
        Holder().shared["worker"] = 2

    worker = threading.Thread(target=other_thread)
    worker.start()
    worker.join()
    assert seen["keys"] == [] and list(Holder.shared) == ["main"]


def test_generate_device_map_spreads_blocks_evenly():
    from accelerate_hpc_test_b200.inference import generate_device_map

    model = nn.Sequential(*[nn.Linear(16, 16) for _ in range(4)])
    one = generate_device_map(model, num_processes=1)
    assert set(one.values()) <= {0, "cpu"} and len(set(one.values())) == 1
    two = generate_device_map(model, num_processes=2)
    placed = [two[name] for name in ("0", "1", "2", "3")]
    assert placed == sorted(placed) and placed[0] != placed[-1]         # contiguous halves on two devices
