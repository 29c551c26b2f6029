"""ParallelismConfig: axis bookkeeping, env defaults, validation, mesh construction on a 4-process gloo job (behaviour of reference
tests/test_parallelism_config.py)."""

import types
import warnings

import pytest

from accelerate_hpc_test_b200 import ParallelismConfig
from accelerate_hpc_test_b200.utils.dataclasses import DeepSpeedSequenceParallelConfig, TorchContextParallelConfig, TorchTensorParallelConfig
from accelerate_hpc_test_b200.utils.environment import patch_environment


def test_defaults_are_all_one_and_no_mesh():
    pc = ParallelismConfig()
    assert pc.total_size == 1 and pc.active_mesh_dims == [] and pc._get_mesh() == ()
    assert pc.build_device_mesh("cpu") is None
    assert pc.cp_backend == "torch" and pc.sp_backend == "deepspeed"
    assert pc.tp_handler is None and pc.cp_handler is None and pc.sp_handler is None


@pytest.mark.parametrize("sizes,dp,non_dp,shard_cp,dp_cp,fsdp,mesh", [
    (dict(dp_shard_size=4, tp_size=2), ["dp_shard"], ["tp"], ["dp_shard"], ["dp_shard"], ["dp_shard_cp"], (("dp_shard", "tp"), (4, 2))),
    (dict(dp_replicate_size=2, dp_shard_size=2, cp_size=2), ["dp_replicate", "dp_shard"], ["cp"], ["dp_shard", "cp"],
     ["dp_replicate", "dp_shard", "cp"], ["dp_replicate", "dp_shard_cp"], (("dp_replicate", "dp_shard", "cp"), (2, 2, 2))),
    (dict(sp_size=8), [], ["sp"], ["sp"], [], ["dp_shard_cp"], (("sp",), (8,))),
    (dict(dp_shard_size=2, sp_size=2, tp_size=2), ["dp_shard"], ["tp", "sp"], ["dp_shard", "sp"], ["dp_shard"], ["dp_shard_cp"],
     (("dp_shard", "sp", "tp"), (2, 2, 2))),
])
def test_axis_groups_and_mesh_order(sizes, dp, non_dp, shard_cp, dp_cp, fsdp, mesh):
    pc = ParallelismConfig(**sizes)
    assert pc.dp_dim_names == dp and pc.non_dp_dim_names == non_dp
    assert pc.dp_shard_cp_dim_names == shard_cp and pc.dp_cp_dim_names == dp_cp and pc.fsdp_dim_names == fsdp
    assert pc._get_mesh() == mesh and pc.active_mesh_dims == dp + non_dp
    assert pc.total_size == pc.data_parallel_size * pc.non_data_parallel_size


def test_handlers_are_created_for_enabled_axes_only():
    pc = ParallelismConfig(dp_shard_size=2, tp_size=2, cp_size=2)
    assert isinstance(pc.tp_handler, TorchTensorParallelConfig) and isinstance(pc.cp_handler, TorchContextParallelConfig)
    assert pc.sp_handler is None
    assert isinstance(ParallelismConfig(sp_size=2).sp_handler, DeepSpeedSequenceParallelConfig)


def test_environment_defaults_and_explicit_arguments_win():
    with patch_environment(parallelism_config_dp_shard_size="4", parallelism_config_tp_size="2", parallelism_config_cp_backend="b200",
                           parallelism_config_sp_backend="b200"):
        pc = ParallelismConfig()
        assert (pc.dp_shard_size, pc.tp_size, pc.cp_backend, pc.sp_backend) == (4, 2, "b200", "b200")
        assert ParallelismConfig(tp_size=1).tp_size == 1


@pytest.mark.parametrize("kwargs,message", [
    (dict(tp_size=0), "tp_size must be at least 1"),
    (dict(cp_size=2, sp_size=2), "mutually exclusive"),
    (dict(dp_replicate_size=2, tp_size=2), "cannot be used with pure data parallelism"),
    (dict(dp_replicate_size=2, cp_size=2), "cannot be used with pure data parallelism"),
    (dict(cp_size=2, cp_backend="ring"), "Invalid cp_backend"),
    (dict(sp_size=2, sp_backend="megatron"), "Invalid sp_backend"),
])
def test_invalid_configurations(kwargs, message):
    with pytest.raises(ValueError, match=message):
        ParallelismConfig(**kwargs)


def test_repr_and_json():
    pc = ParallelismConfig(dp_shard_size=2, tp_size=2)
    text = repr(pc)
    assert "dp_shard_size=2" in text and "tp_size=2" in text and "total_size=4" in text
    blob = pc.to_json()
    assert blob["tp_size"] == 2 and "device_mesh" not in blob and isinstance(blob["tp_handler"], dict)


def _fake_accelerator(num_processes, multi_device=True, is_fsdp2=False):
    return types.SimpleNamespace(num_processes=num_processes, multi_device=multi_device, is_fsdp2=is_fsdp2, is_main_process=True,
                                 distributed_type="MULTI_GPU" if multi_device else "NO")


def test_validation_against_the_job():
    pc = ParallelismConfig()
    pc._validate_accelerator(_fake_accelerator(1, multi_device=False))
    assert pc.total_size == 1
    pc._validate_accelerator(_fake_accelerator(4))
    assert pc.dp_replicate_size == 4  # nothing asked on a 4-process job: plain data parallelism
    with pytest.raises(ValueError, match="does not match num_processes"):
        ParallelismConfig(dp_shard_size=2, tp_size=2)._validate_accelerator(_fake_accelerator(8))
    with pytest.raises(ValueError, match="only compatible"):
        ParallelismConfig(dp_shard_size=2)._validate_accelerator(_fake_accelerator(2, multi_device=False))
    idle = ParallelismConfig(dp_shard_size=2, tp_handler=TorchTensorParallelConfig())
    with pytest.warns(UserWarning, match="tp_handler is set, but tp_size is set to 1"):
        idle._validate_accelerator(_fake_accelerator(2))
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        ParallelismConfig(dp_shard_size=2)._validate_accelerator(_fake_accelerator(2))


def test_sizes_are_frozen_once_the_mesh_exists():
    pc = ParallelismConfig(dp_shard_size=2)
    pc.device_mesh = object()
    pc._set_size("tp", 1)
    with pytest.raises(AssertionError, match="after the device mesh was built"):
        pc._set_size("tp", 2)
    with pytest.raises(AssertionError, match="Parallelism must be one of"):
        pc._set_size("pp", 2)
    with pytest.raises(ValueError, match="need to pass a device_type"):
        ParallelismConfig(dp_shard_size=2).get_device_mesh()


def _mesh_job():
    import torch.distributed as dist

    from accelerate_hpc_test_b200.state import PartialState

    state = PartialState()
    pc = ParallelismConfig(dp_replicate_size=2, dp_shard_size=2)
    mesh = pc.get_device_mesh("cpu")
    assert mesh.mesh_dim_names == ("dp_replicate", "dp_shard") and tuple(mesh.shape) == (2, 2)
    assert mesh["dp"].size() == 4 and mesh["dp_shard_cp"].size() == 2 and mesh["dp_cp"].size() == 4
    assert mesh.get_local_rank("dp_replicate") == state.process_index // 2 and mesh.get_local_rank("dp_shard") == state.process_index % 2
    assert pc.get_device_mesh("cpu") is mesh
    try:
        pc.get_device_mesh("cuda")
    except ValueError as err:
        assert "already created with device type cpu" in str(err)
    else:
        raise AssertionError("a second device type must be refused")
    dist.barrier()


def test_mesh_on_four_cpu_processes():
    from accelerate_hpc_test_b200.launchers import debug_launcher

    debug_launcher(_mesh_job, num_processes=4)


def _mesh_ranks():
    import pytest

    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig
    from accelerate_hpc_test_b200.state import AcceleratorState

    acc = Accelerator(parallelism_config=ParallelismConfig(cp_size=2))
    assert acc.context_parallel_rank == acc.process_index and acc.parallelism_config.cp_enabled
    with pytest.raises(RuntimeError, match="Tensor parallelism is not enabled"):
        acc.tensor_parallel_rank
    with pytest.raises(RuntimeError, match="Sequence parallelism is not enabled"):
        acc.sequence_parallel_rank
    with pytest.raises(NotImplementedError, match="Pipeline parallelism"):
        acc.pipeline_parallel_rank
    assert acc.should_save_model == (acc.process_index == 0)        # the second cp rank holds a replica of what the first saves
    AcceleratorState._reset_state(reset_partial_state=False)
    acc = Accelerator(parallelism_config=ParallelismConfig(sp_size=2))
    assert acc.sequence_parallel_rank == acc.process_index
    AcceleratorState._reset_state(reset_partial_state=False)
    acc = Accelerator(parallelism_config=ParallelismConfig(tp_size=2))
    assert acc.tensor_parallel_rank == acc.process_index and acc.should_save_model


def test_rank_properties_follow_the_mesh_two_processes():
    from accelerate_hpc_test_b200 import debug_launcher

    debug_launcher(_mesh_ranks, num_processes=2)
