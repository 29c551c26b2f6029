"""``load_checkpoint_and_dispatch(..., broadcast_from_rank0=True)``: only rank 0 reads the checkpoint, every rank ends with the weights —
for a plain module (also the module inside a DDP wrapper), for ``DTensor`` parameters of torch ``fully_shard`` and for the native sharded
engine with a checkpoint split over several files.  Behaviour of reference tests/test_load_checkpoint_and_dispatch_with_broadcast.py,
on two gloo processes."""

import os

import torch
import torch.nn as nn

from accelerate_hpc_test_b200 import debug_launcher


def _net(seed):
    torch.manual_seed(seed)
    net = nn.Sequential(nn.Linear(16, 32), nn.Tanh(), nn.Linear(32, 32), nn.BatchNorm1d(32), nn.Linear(32, 8))
    net[3].running_mean.uniform_(-1, 1)
    return net


def _only_rank0_may_read(monkeypatch_target):
    """Make reading a checkpoint file an error on every rank but 0."""
    import torch.distributed as dist

    import accelerate_hpc_test_b200.utils.modeling as modeling

    original = modeling.load_state_dict

    def guarded(*args, **kwargs):
        assert dist.get_rank() == 0, "a rank other than 0 read the checkpoint although broadcast_from_rank0 was set"
        return original(*args, **kwargs)

    modeling.load_state_dict = guarded
    return lambda: setattr(modeling, "load_state_dict", original)


def _worker(folder):
    import torch.distributed as dist
    from safetensors.torch import save_file
    from torch.distributed.fsdp import fully_shard
    from torch.nn.parallel import DistributedDataParallel

    from accelerate_hpc_test_b200 import Accelerator, PartialState, load_checkpoint_and_dispatch
    from accelerate_hpc_test_b200.state import AcceleratorState
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    state = PartialState()
    source = _net(0)
    truth = {k: v.clone() for k, v in source.state_dict().items()}
    path = os.path.join(folder, "model.safetensors")
    if state.is_main_process:
        save_file({k: v.contiguous() for k, v in truth.items()}, path)
    state.wait_for_everyone()

    def same(model_state):
        for name, value in truth.items():
            got = model_state[name]
            got = got.full_tensor() if hasattr(got, "full_tensor") else got
            assert torch.equal(got, value), name

    restore = _only_rank0_may_read(None)
    try:
        # 1. plain module, and the module inside a DDP wrapper
        plain = _net(1 + state.process_index)
        load_checkpoint_and_dispatch(plain, path, strict=True, broadcast_from_rank0=True)
        same(plain.state_dict())
        wrapped = DistributedDataParallel(_net(5))
        load_checkpoint_and_dispatch(wrapped.module, path, strict=True, broadcast_from_rank0=True)
        same(wrapped.module.state_dict())
        # 2. DTensor parameters (torch fully_shard): each rank keeps its shard of what rank 0 read
        sharded = _net(7)
        for index in (0, 2, 4):
            fully_shard(sharded[index])
        fully_shard(sharded)
        load_checkpoint_and_dispatch(sharded, path, strict=True, broadcast_from_rank0=True)
        assert all(hasattr(p, "full_tensor") for p in sharded.parameters())
        same(sharded.state_dict())
    finally:
        restore()
    # without the flag every rank reads for itself and gets the same result
    again = _net(9)
    load_checkpoint_and_dispatch(again, path, strict=True, broadcast_from_rank0=False)
    same(again.state_dict())

    # 3. the native sharded engine and a checkpoint in several files
    AcceleratorState._reset_state(reset_partial_state=False)
    acc = Accelerator(fsdp_plugin=FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="no_wrap"))
    split = os.path.join(folder, "split")
    acc.save_model(source, split, max_shard_size="3KB", safe_serialization=True)
    acc.wait_for_everyone()
    assert len([f for f in os.listdir(split) if f.endswith(".safetensors")]) > 1
    engine = acc.prepare(_net(11))
    assert type(engine).__name__ == "B200ShardedModule"
    restore = _only_rank0_may_read(None)
    try:
        load_checkpoint_and_dispatch(engine, split, broadcast_from_rank0=True)
    finally:
        restore()
    full = engine.full_state_dict()
    for name, value in truth.items():
        if name in full:
            assert torch.equal(full[name].cpu().to(value.dtype), value), name
    assert torch.equal(dict(engine.module.named_buffers())["3.running_mean"], truth["3.running_mean"])
    dist.barrier()


def test_only_rank0_reads_and_every_rank_receives(tmp_path):
    debug_launcher(_worker, args=(str(tmp_path),), num_processes=2)
