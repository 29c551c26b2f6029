"""Sharded-engine generality: foreign optimizers on shard views, DCP checkpoints that interchange with torch FSDP2
(what the reference writes, utils/fsdp_utils.py:147-158), and the torch FSDP2 fallback of ``Accelerator`` (2 gloo processes)."""

import os

import torch
import torch.nn as nn

from accelerate_hpc_test_b200.launchers import debug_launcher


def _mlp(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(16, 32), nn.Tanh(), nn.Linear(32, 32), nn.Tanh(), nn.Linear(32, 4))


def _batch(step, rank=None):
    g = torch.Generator().manual_seed(100 + step)
    x, y = torch.randn(8, 16, generator=g), torch.randn(8, 4, generator=g)
    if rank is None:
        return x, y
    return x[rank * 4 : rank * 4 + 4], y[rank * 4 : rank * 4 + 4]


def _plugin():
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    return FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="no_wrap")


def _sgd_two_groups_parity():
    """SGD with momentum and two parameter groups (different lr / weight decay): the engine steps the user's own optimizer on
    shard-shaped views; the result must equal one process training on the whole batch."""
    from accelerate_hpc_test_b200 import Accelerator

    acc = Accelerator(fsdp_plugin=_plugin())
    model, ref = _mlp(), _mlp()

    def groups(m):
        decay = [p for p in m.parameters() if p.dim() > 1]
        rest = [p for p in m.parameters() if p.dim() <= 1]
        return [{"params": decay, "lr": 0.05, "weight_decay": 0.1}, {"params": rest, "lr": 0.01, "weight_decay": 0.0}]

    opt = torch.optim.SGD(groups(model), momentum=0.9)
    ropt = torch.optim.SGD(groups(ref), momentum=0.9)
    model, opt = acc.prepare(model, opt)
    assert type(model).__name__ == "B200ShardedModule" and model._generic
    for step in range(4):
        x, y = _batch(step, acc.process_index)
        opt.zero_grad()
        loss = nn.functional.mse_loss(model(x), y)
        acc.backward(loss)
        norm = acc.clip_grad_norm_(model.parameters(), 0.5)
        opt.step()
        gx, gy = _batch(step)
        ropt.zero_grad()
        nn.functional.mse_loss(ref(gx), gy).backward()
        rnorm = torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.5)
        ropt.step()
        assert abs(float(norm) - float(rnorm)) < 1e-5 * max(1.0, float(rnorm)), (step, float(norm), float(rnorm))
    sd = model.full_state_dict()
    for name, p in ref.named_parameters():
        assert torch.allclose(sd[name], p.detach(), atol=1e-6), name
    # optimizer.state_dict() is the gathered, torch-layout state (momentum buffers of full shape)
    osd = opt.state_dict()
    shapes = {tuple(v["momentum_buffer"].shape) for v in osd["state"].values()}
    assert (32, 16) in shapes and (4,) in shapes and len(osd["param_groups"]) == 2


def test_generic_optimizer_sgd_two_groups_matches_single_process():
    debug_launcher(_sgd_two_groups_parity, num_processes=2)


def _native_checkpoint_read_by_torch_fsdp2(tmpdir):
    """save_state with the native engine -> a DCP directory that torch FSDP2 (the reference's engine) loads."""
    import torch.distributed.checkpoint as dist_cp
    from torch.distributed.checkpoint.state_dict import get_model_state_dict, set_model_state_dict
    from torch.distributed.fsdp import fully_shard

    from accelerate_hpc_test_b200 import Accelerator

    acc = Accelerator(fsdp_plugin=_plugin())
    model = _mlp()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2)
    model, opt = acc.prepare(model, opt)
    for step in range(2):
        x, y = _batch(step, acc.process_index)
        opt.zero_grad()
        acc.backward(nn.functional.mse_loss(model(x), y))
        opt.step()
    acc.save_state(tmpdir)
    want = model.full_state_dict()
    d = os.path.join(tmpdir, "pytorch_model_fsdp_0")
    assert os.path.isfile(os.path.join(d, ".metadata")) and any(f.endswith(".distcp") for f in os.listdir(d))
    assert os.path.isdir(os.path.join(tmpdir, "optimizer_0"))
    # the reference side: torch fully_shard + dist_cp.load({"model": sd})
    other = _mlp(seed=123)
    fully_shard(other)
    holder = {"model": get_model_state_dict(other)}
    dist_cp.load(holder, storage_reader=dist_cp.FileSystemReader(d))
    set_model_state_dict(other, holder["model"])
    for name, p in other.named_parameters():
        assert torch.allclose(p.full_tensor(), want[name], atol=0), name
    # and the single-process merge used by `accelerate merge-weights`
    from accelerate_hpc_test_b200.utils.fsdp_utils import merge_fsdp_weights

    merge_fsdp_weights(tmpdir, os.path.join(tmpdir, "merged"))  # every rank calls it: the main process works, the others wait
    from safetensors.torch import load_file

    merged = load_file(os.path.join(tmpdir, "merged", "model.safetensors"))
    assert set(merged) == set(want) and all(torch.equal(merged[k], want[k]) for k in want)


def test_native_sharded_checkpoint_loads_in_torch_fsdp2(tmp_path):
    debug_launcher(_native_checkpoint_read_by_torch_fsdp2, args=(str(tmp_path),), num_processes=2)


def _torch_checkpoint_read_by_native(tmpdir):
    """A DCP directory written the reference's way (torch FSDP2 + dist_cp.save) -> load_state of the native engine."""
    import torch.distributed as dist
    import torch.distributed.checkpoint as dist_cp
    from torch.distributed.checkpoint.state_dict import get_model_state_dict, get_optimizer_state_dict
    from torch.distributed.fsdp import fully_shard

    from accelerate_hpc_test_b200 import Accelerator

    acc = Accelerator(fsdp_plugin=_plugin())
    theirs = _mlp(seed=7)
    fully_shard(theirs)
    topt = torch.optim.AdamW(theirs.parameters(), lr=1e-2)
    for step in range(2):
        x, y = _batch(step, acc.process_index)
        topt.zero_grad()
        nn.functional.mse_loss(theirs(x), y).backward()
        topt.step()
    os.makedirs(os.path.join(tmpdir, "pytorch_model_fsdp_0"), exist_ok=True)
    os.makedirs(os.path.join(tmpdir, "optimizer_0"), exist_ok=True)
    dist_cp.save({"model": get_model_state_dict(theirs)}, storage_writer=dist_cp.FileSystemWriter(os.path.join(tmpdir, "pytorch_model_fsdp_0")))
    dist_cp.save({"optimizer": get_optimizer_state_dict(theirs, topt)}, storage_writer=dist_cp.FileSystemWriter(os.path.join(tmpdir, "optimizer_0")))
    dist.barrier()
    mine = _mlp(seed=99)
    opt = torch.optim.AdamW(mine.parameters(), lr=1e-2)
    mine, opt = acc.prepare(mine, opt)
    mine.load_sharded_state(tmpdir, 0)
    mine.load_sharded_optimizer(opt, tmpdir, 0)
    got = mine.full_state_dict()
    for name, p in theirs.named_parameters():
        assert torch.allclose(got[name], p.full_tensor(), atol=0), name
    # the moments arrived too: one more identical step on both sides keeps them equal
    x, y = _batch(5, acc.process_index)
    topt.zero_grad()
    nn.functional.mse_loss(theirs(x), y).backward()
    topt.step()
    opt.zero_grad()
    acc.backward(nn.functional.mse_loss(mine(x), y))
    opt.step()
    got = mine.full_state_dict()
    for name, p in theirs.named_parameters():
        assert torch.allclose(got[name], p.full_tensor(), atol=2e-6), (name, (got[name] - p.full_tensor()).abs().max())


def test_torch_fsdp2_checkpoint_loads_in_native_engine(tmp_path):
    debug_launcher(_torch_checkpoint_read_by_native, args=(str(tmp_path),), num_processes=2)


def _fallback_engine(tmpdir):
    """ACCELERATE_B200_FSDP=0: Accelerator shards with torch fully_shard, re-points the optimizer, trains, saves and resumes."""
    from accelerate_hpc_test_b200 import Accelerator

    os.environ["ACCELERATE_B200_FSDP"] = "0"
    try:
        acc = Accelerator(fsdp_plugin=_plugin())
        model = _mlp()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-2)
        model, opt = acc.prepare(model, opt)
        assert getattr(model, "_b200_torch_fsdp2", False)
        losses = []
        for step in range(6):
            x, y = _batch(0, acc.process_index)
            opt.zero_grad()
            loss = nn.functional.mse_loss(model(x), y)
            acc.backward(loss)
            opt.step()
            losses.append(loss.item())
            if step == 2:
                acc.save_state(tmpdir)
        assert losses[-1] < losses[0]
        after = [p.full_tensor().clone() for p in model.parameters()]
        acc.load_state(tmpdir)
        for step in range(3, 6):
            x, y = _batch(0, acc.process_index)
            opt.zero_grad()
            acc.backward(nn.functional.mse_loss(model(x), y))
            opt.step()
        for a, p in zip(after, model.parameters()):
            assert torch.allclose(a, p.full_tensor(), atol=1e-7)
    finally:
        os.environ.pop("ACCELERATE_B200_FSDP", None)


def test_torch_fsdp2_fallback_trains_and_resumes(tmp_path):
    debug_launcher(_fallback_engine, args=(str(tmp_path),), num_processes=2)
