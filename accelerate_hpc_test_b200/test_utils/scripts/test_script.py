"""The end-to-end sanity program behind ``accelerate test``: run it under any launcher and configuration and it walks through what a
training script relies on — which rank executes what, RNG synchronisation, dataloader sharding (per-rank and rank-0 dispatch) against
the unsharded data, ``split_between_processes``, training that matches a single-process baseline, the cross-rank breakpoint trigger.
Behaviour of reference ``src/accelerate/test_utils/scripts/test_script.py:827-905``.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.test_script
"""

import contextlib
import io
import math

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.data_loader import prepare_data_loader
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState
from accelerate_hpc_test_b200.test_utils.training import RegressionDataset, RegressionModel
from accelerate_hpc_test_b200.utils import DataLoaderConfiguration, DistributedType, gather, set_seed, synchronize_rng_states


def announce(title):
    if PartialState().local_process_index == 0:
        print(f"\n**{title}**")


def same_on_every_rank(tensor) -> bool:
    everyone = gather(tensor[None].clone().to(PartialState().device)).cpu()
    return bool((everyone == everyone[0]).all())


are_the_same_tensors = same_on_every_rank  # name used by other scripts


# ------------------------------------------------------------------------------------------------- who runs what
def _captured(call) -> str:
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        call()
    return sink.getvalue().strip()


def process_execution_check():
    accelerator = Accelerator()
    me, world = accelerator.process_index, accelerator.num_processes
    with accelerator.main_process_first():
        pass
    accelerator.wait_for_everyone()
    cases = [
        (accelerator.on_main_process, {}, accelerator.is_main_process, "main"),
        (accelerator.on_local_main_process, {}, accelerator.is_local_main_process, "local main"),
        (accelerator.on_last_process, {}, accelerator.is_last_process, "last"),
    ] + [(accelerator.on_process, {"process_index": rank}, me == rank, f"number {rank}") for rank in range(world)]
    for decorator, options, mine, who in cases:
        text = f"the {who} process speaking: {me}"
        wrapped = decorator(lambda: print(text), **options) if options else decorator(lambda: print(text))
        said = _captured(wrapped)
        assert said == (text if mine else ""), f"rank {me}, decorator for the {who} process printed {said!r}"


def init_state_check():
    state = AcceleratorState()
    if state.local_process_index == 0:
        print("Testing, testing. 1, 2, 3.")
        print(state)


# ------------------------------------------------------------------------------------------------- randomness
def rng_sync_check():
    state = AcceleratorState()
    synchronize_rng_states(["torch"])
    assert same_on_every_rank(torch.get_rng_state()), "RNG states improperly synchronized on CPU."
    if state.distributed_type == DistributedType.MULTI_GPU:
        synchronize_rng_states(["cuda"])
        assert same_on_every_rank(torch.cuda.get_rng_state()), "RNG states improperly synchronized on GPU."
    generator = torch.Generator()
    synchronize_rng_states(["generator"], generator=generator)
    assert same_on_every_rank(generator.get_state()), "RNG states improperly synchronized in generator."
    if state.local_process_index == 0:
        print("All rng are properly synched.")


# ------------------------------------------------------------------------------------------------- dataloaders
def _everything_the_ranks_saw(loader):
    return torch.cat([gather(batch) for batch in loader]).cpu()


def dl_preparation_check():
    state = AcceleratorState()
    samples = 32 * state.num_processes
    for split_batches in (False, True):
        for shuffle in (False, True):
            batch_size = 8 * (state.num_processes if split_batches else 1)
            loader = prepare_data_loader(DataLoader(range(samples), batch_size=batch_size, shuffle=shuffle), state.device, state.num_processes,
                                         state.process_index, put_on_device=True, split_batches=split_batches)
            seen = _everything_the_ranks_saw(loader)
            expected = torch.arange(samples)
            assert torch.equal(seen.sort().values if shuffle else seen, expected), "Wrong dataloader sharding."
    if state.process_index == 0:
        print("Non-shuffled and shuffled dataloaders passing.")


def central_dl_preparation_check():
    state = AcceleratorState()
    samples = 32 * state.num_processes
    for split_batches in (False, True):
        batch_size = 8 * (state.num_processes if split_batches else 1)
        loader = prepare_data_loader(DataLoader(range(samples), batch_size=batch_size), state.device, state.num_processes, state.process_index,
                                     put_on_device=True, dispatch_batches=True, split_batches=split_batches)
        assert torch.equal(_everything_the_ranks_saw(loader), torch.arange(samples)), "Wrong central dataloader."
    if state.process_index == 0:
        print("Central (dispatch) dataloaders passing.")


# ------------------------------------------------------------------------------------------------- training
def _fit(model, optimizer, loader, backward, epochs=3):
    for _ in range(epochs):
        for batch in loader:
            model.zero_grad()
            backward(torch.nn.functional.mse_loss(model(batch["x"]), batch["y"]))
            optimizer.step()


def mock_training(length, batch_size, generator, use_seedable_sampler=False):
    """The baseline: plain PyTorch on one process over the whole batch."""
    set_seed(42)
    generator.manual_seed(42)
    data = RegressionDataset(length=length, seed=42)
    model = RegressionModel()
    _fit(model, torch.optim.SGD(model.parameters(), lr=0.1), DataLoader(data, batch_size=batch_size, shuffle=True, generator=generator),
         lambda loss: loss.backward())
    return data, model


def training_check(use_seedable_sampler=False):
    state = AcceleratorState()
    world = state.num_processes
    generator = torch.Generator()
    per_rank = 8
    data, baseline = mock_training(per_rank * 4 * world, per_rank * world, generator, use_seedable_sampler)
    assert same_on_every_rank(baseline.a) and same_on_every_rank(baseline.b), "Did not obtain the same model on both processes."
    for split_batches in (False, True):
        AcceleratorState._reset_state(True)
        GradientState._reset_state()
        accelerator = Accelerator(dataloader_config=DataLoaderConfiguration(split_batches=split_batches))
        loader = DataLoader(data, batch_size=per_rank * (world if split_batches else 1), shuffle=True, generator=generator)
        model = RegressionModel()
        optimizer = torch.optim.SGD(model.parameters(), lr=0.1)
        loader, model, optimizer = accelerator.prepare(loader, model, optimizer)
        set_seed(42)
        generator.manual_seed(42)
        _fit(model, optimizer, loader, accelerator.backward)
        trained = accelerator.unwrap_model(model).cpu()
        for name in ("a", "b"):
            want, got = getattr(baseline, name), getattr(trained, name)
            assert torch.allclose(want, got, atol=1e-4), f"Did not obtain the same model on CPU or distributed training ({want} vs {got})."
        accelerator.print(f"Training yielded the same results on one CPU or distributed setup with {'batch split' if split_batches else 'no batch split'}.")


# ------------------------------------------------------------------------------------------------- splitting work
def test_split_between_processes_list():
    state = AcceleratorState()
    world, me = state.num_processes, state.process_index
    with state.split_between_processes(list(range(2 * world))) as mine:
        assert len(mine) == 2, f"Each process did not have two items. Process index: {me}; Length: {len(mine)}"
    uneven = list(range(3 * world - 1))
    with state.split_between_processes(uneven, apply_padding=True) as mine:
        if state.is_last_process:
            assert len(mine) == math.ceil(len(uneven) / world), f"Last process did not get the extra item(s). Process index: {me}; Length: {len(mine)}"
    state.wait_for_everyone()


def test_split_between_processes_nested_dict():
    state = AcceleratorState()
    world, me = state.num_processes, state.process_index
    if 8 % world == 0:
        whole = {"a": list(range(1, 9)), "b": list("abcdefgh"), "c": torch.arange(1, 9)}
        share = 8 // world
        window = slice(me * share, (me + 1) * share)
        with state.split_between_processes(dict(whole)) as mine:
            assert mine["a"] == whole["a"][window] and mine["b"] == whole["b"][window] and torch.equal(mine["c"], whole["c"][window])
    state.wait_for_everyone()


def test_trigger():
    accelerator = Accelerator()
    assert accelerator.check_trigger() is False
    if accelerator.is_main_process:
        accelerator.set_trigger()
    # one rank raised the flag: every rank sees it, once
    assert accelerator.check_trigger() is True
    assert accelerator.check_trigger() is False


def test_reinstantiated_state():
    """After a state reset an Accelerator built earlier must fail with a message that says so, not with a bare AttributeError."""
    AcceleratorState._reset_state()
    accelerator = Accelerator()
    try:
        accelerator.prepare(torch.nn.Linear(1, 1))
    except AttributeError as error:
        assert "`AcceleratorState` object has no attribute" in str(error)


def main():
    accelerator = Accelerator()
    state = accelerator.state
    announce("Initialization")
    init_state_check()
    state.wait_for_everyone()
    for title, checks in (
        ("Test process execution", (process_execution_check,)),
        ("Test split between processes as a list", (test_split_between_processes_list,)),
        ("Test split between processes as a dict", (test_split_between_processes_nested_dict,)),
        ("Test random number generator synchronization", (rng_sync_check,)),
        ("DataLoader integration test", (dl_preparation_check, central_dl_preparation_check)),
        ("Training integration test", (training_check,)),
        ("Breakpoint trigger test", (test_trigger,)),
    ):
        announce(title)
        for check in checks:
            check()
    PartialState().wait_for_everyone()
    if PartialState().is_main_process:
        print("test_script: all passed")
    PartialState().destroy_process_group()


if __name__ == "__main__":
    main()
