"""End-to-end sanity script run by ``accelerate test`` and the launcher tests (reference
src/accelerate/test_utils/scripts/test_script.py:827-905): process control, RNG sync, dataloader
sharding against a baseline, ``split_between_processes``, single-vs-distributed training parity, triggers."""

import io
import math
from contextlib import redirect_stdout

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.data_loader import SeedableRandomSampler, prepare_data_loader
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState
from accelerate_hpc_test_b200.test_utils.training import RegressionDataset, RegressionModel
from accelerate_hpc_test_b200.utils import DistributedType, gather, set_seed, synchronize_rng_states


def print_main(state):
    print(f"Printing from the main process {state.process_index}")


def print_local_main(state):
    print(f"Printing from the local main process {state.local_process_index}")


def print_last(state):
    print(f"Printing from the last process {state.process_index}")


def print_on(state, process_idx):
    print(f"Printing from process {process_idx}: {state.process_index}")


def process_execution_check():
    accelerator = Accelerator()
    num_processes = accelerator.num_processes
    f = io.StringIO()
    with accelerator.main_process_first():
        with redirect_stdout(f):
            pass
    accelerator.wait_for_everyone()
    f = io.StringIO()
    with redirect_stdout(f):
        accelerator.on_main_process(print_main)(accelerator.state)
    result = f.getvalue().rstrip()
    if accelerator.is_main_process:
        assert result == "Printing from the main process 0", f"{result} != Printing from the main process 0"
    else:
        assert f.getvalue().rstrip() == "", f'{result} != ""'
    f = io.StringIO()
    with redirect_stdout(f):
        accelerator.on_last_process(print_last)(accelerator.state)
    if accelerator.is_last_process:
        assert f.getvalue().rstrip() == f"Printing from the last process {num_processes - 1}"
    else:
        assert f.getvalue().rstrip() == ""
    for process_idx in range(num_processes):
        f = io.StringIO()
        with redirect_stdout(f):
            accelerator.on_process(print_on, process_index=process_idx)(accelerator.state, process_idx)
        if accelerator.process_index == process_idx:
            assert f.getvalue().rstrip() == f"Printing from process {process_idx}: {accelerator.process_index}"
        else:
            assert f.getvalue().rstrip() == ""


def init_state_check():
    state = AcceleratorState()
    if state.local_process_index == 0:
        print("Testing, testing. 1, 2, 3.")
        print(state)


def rng_sync_check():
    state = AcceleratorState()
    synchronize_rng_states(["torch"])
    assert are_the_same_tensors(torch.get_rng_state()), "RNG states improperly synchronized on CPU."
    if state.distributed_type == DistributedType.MULTI_GPU:
        synchronize_rng_states(["cuda"])
        assert are_the_same_tensors(torch.cuda.get_rng_state()), "RNG states improperly synchronized on GPU."
    generator = torch.Generator()
    synchronize_rng_states(["generator"], generator=generator)
    assert are_the_same_tensors(generator.get_state()), "RNG states improperly synchronized in generator."
    if state.local_process_index == 0:
        print("All rng are properly synched.")


def are_the_same_tensors(tensor):
    state = AcceleratorState()
    tensor = tensor[None].clone().to(state.device)
    tensors = gather(tensor).cpu()
    tensor = tensor[0].cpu()
    for i in range(tensors.shape[0]):
        if not torch.equal(tensors[i], tensor):
            return False
    return True


def dl_preparation_check():
    state = AcceleratorState()
    length = 32 * state.num_processes
    for split_batches in (False, True):
        for shuffle in (False, True):
            dl = DataLoader(range(length), batch_size=8 * (state.num_processes if split_batches else 1), shuffle=shuffle)
            dl = prepare_data_loader(dl, state.device, state.num_processes, state.process_index, put_on_device=True, split_batches=split_batches)
            result = []
            for batch in dl:
                result.append(gather(batch))
            result = torch.cat(result)
            assert torch.equal(result.cpu().sort().values if shuffle else result.cpu(), torch.arange(0, length).long()), "Wrong dataloader sharding."
    if state.process_index == 0:
        print("Non-shuffled and shuffled dataloaders passing.")


def central_dl_preparation_check():
    state = AcceleratorState()
    length = 32 * state.num_processes
    for split_batches in (False, True):
        dl = DataLoader(range(length), batch_size=8 * (state.num_processes if split_batches else 1))
        dl = prepare_data_loader(dl, state.device, state.num_processes, state.process_index, put_on_device=True, dispatch_batches=True,
                                 split_batches=split_batches)
        result = []
        for batch in dl:
            result.append(gather(batch))
        result = torch.cat(result)
        assert torch.equal(result.cpu(), torch.arange(0, length).long()), "Wrong central dataloader."
    if state.process_index == 0:
        print("Central (dispatch) dataloaders passing.")


def mock_training(length, batch_size, generator, use_seedable_sampler=False):
    set_seed(42)
    generator.manual_seed(42)
    train_set = RegressionDataset(length=length, seed=42)
    train_dl = DataLoader(train_set, batch_size=batch_size, shuffle=True, generator=generator)
    model = RegressionModel()
    optimizer = torch.optim.SGD(model.parameters(), lr=0.1)
    for epoch in range(3):
        for batch in train_dl:
            model.zero_grad()
            output = model(batch["x"])
            loss = torch.nn.functional.mse_loss(output, batch["y"])
            loss.backward()
            optimizer.step()
    return train_set, model


def training_check(use_seedable_sampler=False):
    state = AcceleratorState()
    generator = torch.Generator()
    batch_size = 8
    length = batch_size * 4 * state.num_processes
    train_set, old_model = mock_training(length, batch_size * state.num_processes, generator, use_seedable_sampler)
    assert are_the_same_tensors(old_model.a), "Did not obtain the same model on both processes."
    assert are_the_same_tensors(old_model.b), "Did not obtain the same model on both processes."
    from accelerate_hpc_test_b200.utils import DataLoaderConfiguration

    for split_batches in (False, True):
        AcceleratorState._reset_state(True)
        GradientState._reset_state()
        accelerator = Accelerator(dataloader_config=DataLoaderConfiguration(split_batches=split_batches))
        train_dl = DataLoader(train_set, batch_size=batch_size * (state.num_processes if split_batches else 1), shuffle=True, generator=generator)
        model = RegressionModel()
        optimizer = torch.optim.SGD(model.parameters(), lr=0.1 if split_batches else 0.1)
        train_dl, model, optimizer = accelerator.prepare(train_dl, model, optimizer)
        set_seed(42)
        generator.manual_seed(42)
        for epoch in range(3):
            for batch in train_dl:
                model.zero_grad()
                output = model(batch["x"])
                loss = torch.nn.functional.mse_loss(output, batch["y"])
                accelerator.backward(loss)
                optimizer.step()
        model = accelerator.unwrap_model(model).cpu()
        assert torch.allclose(old_model.a, model.a, atol=1e-4), f"Did not obtain the same model on CPU or distributed training ({old_model.a} vs {model.a})."
        assert torch.allclose(old_model.b, model.b, atol=1e-4), f"Did not obtain the same model on CPU or distributed training ({old_model.b} vs {model.b})."
        accelerator.print(f"Training yielded the same results on one CPU or distributed setup with {'batch split' if split_batches else 'no batch split'}.")


def test_split_between_processes_list():
    state = AcceleratorState()
    data = list(range(0, 2 * state.num_processes))
    with state.split_between_processes(data) as results:
        assert len(results) == 2, f"Each process did not have two items. Process index: {state.process_index}; Length: {len(results)}"
    data = list(range(0, (3 * state.num_processes) - 1))
    with state.split_between_processes(data, apply_padding=True) as results:
        if state.is_last_process:
            num_samples_per_device = math.ceil(len(data) / state.num_processes)
            assert len(results) == num_samples_per_device, f"Last process did not get the extra item(s). Process index: {state.process_index}; Length: {len(results)}"
    state.wait_for_everyone()


def test_split_between_processes_nested_dict():
    state = AcceleratorState()
    a = [1, 2, 3, 4, 5, 6, 7, 8]
    b = ["a", "b", "c", "d", "e", "f", "g", "h"]
    c = torch.tensor([1, 2, 3, 4, 5, 6, 7, 8])
    if state.num_processes in (1, 2, 4):
        data = {"a": a, "b": b, "c": c}
        data_copy = dict(data)
        with state.split_between_processes(data) as results:
            per = 8 // state.num_processes
            lo = state.process_index * per
            assert results["a"] == data_copy["a"][lo : lo + per]
            assert results["b"] == data_copy["b"][lo : lo + per]
            assert torch.allclose(results["c"], data_copy["c"][lo : lo + per])
    state.wait_for_everyone()


def test_trigger():
    accelerator = Accelerator()
    assert accelerator.check_trigger() is False
    if accelerator.is_main_process:
        accelerator.set_trigger()
    assert accelerator.check_trigger() is True
    assert accelerator.check_trigger() is False


def test_reinstantiated_state():
    AcceleratorState._reset_state()
    simple_model = torch.nn.Linear(1, 1)
    accelerator = Accelerator()
    try:
        accelerator.prepare(simple_model)
    except AttributeError as e:
        assert "`AcceleratorState` object has no attribute" in str(e)
    else:
        pass


def main():
    accelerator = Accelerator()
    state = accelerator.state
    if state.local_process_index == 0:
        print("**Initialization**")
    init_state_check()
    state.wait_for_everyone()
    if state.local_process_index == 0:
        print("\n**Test process execution**")
    process_execution_check()
    if state.local_process_index == 0:
        print("\n**Test split between processes as a list**")
    test_split_between_processes_list()
    if state.local_process_index == 0:
        print("\n**Test split between processes as a dict**")
    test_split_between_processes_nested_dict()
    if state.local_process_index == 0:
        print("\n**Test random number generator synchronization**")
    rng_sync_check()
    if state.local_process_index == 0:
        print("\n**DataLoader integration test**")
    dl_preparation_check()
    central_dl_preparation_check()
    if state.local_process_index == 0:
        print("\n**Training integration test**")
    training_check()
    if state.local_process_index == 0:
        print("\n**Breakpoint trigger test**")
    test_trigger()
    PartialState().wait_for_everyone()
    if PartialState().is_main_process:
        print("test_script: all passed")
    PartialState().destroy_process_group()


if __name__ == "__main__":
    main()
