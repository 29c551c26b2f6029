"""Distributed checks of the collective helpers; run under any launcher (reference
src/accelerate/test_utils/scripts/test_ops.py:37-148)."""

import torch

from accelerate_hpc_test_b200 import PartialState
from accelerate_hpc_test_b200.utils.operations import (
    DistributedOperationException,
    broadcast,
    copy_tensor_to_devices,
    gather,
    gather_object,
    pad_across_processes,
    reduce,
)


def create_tensor(state):
    return (torch.arange(state.num_processes) + 1.0 + (state.num_processes * state.process_index)).to(state.device)


def test_gather(state):
    tensor = create_tensor(state)
    gathered = gather(tensor)
    assert gathered.tolist() == list(range(1, state.num_processes**2 + 1))


def test_gather_object(state):
    obj = [state.process_index]
    gathered = gather_object(obj)
    assert len(gathered) == state.num_processes and gathered == list(range(state.num_processes))


def test_gather_non_contiguous(state):
    tensor = torch.arange(12, device=state.device).view(4, 3).t()
    assert not tensor.is_contiguous()
    out = gather(tensor)
    assert out.shape[0] == 3 * state.num_processes


def test_broadcast(state):
    tensor = create_tensor(state)
    out = broadcast(tensor)
    assert out.shape == torch.Size([state.num_processes]) and out.tolist() == list(range(1, state.num_processes + 1))


def test_pad_across_processes(state):
    if state.is_main_process:
        tensor = torch.arange(state.num_processes + 1).to(state.device)
    else:
        tensor = torch.arange(state.num_processes).to(state.device)
    padded = pad_across_processes(tensor)
    assert padded.shape == torch.Size([state.num_processes + 1])
    if not state.is_main_process:
        assert padded.tolist() == list(range(0, state.num_processes)) + [0]
    first = pad_across_processes(tensor, pad_first=True, pad_index=-1)
    if not state.is_main_process:
        assert first.tolist() == [-1] + list(range(0, state.num_processes))


def test_reduce_sum(state):
    tensor = create_tensor(state)
    out = reduce(tensor, "sum")
    n = state.num_processes
    expected = sum(torch.arange(n) + 1.0 + n * r for r in range(n))
    assert torch.allclose(out.cpu(), expected), (out, expected)


def test_reduce_mean(state):
    tensor = create_tensor(state)
    out = reduce(tensor, "mean")
    n = state.num_processes
    expected = sum(torch.arange(n) + 1.0 + n * r for r in range(n)) / n
    assert torch.allclose(out.cpu(), expected)


def test_op_checker(state):
    if state.num_processes < 2:
        return
    state.debug = True
    if state.process_index == 0:
        data = {"tensor": torch.tensor([[0.0, 1, 2, 3, 4]]).to(state.device)}
    else:
        data = {"tensor": torch.tensor([[[0.0, 1, 2, 3, 4, 5]]]).to(state.device)}
    for fn in (lambda: gather(data), lambda: broadcast(data), lambda: reduce(data), lambda: pad_across_processes(data, dim=0)):
        try:
            fn()
            raise AssertionError("shape mismatch went unnoticed in debug mode")
        except DistributedOperationException:
            pass
    state.debug = False


def test_copy_tensor_to_devices(state):
    tensor = torch.tensor([1, 2, 3], dtype=torch.int).to(state.device) if state.is_main_process else None
    out = copy_tensor_to_devices(tensor)
    assert torch.allclose(out.cpu(), torch.tensor([1, 2, 3], dtype=torch.int))


def _mp_fn(index):
    main()


def main():
    state = PartialState()
    state.print(f"State: {state}")
    for fn in (test_gather, test_gather_object, test_gather_non_contiguous, test_broadcast, test_pad_across_processes,
               test_reduce_sum, test_reduce_mean, test_op_checker, test_copy_tensor_to_devices):
        state.print(f"running {fn.__name__}")
        fn(state)
    state.wait_for_everyone()
    state.print("test_ops: all passed")


if __name__ == "__main__":
    main()
