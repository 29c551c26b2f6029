"""Launched checks of the tensor-tree collectives (``gather``, ``gather_object``, ``broadcast``, ``pad_across_processes``, ``reduce``,
``copy_tensor_to_devices``) and of the debug-mode shape verification; works on any number of processes, including one.
Behaviour of reference ``src/accelerate/test_utils/scripts/test_ops.py:37-148``.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.test_ops
"""

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


def ramp(state):
    """Rank r holds ``[r*P + 1, ..., r*P + P]``: gathered in rank order the values read 1 .. P*P."""
    world = state.num_processes
    return (torch.arange(world) + 1.0 + world * state.process_index).to(state.device)


def every_rank_sum(world):
    return sum(torch.arange(world) + 1.0 + world * rank for rank in range(world))


def test_gather(state):
    assert gather(ramp(state)).tolist() == list(range(1, state.num_processes**2 + 1))


def test_gather_object(state):
    assert gather_object([state.process_index]) == list(range(state.num_processes))


def test_gather_non_contiguous(state):
    view = torch.arange(12, device=state.device).view(4, 3).t()
    assert not view.is_contiguous()
    assert gather(view).shape[0] == 3 * state.num_processes  # made contiguous on the way, not refused


def test_broadcast(state):
    out = broadcast(ramp(state))
    assert out.tolist() == list(range(1, state.num_processes + 1))  # everybody now holds rank 0's values


def test_pad_across_processes(state):
    world = state.num_processes
    mine = torch.arange(world + 1 if state.is_main_process else world).to(state.device)  # rank 0 is one longer
    # (collectives: every rank makes both calls, only the shorter ranks have something to check)
    at_the_end = pad_across_processes(mine)
    in_front = pad_across_processes(mine, pad_first=True, pad_index=-1)
    assert at_the_end.shape == in_front.shape == (world + 1,)
    if not state.is_main_process:
        assert at_the_end.tolist() == [*range(world), 0] and in_front.tolist() == [-1, *range(world)]


def test_reduce_sum(state):
    assert torch.allclose(reduce(ramp(state), "sum").cpu(), every_rank_sum(state.num_processes))


def test_reduce_mean(state):
    world = state.num_processes
    assert torch.allclose(reduce(ramp(state), "mean").cpu(), every_rank_sum(world) / world)


def test_op_checker(state):
    """In debug mode the collectives compare operand shapes across ranks first and raise instead of hanging."""
    if state.num_processes < 2:
        return
    state.debug = True
    try:
        width = 5 if state.process_index == 0 else 6
        shape = (1, width) if state.process_index == 0 else (1, 1, width)
        data = {"tensor": torch.arange(width, dtype=torch.float32).reshape(shape).to(state.device)}
        for call in (lambda: gather(data), lambda: broadcast(data), lambda: reduce(data), lambda: pad_across_processes(data, dim=0)):
            try:
                call()
            except DistributedOperationException:
                continue
            raise AssertionError("shape mismatch went unnoticed in debug mode")
    finally:
        state.debug = False


def test_copy_tensor_to_devices(state):
    only_on_main = torch.tensor([1, 2, 3], dtype=torch.int).to(state.device) if state.is_main_process else None
    assert copy_tensor_to_devices(only_on_main).cpu().tolist() == [1, 2, 3]


CHECKS = (test_gather, test_gather_object, test_gather_non_contiguous, test_broadcast, test_pad_across_processes, test_reduce_sum,
          test_reduce_mean, test_op_checker, test_copy_tensor_to_devices)


def _mp_fn(index):
    main()


def main():
    state = PartialState()
    state.print(f"State: {state}")
    for check in CHECKS:
        state.print(f"running {check.__name__}")
        check(state)
    state.wait_for_everyone()
    state.print("test_ops: all passed")


if __name__ == "__main__":
    main()
