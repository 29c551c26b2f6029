"""Launched checks of how prepared dataloaders deal with data that does not divide evenly between ranks: ``even_batches`` on / off,
``Accelerator.join_uneven_inputs`` (with and without overriding ``even_batches``, with iterable datasets in the mix), pickling an
``Accelerator``, and a prepared loader yielding exactly what the plain one does.  Run with >= 2 processes:

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.test_distributed_data_loop
"""

import pickle
import warnings

import torch
from torch.utils.data import DataLoader, IterableDataset, TensorDataset

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.utils import DataLoaderConfiguration, DistributedType, set_seed


class _Counting(IterableDataset):
    def __init__(self, n):
        self.n = n

    def __iter__(self):
        return iter(torch.arange(self.n).float())


def fresh_accelerator(even_batches=True):
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    accelerator = Accelerator(dataloader_config=DataLoaderConfiguration(even_batches=even_batches))
    assert accelerator.num_processes >= 2, "this script checks behaviour between ranks: launch it with at least two processes"
    return accelerator


def loader(accelerator, size, batch_size, iterable=False):
    data = _Counting(size) if iterable else TensorDataset(torch.arange(size).float())
    return accelerator.prepare(DataLoader(data, batch_size=batch_size))


def batch_sizes_seen(accelerator, size, batch_size):
    return [len(batch[0]) for batch in loader(accelerator, size, batch_size)]


def expect(per_rank, seen, accelerator):
    mine = per_rank[accelerator.process_index] if accelerator.process_index < len(per_rank) else per_rank[-1]
    assert seen == mine, (accelerator.process_index, seen, mine)


def check_even_batches_is_the_default():
    acc = fresh_accelerator()
    # 3 samples, batch 1, two ranks: the second rank's second batch is recycled from the start
    expect([[1, 1], [1, 1]], batch_sizes_seen(acc, 3, 1), acc)
    # 7 samples, batch 2: rank 0 gets [0,1],[4,5]; rank 1 gets [2,3],[6,+0]
    expect([[2, 2], [2, 2]], batch_sizes_seen(acc, 7, 2), acc)


def check_even_batches_can_be_disabled():
    acc = fresh_accelerator(even_batches=False)
    expect([[1, 1], [1]], batch_sizes_seen(acc, 3, 1), acc)
    expect([[2, 2], [2, 1]], batch_sizes_seen(acc, 7, 2), acc)


def check_join_trains_on_uneven_inputs():
    acc = fresh_accelerator(even_batches=False)
    model = acc.prepare(torch.nn.Linear(1, 1))
    dl = loader(acc, 3, 1)
    seen = []
    with acc.join_uneven_inputs([model]):
        for (x,) in dl:
            model(x[:, None]).sum().backward()  # the rank that runs out of data shadows the collectives of the others
            seen.append(int(x[0]))
    acc.wait_for_everyone()
    expect([[0, 2], [1]], seen, acc)


def check_join_overrides_even_batches():
    acc = fresh_accelerator(even_batches=True)
    model = acc.prepare(torch.nn.Linear(1, 1))
    train, valid = loader(acc, 3, 1), loader(acc, 3, 1)
    with acc.join_uneven_inputs([model], even_batches=False):
        inside = (train.batch_sampler.even_batches, valid.batch_sampler.even_batches)
    assert inside == (False, False) and train.batch_sampler.even_batches and valid.batch_sampler.even_batches  # restored on exit


def check_join_with_iterable_loaders_in_the_mix():
    acc = fresh_accelerator(even_batches=True)
    model = acc.prepare(torch.nn.Linear(1, 1))
    loader(acc, 3, 1, iterable=True)
    mapped = loader(acc, 3, 1)
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        with acc.join_uneven_inputs([model], even_batches=False):
            inside = mapped.batch_sampler.even_batches
    assert inside is False and mapped.batch_sampler.even_batches is True
    assert any("only supported for map-style datasets" in str(w.message) for w in caught)  # the iterable one cannot be overridden


def check_accelerator_pickles():
    acc = fresh_accelerator()
    dl = loader(acc, 8, 2)
    clone = pickle.loads(pickle.dumps(acc))
    assert clone.state.__dict__ == acc.state.__dict__ and len(clone._dataloaders) == 1
    assert [b[0].tolist() for b in pickle.loads(pickle.dumps(dl))] == [b[0].tolist() for b in dl]


def check_prepared_loader_covers_the_data_once():
    acc = fresh_accelerator()
    set_seed(3)
    data = TensorDataset(torch.arange(16 * acc.num_processes).float())
    seen = torch.cat([acc.gather(batch[0]) for batch in acc.prepare(DataLoader(data, batch_size=4, shuffle=True))])
    assert sorted(seen.tolist()) == list(range(16 * acc.num_processes))


def main():
    acc = fresh_accelerator()
    if acc.distributed_type not in (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
        raise SystemExit("join_uneven_inputs is a data-parallel feature: run this script with plain multi-GPU / multi-CPU")
    checks = (check_even_batches_is_the_default, check_even_batches_can_be_disabled, check_join_trains_on_uneven_inputs,
              check_join_overrides_even_batches, check_join_with_iterable_loaders_in_the_mix, check_accelerator_pickles,
              check_prepared_loader_covers_the_data_once)
    for check in checks:
        check()
        AcceleratorState().wait_for_everyone()
        if AcceleratorState().is_main_process:
            print(f"{check.__name__}: ok", flush=True)
    if AcceleratorState().is_main_process:
        print("test_distributed_data_loop: all passed")


if __name__ == "__main__":
    main()
