"""Index arithmetic of the sharded samplers / datasets (modelled on reference tests/test_data_loader.py:
exact expected index lists for every rank)."""

import random

import pytest
import torch
from torch.utils.data import BatchSampler, DataLoader, IterableDataset

from accelerate_hpc_test_b200.data_loader import (
    BatchSamplerShard,
    DataLoaderShard,
    IterableDatasetShard,
    SeedableRandomSampler,
    SkipBatchSampler,
    SkipDataLoader,
    prepare_data_loader,
    skip_first_batches,
)
from accelerate_hpc_test_b200.state import GradientState, PartialState


def shards(n, batch_size, drop_last, **kw):
    bs = BatchSampler(range(n), batch_size=batch_size, drop_last=drop_last)
    return [list(BatchSamplerShard(bs, 2, i, **kw)) for i in range(2)]


def test_no_split_even_division():
    assert shards(24, 3, False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17], [21, 22, 23]],
    ]
    assert shards(24, 3, True) == shards(24, 3, False)


def test_no_split_wraps_around_when_batches_do_not_divide():
    # 7 batches: rank 1's last batch is minted from the start of the epoch
    assert shards(21, 3, False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17], [0, 1, 2]],
    ]
    # with drop_last the incomplete round disappears
    assert shards(21, 3, True) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17]],
    ]


def test_no_split_tops_up_a_short_last_batch():
    assert shards(22, 3, False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17], [21, 0, 1]],
    ]
    assert shards(20, 3, False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 0]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17], [1, 2, 3]],
    ]
    assert shards(22, 3, True) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17]],
    ]


def test_no_split_tiny_datasets():
    assert shards(2, 3, False) == [[[0, 1, 0]], [[1, 0, 1]]]
    assert shards(2, 3, True) == [[], []]


def test_no_split_uneven_batches():
    assert shards(21, 3, False, even_batches=False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17]],
    ]
    assert shards(22, 3, False, even_batches=False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19, 20]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17], [21]],
    ]
    assert shards(20, 3, False, even_batches=False) == [
        [[0, 1, 2], [6, 7, 8], [12, 13, 14], [18, 19]],
        [[3, 4, 5], [9, 10, 11], [15, 16, 17]],
    ]
    assert shards(2, 3, False, even_batches=False) == [[[0, 1]], []]


def test_split_batches():
    assert shards(24, 4, False, split_batches=True) == [
        [[0, 1], [4, 5], [8, 9], [12, 13], [16, 17], [20, 21]],
        [[2, 3], [6, 7], [10, 11], [14, 15], [18, 19], [22, 23]],
    ]
    # last batch [20, 21] is short: completed with [0, 1]
    assert shards(22, 4, False, split_batches=True) == [
        [[0, 1], [4, 5], [8, 9], [12, 13], [16, 17], [20, 21]],
        [[2, 3], [6, 7], [10, 11], [14, 15], [18, 19], [0, 1]],
    ]
    assert shards(21, 4, False, split_batches=True) == [
        [[0, 1], [4, 5], [8, 9], [12, 13], [16, 17], [20, 0]],
        [[2, 3], [6, 7], [10, 11], [14, 15], [18, 19], [1, 2]],
    ]
    assert shards(22, 4, True, split_batches=True) == [
        [[0, 1], [4, 5], [8, 9], [12, 13], [16, 17]],
        [[2, 3], [6, 7], [10, 11], [14, 15], [18, 19]],
    ]
    assert shards(22, 4, False, split_batches=True, even_batches=False) == [
        [[0, 1], [4, 5], [8, 9], [12, 13], [16, 17], [20, 21]],
        [[2, 3], [6, 7], [10, 11], [14, 15], [18, 19]],
    ]
    assert shards(2, 4, False, split_batches=True) == [[[0, 1]], [[0, 1]]]


def test_split_batches_needs_divisible_batch_size():
    with pytest.raises(ValueError):
        BatchSamplerShard(BatchSampler(range(10), 3, False), 2, 0, split_batches=True)


def test_batch_sampler_lengths_match_iteration():
    for n in (2, 7, 20, 21, 22, 24, 25):
        for bs in (3, 4):
            for drop in (False, True):
                for even in (False, True):
                    for split in (False, True):
                        if split and (bs % 2 or not even):
                            continue  # split + uneven: late ranks may get an empty tail slice (as in the reference)
                        for i in range(2):
                            sh = BatchSamplerShard(BatchSampler(range(n), bs, drop), 2, i, split_batches=split, even_batches=even)
                            assert len(sh) == len(list(sh)), (n, bs, drop, even, split, i)


def test_batch_sampler_with_varying_batch_size():
    batch_sampler = [[0, 1, 2], [3, 4], [5, 6, 7, 8], [9, 10, 11], [12, 13]]
    got = [list(BatchSamplerShard(batch_sampler, 2, i, even_batches=False)) for i in range(2)]
    assert got == [[[0, 1, 2], [5, 6, 7, 8], [12, 13]], [[3, 4], [9, 10, 11]]]


class RandomIterableDataset(IterableDataset):
    def __init__(self, p_stop=0.01, max_length=1000):
        self.p_stop, self.max_length = p_stop, max_length

    def __iter__(self):
        count, stop = 0, False
        while not stop and count < self.max_length:
            yield count
            count += 1
            stop = random.random() < self.p_stop


@pytest.mark.parametrize("batch_size,drop_last,split_batches", [(4, False, False), (4, True, False), (4, False, True), (4, True, True)])
def test_iterable_dataset_shard(batch_size, drop_last, split_batches):
    random.seed(42)
    dataset = RandomIterableDataset(max_length=2 if False else 1000)
    num_processes = 2
    random.seed(7)
    reference = list(dataset)
    outs = []
    for i in range(num_processes):
        random.seed(7)
        outs.append(list(IterableDatasetShard(dataset, batch_size=batch_size, drop_last=drop_last, num_processes=num_processes,
                                              process_index=i, split_batches=split_batches)))
    shard_bs = batch_size // num_processes if split_batches else batch_size
    assert len(outs[0]) == len(outs[1]) and len(outs[0]) % shard_bs == 0
    observed = []
    for idx in range(0, len(outs[0]), shard_bs):
        for o in outs:
            observed += o[idx : idx + shard_bs]
    if not drop_last:
        while len(reference) < len(observed):
            reference += reference
    assert observed == reference[: len(observed)]


def test_iterable_dataset_shard_tiny():
    class Two(IterableDataset):
        def __iter__(self):
            yield from range(2)

    outs = [list(IterableDatasetShard(Two(), batch_size=4, num_processes=2, process_index=i)) for i in range(2)]
    assert outs == [[0, 1, 0, 1], [0, 1, 0, 1]]


def test_seedable_sampler_is_reproducible_per_epoch():
    a = SeedableRandomSampler(range(20), data_seed=3)
    b = SeedableRandomSampler(range(20), data_seed=3)
    e0 = list(a)
    assert e0 == list(b)
    e1 = list(a)
    assert e1 != e0 and sorted(e1) == list(range(20))
    b.set_epoch(1)
    assert list(b) == e1


def test_skip_batch_sampler_and_loader():
    bs = SkipBatchSampler(BatchSampler(range(16), 4, False), 2)
    assert list(bs) == [[8, 9, 10, 11], [12, 13, 14, 15]] and len(bs) == 2
    dl = SkipDataLoader(list(range(16)), batch_size=4, skip_batches=2)
    assert [t.tolist() for t in dl] == [[8, 9, 10, 11], [12, 13, 14, 15]]
    new = skip_first_batches(DataLoader(list(range(16)), batch_size=4), num_batches=2)
    assert [t.tolist() for t in new] == [[8, 9, 10, 11], [12, 13, 14, 15]]


def test_dataloader_shard_end_of_dataloader_and_remainder():
    PartialState()
    dl = DataLoaderShard(list(range(10)), batch_size=4)
    seen = []
    for i, batch in enumerate(dl):
        seen.append((batch.tolist(), dl.end_of_dataloader))
    assert [s[1] for s in seen] == [False, False, True]
    assert dl.remainder == 10 % 4
    # state is shared with GradientState while iterating
    gs = GradientState()
    assert not gs.in_dataloader
    it = iter(dl)
    next(it)
    assert gs.in_dataloader and gs.active_dataloader is dl
    list(it)
    assert not gs.in_dataloader


def test_prepare_data_loader_single_process_roundtrip():
    PartialState()
    dl = DataLoader(list(range(12)), batch_size=4, shuffle=False)
    prepared = prepare_data_loader(dl, device=torch.device("cpu"), put_on_device=True)
    assert isinstance(prepared, DataLoader)  # adapters keep the isinstance contract
    assert [b.tolist() for b in prepared] == [[0, 1, 2, 3], [4, 5, 6, 7], [8, 9, 10, 11]]
    again = skip_first_batches(prepared, 1)
    assert [b.tolist() for b in again] == [[4, 5, 6, 7], [8, 9, 10, 11]]


def test_prepare_data_loader_shards_for_virtual_ranks():
    PartialState()
    dl = DataLoader(list(range(16)), batch_size=2, shuffle=False)
    outs = []
    for r in range(4):
        p = prepare_data_loader(dl, num_processes=4, process_index=r)
        outs.append([b.tolist() for b in p])
    assert outs[0] == [[0, 1], [8, 9]] and outs[3] == [[6, 7], [14, 15]]
    assert p.total_batch_size == 8
