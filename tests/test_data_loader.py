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


# ------------------------------------------------------------------------------------------------- more of the layer
class _Stream(IterableDataset):
    def __init__(self, n):
        self.n = n

    def __iter__(self):
        yield from range(self.n)

    def __len__(self):
        return self.n


def test_refill_cycles_over_what_it_remembered():
    from accelerate_hpc_test_b200.data_loader import _Refill

    r = _Refill(4)
    r.remember([0, 1, 2])
    r.remember([3, 4, 5])  # only one more fits
    assert r.pool == [0, 1, 2, 3]
    assert r.take(6) == [0, 1, 2, 3, 0, 1] and r.take(1) == [2]


def test_flag_last_marks_only_the_last_item_and_calls_back_before_each_read_ahead():
    from accelerate_hpc_test_b200.data_loader import _flag_last

    reads = []
    out = list(_flag_last(iter("abc"), before_fetch=lambda: reads.append(1), on_fetch=str.upper))
    assert out == [("A", False), ("B", False), ("C", True)] and len(reads) == 3
    assert list(_flag_last(iter(()))) == []


def test_get_and_set_sampler_reach_through_the_shard_wrapper():
    PartialState()
    base = DataLoader(list(range(12)), batch_size=2, shuffle=True)
    prepared = prepare_data_loader(base, num_processes=2, process_index=0)
    assert prepared.get_sampler() is base.sampler  # found under BatchSamplerShard -> BatchSampler
    new = SeedableRandomSampler(list(range(12)), data_seed=5)
    prepared.set_sampler(new)
    assert prepared.get_sampler() is new and prepared.batch_sampler.batch_sampler.sampler is new


def test_seedable_sampler_through_prepare_gives_disjoint_shards_that_change_with_the_epoch():
    PartialState()
    base = DataLoader(list(range(16)), batch_size=2, shuffle=True)
    loaders = [prepare_data_loader(base, num_processes=2, process_index=r, use_seedable_sampler=True, data_seed=11) for r in range(2)]
    epoch0 = [[b.tolist() for b in dl] for dl in loaders]
    flat = sorted(i for shard in epoch0 for b in shard for i in b)
    assert flat == list(range(16))  # same permutation on both ranks, dealt without overlap
    epoch1 = [[b.tolist() for b in dl] for dl in loaders]
    assert sorted(i for shard in epoch1 for b in shard for i in b) == list(range(16)) and epoch1 != epoch0
    for dl in loaders:
        dl.set_epoch(0)
    assert [[b.tolist() for b in dl] for dl in loaders] == epoch0  # the order is a function of seed + epoch only


def test_set_epoch_reaches_the_dataset_when_nothing_else_listens():
    class Data(_Stream):
        epoch = None

        def set_epoch(self, e):
            self.epoch = e

    PartialState()
    data = Data(8)
    dl = prepare_data_loader(DataLoader(data, batch_size=2), num_processes=2, process_index=1)
    dl.set_epoch(3)
    assert data.epoch == 3 and dl.iteration == 3
    assert [b.tolist() for b in dl] == [[2, 3], [6, 7]]


def test_split_batches_checks():
    PartialState()
    with pytest.raises(ValueError, match="round multiple"):
        prepare_data_loader(DataLoader(list(range(12)), batch_size=3), num_processes=2, process_index=0, split_batches=True)
    with pytest.raises(ValueError, match="requires `put_on_device=True`"):
        prepare_data_loader(DataLoader(list(range(12)), batch_size=2), dispatch_batches=True)
    dl = prepare_data_loader(DataLoader(list(range(12)), batch_size=4), num_processes=2, process_index=1, split_batches=True)
    assert [b.tolist() for b in dl] == [[2, 3], [6, 7], [10, 11]] and dl.total_batch_size == 4 and len(dl) == 3


def test_batch_sampler_passed_as_sampler_keeps_that_convention():
    PartialState()
    base = DataLoader(torch.arange(12), sampler=BatchSampler(range(12), 3, False), batch_size=None)
    dl = prepare_data_loader(base, num_processes=2, process_index=1)
    assert isinstance(dl.sampler, BatchSamplerShard) and dl.batch_size is None
    assert [b.tolist() for b in dl] == [[3, 4, 5], [9, 10, 11]]
    assert [b.tolist() for b in skip_first_batches(dl, 1)] == [[9, 10, 11]]


def test_skip_first_batches_on_a_streaming_shard_reads_through():
    PartialState()
    dl = prepare_data_loader(DataLoader(_Stream(16), batch_size=2), num_processes=2, process_index=0)
    assert [b.tolist() for b in dl] == [[0, 1], [4, 5], [8, 9], [12, 13]]
    resumed = skip_first_batches(dl, 2)
    assert type(resumed) is DataLoaderShard and resumed.skip_batches == 2
    assert [b.tolist() for b in resumed] == [[8, 9], [12, 13]]
    plain = skip_first_batches(DataLoader(_Stream(8), batch_size=2), 3)
    assert type(plain) is SkipDataLoader and [b.tolist() for b in plain] == [[6, 7]] and len(plain) == 1


def test_inherited_kwargs_follow_the_installed_dataloader_signature():
    from accelerate_hpc_test_b200.data_loader import _PYTORCH_DATALOADER_KWARGS, _inherited_kwargs

    assert {"num_workers", "collate_fn", "pin_memory", "timeout", "generator", "persistent_workers"} <= set(_PYTORCH_DATALOADER_KWARGS)
    collate = lambda rows: torch.tensor(rows) * 10  # noqa: E731
    kw = _inherited_kwargs(DataLoader(list(range(4)), batch_size=2, collate_fn=collate))
    assert kw["collate_fn"] is collate and not {"batch_size", "sampler", "batch_sampler", "shuffle", "drop_last"} & set(kw)
    PartialState()
    dl = prepare_data_loader(DataLoader(list(range(8)), batch_size=2, collate_fn=collate), num_processes=2, process_index=1)
    assert [b.tolist() for b in dl] == [[20, 30], [60, 70]]


def test_mesh_data_ranks():
    from accelerate_hpc_test_b200.data_loader import _mesh_data_ranks

    class Axis:
        def __init__(self, n):
            self.n = n

        def size(self):
            return self.n

    class Mesh:
        def __init__(self, extents, coords):
            self.mesh_dim_names, self.extents, self.coords = tuple(extents), extents, coords

        def __getitem__(self, name):
            return Axis(self.extents[name])

        def get_local_rank(self, name):
            return self.coords[name]

        def size(self):
            return math.prod(self.extents.values())

    import math

    # dp_replicate 2 x dp_shard 2 x tp 2: the two tp ranks of a group read the same data; readers are numbered replicate-major
    mesh = Mesh({"dp_replicate": 2, "dp_shard": 2, "tp": 2}, {"dp_replicate": 1, "dp_shard": 0, "tp": 1})
    assert _mesh_data_ranks(mesh, 5, 8) == (2, 4)
    assert _mesh_data_ranks(Mesh({"cp": 4}, {"cp": 3}), 3, 4) == (0, 1)
    assert _mesh_data_ranks(Mesh({"pp": 2, "dp_shard": 2}, {"pp": 1, "dp_shard": 1}), 3, 4) == (3, 4)  # unknown axis: untouched


def _dispatch_ragged():
    from accelerate_hpc_test_b200.utils.operations import gather

    state = PartialState()
    # 22 samples, batches of 4 -> loader batches 4,4,4,4,4,2; two ranks => global batches of 8, 8, and a last one of 6 = 3 + 3
    dl = prepare_data_loader(DataLoader(_Stream(22), batch_size=4), state.device, put_on_device=True, dispatch_batches=True)
    got, flags = [], []
    for batch in dl:
        flags.append(dl.end_of_dataloader)
        got.append(gather(batch).tolist())
    assert got == [list(range(8)), list(range(8, 16)), list(range(16, 22))], got
    assert flags == [False, False, True] and dl.remainder == 6
    # 21 samples: the last global batch has 5 rows -> completed with the first row of the epoch, 3 rows per rank
    dl = prepare_data_loader(DataLoader(_Stream(21), batch_size=4), state.device, put_on_device=True, dispatch_batches=True)
    got = [gather(b).tolist() for b in dl]
    assert got[-1] == [16, 17, 18, 19, 20, 0] and dl.remainder == 5, got
    # drop_last on the loader drops the incomplete round as a whole
    dl = prepare_data_loader(DataLoader(_Stream(22), batch_size=4, drop_last=True), state.device, put_on_device=True, dispatch_batches=True)
    assert [gather(b).tolist() for b in dl] == [list(range(8)), list(range(8, 16))] and len(dl) == 2
    # an empty stream ends the pass on every rank without a batch
    dl = prepare_data_loader(DataLoader(_Stream(0), batch_size=4), state.device, put_on_device=True, dispatch_batches=True)
    assert list(dl) == []
    # resuming: skip the first global batch
    dl = prepare_data_loader(DataLoader(_Stream(22), batch_size=4), state.device, put_on_device=True, dispatch_batches=True)
    assert [gather(b).tolist() for b in skip_first_batches(dl, 1)] == [list(range(8, 16)), list(range(16, 22))]


def test_dispatcher_ragged_tail_two_processes():
    from accelerate_hpc_test_b200.launchers import debug_launcher

    debug_launcher(_dispatch_ragged, num_processes=2)


@pytest.mark.gpu
def test_overlapped_host_to_device_copies_deliver_the_right_batches():
    PartialState()
    data = torch.arange(64 * 1024, dtype=torch.int64).view(64, 1024)
    dl = prepare_data_loader(DataLoader(data, batch_size=8), device=torch.device("cuda", 0), put_on_device=True, non_blocking=True)
    seen = []
    for batch in dl:
        assert batch.is_cuda
        # keep the device busy so that the next copy really runs beside compute
        x = torch.randn(2048, 2048, device="cuda")
        for _ in range(4):
            x = x @ x
        seen.append(batch.sum().item())
    assert seen == [data[i : i + 8].sum().item() for i in range(0, 64, 8)]
