"""Data-loading layer: per-rank sharding, rank-0 dispatch, seedable sampling, mid-epoch skip.

Behavioural parity with reference src/accelerate/data_loader.py: ``SeedableRandomSampler`` :73-107,
``BatchSamplerShard`` :110-263, ``IterableDatasetShard`` :266-362, ``DataLoaderStateMixin`` :365-405,
``DataLoaderAdapter`` :408-497, ``DataLoaderShard`` :500-648 (iterates one batch ahead to know
``end_of_dataloader``), ``DataLoaderDispatcher`` :704-959, ``prepare_data_loader`` :996-1309 (rank
arithmetic from the device mesh :1109-1145) and the skip helpers :1312-1451.  Written from the
behaviour up (rounds of ``num_processes`` batches rather than the reference's index bookkeeping);
``tests/test_data_loader.py`` pins the exact index lists.

On CUDA the host->device copy of ``DataLoaderShard`` is issued one batch ahead on a side stream from
pinned memory (``non_blocking``), so the step never waits for its inputs.
"""

from __future__ import annotations

import importlib
import math
from contextlib import suppress
from typing import Callable, Optional, Union

import torch
from torch.utils.data import BatchSampler, DataLoader, IterableDataset, RandomSampler

from .state import GradientState, PartialState
from .utils.dataclasses import DistributedType, RNGType
from .utils.imports import is_torchdata_stateful_dataloader_available
from .utils.operations import (
    broadcast,
    broadcast_object_list,
    concatenate,
    find_batch_size,
    get_data_structure,
    initialize_tensors,
    send_to_device,
    slice_tensors,
)
from .utils.random import synchronize_rng_states

# kwargs copied from the user's DataLoader when it is rebuilt around a sharded sampler
_PYTORCH_DATALOADER_KWARGS = {
    "batch_size": 1, "shuffle": False, "sampler": None, "batch_sampler": None, "num_workers": 0, "collate_fn": None,
    "pin_memory": False, "drop_last": False, "timeout": 0, "worker_init_fn": None, "multiprocessing_context": None,
    "generator": None, "prefetch_factor": None, "persistent_workers": False, "pin_memory_device": "",
}
for _extra in ("in_order",):
    if _extra in DataLoader.__init__.__code__.co_varnames:
        _PYTORCH_DATALOADER_KWARGS[_extra] = True


class SeedableRandomSampler(RandomSampler):
    """``RandomSampler`` whose permutation of epoch ``e`` is a pure function of ``initial_seed + e``,
    so every rank (and a resumed run) draws the same order without broadcasting RNG state."""

    def __init__(self, *args, **kwargs):
        data_seed = kwargs.pop("data_seed", None)
        super().__init__(*args, **kwargs)
        self.initial_seed = data_seed if data_seed is not None else torch.random.initial_seed()
        self.epoch = 0

    def __iter__(self):
        if self.generator is None:
            self.generator = torch.Generator(device=torch.get_default_device() if hasattr(torch, "get_default_device") else "cpu")
            self.generator.manual_seed(self.initial_seed)
        self.generator.manual_seed(self.epoch + self.initial_seed)
        yield from super().__iter__()
        self.set_epoch(self.epoch + 1)

    def set_epoch(self, epoch: int):
        self.epoch = epoch


class BatchSamplerShard(BatchSampler):
    """View of a batch sampler for one of ``num_processes`` ranks.

    ``split_batches=False``: batches are dealt out round-robin (rank r gets batches r, r+P, ...).
    ``split_batches=True``: every batch is cut in P equal slices.
    ``even_batches=True``: when the data does not divide evenly the tail is completed with
    samples from the start of the epoch so every rank sees the same number of equally sized batches.
    """

    def __init__(self, batch_sampler: BatchSampler, num_processes: int = 1, process_index: int = 0,
                 split_batches: bool = False, even_batches: bool = True):
        if split_batches and batch_sampler.batch_size % num_processes != 0:
            raise ValueError(
                f"To use `BatchSamplerShard` in `split_batches` mode, the batch size ({batch_sampler.batch_size}) "
                f"needs to be a round multiple of the number of processes ({num_processes})."
            )
        self.batch_sampler = batch_sampler
        self.num_processes = num_processes
        self.process_index = process_index
        self.split_batches = split_batches
        self.even_batches = even_batches
        self.batch_size = getattr(batch_sampler, "batch_size", None)
        self.drop_last = getattr(batch_sampler, "drop_last", False)
        if self.batch_size is None and self.even_batches:
            raise ValueError(
                "You need to use `even_batches=False` when the batch sampler has no batch size. If you "
                "are not calling this method directly, set `accelerator.even_batches=False` instead."
            )

    @property
    def total_length(self):
        return len(self.batch_sampler)

    def __len__(self):
        n = len(self.batch_sampler)
        if self.split_batches:
            return n
        if n % self.num_processes == 0:
            return n // self.num_processes
        full_rounds = n // self.num_processes
        if self.drop_last:
            return full_rounds
        if self.even_batches:
            return full_rounds + 1
        return full_rounds + 1 if self.process_index < n % self.num_processes else full_rounds

    def __iter__(self):
        return self._iter_with_split() if self.split_batches else self._iter_with_no_split()

    def _is_full(self, batch) -> bool:
        return self.batch_size is None or len(batch) == self.batch_size

    def _iter_with_split(self):
        width = self.batch_sampler.batch_size // self.num_processes
        lo, hi = width * self.process_index, width * (self.process_index + 1)
        first = None
        last = None
        for batch in self.batch_sampler:
            if first is None:
                first = list(batch)
            last = batch
            if len(batch) == self.batch_size:
                yield batch[lo:hi]
        if self.drop_last or first is None or len(last) == self.batch_size:
            return
        # ragged tail
        if not self.even_batches:
            if len(last) > lo:
                yield last[lo:hi]
            return
        filler = list(first)
        while len(filler) < self.batch_size:
            filler += filler
        yield (list(last) + filler)[lo:hi]

    def _iter_with_no_split(self):
        P, me = self.num_processes, self.process_index
        head: list = []  # samples of the first P batches, for the wrap-around completion
        seen_batches = 0
        pending: list = []  # batches of the round being assembled
        for batch in self.batch_sampler:
            if not self.drop_last and seen_batches < P:
                head += list(batch)
            seen_batches += 1
            pending.append(batch)
            if len(pending) == P:
                if self._is_full(batch):
                    yield pending[me]
                    pending = []
                # else: the short batch ends the epoch; handled below
        if self.drop_last or seen_batches == 0 or not pending:
            return
        if not self.even_batches:
            if me < len(pending) and len(pending[me]) > 0:
                yield pending[me]
            return
        # complete the round: top up the short batch, then mint new batches, all from `head`
        while len(head) < P * self.batch_size:
            head += head
        cursor = 0
        completed = [list(b) for b in pending]
        if len(completed[-1]) < self.batch_size:
            need = self.batch_size - len(completed[-1])
            completed[-1] += head[cursor:cursor + need]
            cursor += need
        while len(completed) < P:
            completed.append(head[cursor:cursor + self.batch_size])
            cursor += self.batch_size
        yield completed[me]


class IterableDatasetShard(IterableDataset):
    """Per-rank view of an iterable dataset: buffer one global batch, yield this rank's slice of it."""

    def __init__(self, dataset: IterableDataset, batch_size: int = 1, drop_last: bool = False, num_processes: int = 1,
                 process_index: int = 0, split_batches: bool = False):
        if split_batches and batch_size > 1 and batch_size % num_processes != 0:
            raise ValueError(
                f"To use `IterableDatasetShard` in `split_batches` mode, the batch size ({batch_size}) "
                f"needs to be a round multiple of the number of processes ({num_processes})."
            )
        self.dataset = dataset
        self.batch_size = batch_size
        self.drop_last = drop_last
        self.num_processes = num_processes
        self.process_index = process_index
        self.split_batches = split_batches

    def set_epoch(self, epoch):
        self.epoch = epoch
        if hasattr(self.dataset, "set_epoch"):
            self.dataset.set_epoch(epoch)

    def _sizes(self):
        global_bs = self.batch_size if self.split_batches else self.batch_size * self.num_processes
        local_bs = self.batch_size // self.num_processes if self.split_batches else self.batch_size
        return global_bs, local_bs

    def __len__(self):
        global_bs, local_bs = self._sizes()
        if self.drop_last:
            return (len(self.dataset) // global_bs) * local_bs
        return math.ceil(len(self.dataset) / global_bs) * local_bs

    def __iter__(self):
        if (not hasattr(self.dataset, "set_epoch") and hasattr(self.dataset, "generator")
                and isinstance(self.dataset.generator, torch.Generator)):
            self.dataset.generator.manual_seed(getattr(self, "epoch", 0))
        global_bs, local_bs = self._sizes()
        mine = range(self.process_index * local_bs, (self.process_index + 1) * local_bs)
        first_global = None
        buf = []
        for element in self.dataset:
            buf.append(element)
            if len(buf) == global_bs:
                for i in mine:
                    yield buf[i]
                if first_global is None:
                    first_global = list(buf)
                buf = []
        if not self.drop_last and len(buf) > 0:
            if first_global is None:
                first_global = list(buf)
            while len(buf) < global_bs:
                buf += first_global
            for i in mine:
                yield buf[i]


class DataLoaderStateMixin:
    """``end_of_dataloader`` / ``remainder`` bookkeeping shared by the prepared dataloaders."""

    def __init_subclass__(cls, **kwargs):
        cls.end_of_dataloader = False
        cls.remainder = -1

    def reset(self):
        self.end_of_dataloader = False
        self.remainder = -1

    def begin(self):
        """Register with ``GradientState`` and compute how many samples the last batch duplicates."""
        self.reset()
        with suppress(Exception):
            if not self._drop_last:
                length = getattr(self.dataset, "total_dataset_length", len(self.dataset))
                self.remainder = length % self.total_batch_size
        self.gradient_state._add_dataloader(self)

    def end(self):
        self.gradient_state._remove_dataloader(self)


class DataLoaderAdapter:
    """Composition over a ``DataLoader`` or ``torchdata.StatefulDataLoader`` with attribute forwarding."""

    def __init__(self, dataset, use_stateful_dataloader: bool = False, batch_sampler=None, **kwargs):
        self.use_stateful_dataloader = use_stateful_dataloader
        if use_stateful_dataloader:
            if not is_torchdata_stateful_dataloader_available():
                raise ImportError("StatefulDataLoader is not available. Please install torchdata version 0.8.0 or higher to use it.")
            from torchdata.stateful_dataloader import StatefulDataLoader

            self.base_dataloader = StatefulDataLoader(dataset, batch_sampler=batch_sampler, **kwargs)
        else:
            self.base_dataloader = DataLoader(dataset, batch_sampler=batch_sampler, **kwargs)
        if hasattr(self.base_dataloader, "state_dict"):
            self.dl_state_dict = self.base_dataloader.state_dict()

    def __getattr__(self, name):
        if name == "base_dataloader":
            raise AttributeError()
        return getattr(self.base_dataloader, name)

    def state_dict(self):
        return self.dl_state_dict

    def load_state_dict(self, state_dict):
        self.base_dataloader.load_state_dict(state_dict)

    @property
    def __class__(self):
        # so isinstance(prepared_dl, DataLoader) (or StatefulDataLoader) keeps holding
        return self.base_dataloader.__class__

    def __len__(self):
        return len(self.base_dataloader)

    def adjust_state_dict_for_prefetch(self):
        """A stateful dataloader has already pulled ``prefetch`` batches we have not yielded: rewind the count."""
        if PartialState().distributed_type != DistributedType.NO:
            factor = PartialState().num_processes - 1
            if self.dl_state_dict["_sampler_iter_yielded"] > 0:
                self.dl_state_dict["_sampler_iter_yielded"] -= factor
            if self.dl_state_dict["_num_yielded"] > 0:
                self.dl_state_dict["_num_yielded"] -= factor
            if self.dl_state_dict["_index_sampler_state"] is not None:
                st = self.dl_state_dict["_index_sampler_state"]
                if "samples_yielded" in st and st["samples_yielded"] > 0:
                    st["samples_yielded"] -= self.batch_size * factor

    def _update_state_dict(self):
        if hasattr(self.base_dataloader, "state_dict"):
            self.dl_state_dict = self.base_dataloader.state_dict()
            self.adjust_state_dict_for_prefetch()
            self.dl_state_dict["_iterator_finished"] = self.end_of_dataloader


class DataLoaderShard(DataLoaderAdapter, DataLoaderStateMixin):
    """The per-rank dataloader: synchronises RNG at epoch start, moves batches to the device and knows
    when it is on its last batch (it always holds one batch in hand)."""

    def __init__(self, dataset, device=None, rng_types=None, synchronized_generator=None, skip_batches: int = 0,
                 use_stateful_dataloader: bool = False, _drop_last: bool = False, _non_blocking: bool = False,
                 torch_device_mesh=None, **kwargs):
        super().__init__(dataset, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
        self.device = device
        self.rng_types = rng_types
        self.synchronized_generator = synchronized_generator
        self.skip_batches = skip_batches
        self.gradient_state = GradientState()
        self._drop_last = _drop_last
        self._non_blocking = _non_blocking
        self.iteration = 0

    def __iter__(self):
        if self.rng_types is not None:
            synchronize_rng_states(self.rng_types, self.synchronized_generator)
        self.begin()
        self.set_epoch(self.iteration)
        it = self.base_dataloader.__iter__()
        try:
            current = next(it)
        except StopIteration:
            self.end()
            return
        index = 0
        while True:
            if self.device is not None:
                current = send_to_device(current, self.device, non_blocking=self._non_blocking)
            self._update_state_dict()
            try:
                upcoming = next(it)
            except StopIteration:
                self.end_of_dataloader = True
                self._update_state_dict()
                if index >= self.skip_batches:
                    yield current
                break
            if index >= self.skip_batches:
                yield current
            index += 1
            current = upcoming
        self.iteration += 1
        self.end()

    def __reduce__(self):
        args = super().__reduce__()
        return (DataLoaderShard, *args[1:])

    def set_epoch(self, epoch: int):
        if self.iteration != epoch:
            self.iteration = epoch
        if hasattr(self.batch_sampler, "set_epoch"):
            self.batch_sampler.set_epoch(epoch)
        if hasattr(self.batch_sampler, "sampler") and hasattr(self.batch_sampler.sampler, "set_epoch"):
            self.batch_sampler.sampler.set_epoch(epoch)
        if (hasattr(self.batch_sampler, "batch_sampler") and hasattr(self.batch_sampler.batch_sampler, "sampler")
                and hasattr(self.batch_sampler.batch_sampler.sampler, "set_epoch")):
            self.batch_sampler.batch_sampler.sampler.set_epoch(epoch)
        elif hasattr(self.dataset, "set_epoch"):
            self.dataset.set_epoch(epoch)

    @property
    def total_batch_size(self):
        bs = self.sampler if isinstance(self.sampler, BatchSampler) else self.batch_sampler
        if getattr(bs, "split_batches", False):
            return bs.batch_size
        return bs.batch_size * getattr(bs, "num_processes", 1)

    @property
    def total_dataset_length(self):
        if hasattr(self.dataset, "total_length"):
            return self.dataset.total_length
        return len(self.dataset)

    def get_sampler(self):
        return get_sampler(self)

    def set_sampler(self, sampler):
        sampler_is_batch_sampler = isinstance(self.sampler, BatchSampler)
        if sampler_is_batch_sampler:
            self.sampler.sampler = sampler
        else:
            self.batch_sampler.sampler = sampler
            if hasattr(self.batch_sampler, "batch_sampler"):
                self.batch_sampler.batch_sampler.sampler = sampler


class DataLoaderDispatcher(DataLoaderAdapter, DataLoaderStateMixin):
    """Only rank 0 reads the data: it broadcasts each global batch's structure (pickled) and tensors,
    and every rank keeps its slice.  Useful for iterable datasets that cannot be sharded."""

    def __init__(self, dataset, split_batches: bool = False, skip_batches: int = 0, use_stateful_dataloader: bool = False,
                 _drop_last: bool = False, _non_blocking: bool = False, slice_fn=None, torch_device_mesh=None, **kwargs):
        shuffle = False
        super().__init__(dataset, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
        self.split_batches = split_batches
        self.gradient_state = GradientState()
        self.state = PartialState()
        self._drop_last = _drop_last
        self._non_blocking = _non_blocking
        self.skip_batches = skip_batches
        self.torch_device_mesh = torch_device_mesh
        self.slice_fn = slice_tensors if slice_fn is None else slice_fn
        self.iteration = 0
        # with a mesh, ranks that share tp/cp/sp coordinates must receive the same slice
        self.submesh_tp = self.submesh_dp = self.submesh_fsdp = None
        if torch_device_mesh is not None:
            names = torch_device_mesh.mesh_dim_names or ()
            self.submesh_tp = torch_device_mesh["tp"] if "tp" in names else None
            self.submesh_dp = torch_device_mesh["dp"] if "dp" in names else None
            self.submesh_fsdp = torch_device_mesh["fsdp"] if "fsdp" in names else None
        if self.submesh_tp is not None and (self.submesh_dp or self.submesh_fsdp):
            raise ValueError("TP + (DP/FSDP) is not yet supported in dispatch mode")

    def _fetch_batches(self, iterator):
        """(batch or None, batch_info=[structure, stop_iteration]) — rank 0 reads, everybody learns the structure."""
        batches, batch = None, None
        if self.state.process_index == 0:
            if self.submesh_tp is not None:
                raise ValueError("Dispatch mode with a TP sub-mesh is not supported; use sharded mode")
            try:
                if self.split_batches:
                    self._update_state_dict()
                    batch = next(iterator)
                else:
                    batches = []
                    for _ in range(self.state.num_processes):
                        self._update_state_dict()
                        batches.append(next(iterator))
                    try:
                        batch = concatenate(batches, dim=0)
                    except RuntimeError as e:
                        raise RuntimeError(
                            "You can't use batches of different size with `dispatch_batches=True` or when using an `IterableDataset`."
                            "either pass `dispatch_batches=False` and have each process fetch its own batch "
                            " or pass `split_batches=True`. By doing so, the main process will fetch a full batch and "
                            "slice it into `num_processes` batches for each process."
                        ) from e
                batch_info = [get_data_structure(batch), False]
            except StopIteration:
                batch_info = [None, True]
        else:
            batch_info = [None, self._stop_iteration]
        broadcast_object_list(batch_info)
        self._stop_iteration = batch_info[1]
        if self._stop_iteration:
            # a ragged tail: rank 0 may hold some batches that did not fill the round
            if not self.split_batches and not self._drop_last:
                if self.state.process_index == 0 and batches is not None and len(batches) > 0:
                    batch = concatenate(batches, dim=0)
                    batch_info = [get_data_structure(batch), False]
                else:
                    batch_info = [None, True]
                broadcast_object_list(batch_info)
        return batch, batch_info

    def __iter__(self):
        self.begin()
        self.set_epoch(self.iteration)
        main_iterator = None
        if self.state.process_index == 0:
            main_iterator = self.base_dataloader.__iter__()
        stop_iteration = False
        self._stop_iteration = False
        first_batch = None
        next_batch, next_batch_info = self._fetch_batches(main_iterator)
        batch_index = 0
        while not stop_iteration:
            batch, batch_info = next_batch, next_batch_info
            if self.state.process_index != 0:
                batch = initialize_tensors(batch_info[0])
            batch = send_to_device(batch, self.state.device, non_blocking=self._non_blocking)
            batch = broadcast(batch, from_process=0)
            if not self._drop_last and first_batch is None:
                first_batch = self.slice_fn(batch, slice(0, self.state.num_processes), process_index=self.state.process_index,
                                            num_processes=self.state.num_processes)
            if batch is None:
                raise ValueError(f"Batch does not contain any data (`{batch}`). At the end of all iterable data available before expected stop iteration.")
            observed_batch_size = find_batch_size(batch)
            batch_size = observed_batch_size // self.state.num_processes
            stop_iteration = self._stop_iteration
            if not stop_iteration:
                # peek: is there another global batch?
                next_batch, next_batch_info = self._fetch_batches(main_iterator)
                if self._stop_iteration and next_batch_info[0] is None:
                    stop_iteration = True
            if not self._drop_last and stop_iteration and observed_batch_size % self.state.num_processes != 0:
                # complete the last global batch with samples from the first one
                batch = concatenate([batch, first_batch], dim=0)
                batch_size += 1
            data_slice = slice(self.state.process_index * batch_size, (self.state.process_index + 1) * batch_size)
            batch = self.slice_fn(batch, data_slice, process_index=self.state.process_index, num_processes=self.state.num_processes)
            if stop_iteration:
                self.end_of_dataloader = True
                self._update_state_dict()
                self.remainder = observed_batch_size
            if batch_index >= self.skip_batches:
                yield batch
            batch_index += 1
        self.iteration += 1
        self.end()

    def set_epoch(self, epoch: int):
        if self.iteration != epoch:
            self.iteration = epoch
        if hasattr(self.batch_sampler, "sampler") and hasattr(self.batch_sampler.sampler, "set_epoch"):
            self.batch_sampler.sampler.set_epoch(epoch)
        elif hasattr(self.dataset, "set_epoch"):
            self.dataset.set_epoch(epoch)

    def __len__(self):
        whole_length = len(self.base_dataloader)
        if self.split_batches:
            return whole_length
        if self._drop_last:
            return whole_length // self.state.num_processes
        return math.ceil(whole_length / self.state.num_processes)

    def __reduce__(self):
        args = super().__reduce__()
        return (DataLoaderDispatcher, *args[1:])

    @property
    def total_batch_size(self):
        return self.dataset.batch_size if self.split_batches else (self.dataset.batch_size * self.dataset.num_processes)

    @property
    def total_dataset_length(self):
        return len(self.dataset)

    def get_sampler(self):
        return get_sampler(self)

    def set_sampler(self, sampler):
        sampler_is_batch_sampler = isinstance(self.sampler, BatchSampler)
        if sampler_is_batch_sampler:
            self.sampler.sampler = sampler
        else:
            self.batch_sampler.sampler = sampler
            if hasattr(self.batch_sampler, "batch_sampler"):
                self.batch_sampler.batch_sampler.sampler = sampler


def get_sampler(dataloader):
    """The index sampler underneath whatever batch-sampler wrappers ``dataloader`` carries."""
    sampler_is_batch_sampler = isinstance(dataloader.sampler, BatchSampler)
    if sampler_is_batch_sampler:
        sampler = getattr(dataloader.sampler, "sampler", None)
    else:
        sampler = getattr(dataloader.batch_sampler, "sampler", None)
    return sampler


def _mesh_data_ranks(torch_device_mesh, process_index: int, num_processes: int):
    """(index, count) of this rank among the ranks that must read *different* data.

    Ranks that differ only in tp / cp / sp coordinates consume the same batch; dp_replicate x
    dp_shard ranks consume different ones (reference data_loader.py:1109-1145).
    """
    names = tuple(torch_device_mesh.mesh_dim_names or ())
    shared = 1
    for axis in ("tp", "cp", "sp"):
        if axis in names:
            shared *= torch_device_mesh[axis].size()
    rep = torch_device_mesh["dp_replicate"].size() if "dp_replicate" in names else 1
    shard = torch_device_mesh["dp_shard"].size() if "dp_shard" in names else 1
    rep_rank = torch_device_mesh.get_local_rank("dp_replicate") if "dp_replicate" in names else 0
    shard_rank = torch_device_mesh.get_local_rank("dp_shard") if "dp_shard" in names else 0
    if rep * shard * shared != torch_device_mesh.size():
        return process_index, num_processes  # unknown axes: do not guess
    return rep_rank * shard + shard_rank, rep * shard


def prepare_data_loader(dataloader: DataLoader, device=None, num_processes: Optional[int] = None, process_index: Optional[int] = None,
                        split_batches: bool = False, put_on_device: bool = False, rng_types=None, dispatch_batches: Optional[bool] = None,
                        even_batches: bool = True, slice_fn_for_dispatch: Optional[Callable] = None, use_seedable_sampler: bool = False,
                        data_seed: Optional[int] = None, non_blocking: bool = False, use_stateful_dataloader: bool = False,
                        torch_device_mesh=None) -> DataLoader:
    """Rebuild ``dataloader`` so each rank iterates its share and batches land on ``device``."""
    if dispatch_batches is None:
        if not put_on_device:
            dispatch_batches = False
        else:
            dispatch_batches = isinstance(dataloader.dataset, IterableDataset)
    if dispatch_batches and not put_on_device:
        raise ValueError("Using `dispatch_batches=True` requires `put_on_device=True`.")
    state = PartialState()
    if num_processes is None:
        num_processes = state.num_processes
    if process_index is None:
        process_index = state.process_index
    if torch_device_mesh is not None:
        process_index, num_processes = _mesh_data_ranks(torch_device_mesh, process_index, num_processes)

    if split_batches:
        if dataloader.batch_size is not None:
            batch_size_for_check = dataloader.batch_size
        elif hasattr(dataloader.batch_sampler, "batch_size"):
            batch_size_for_check = dataloader.batch_sampler.batch_size
        else:
            raise ValueError(
                "In order to use `split_batches==True` you must have a `batch_size` attribute either in the passed "
                "`dataloader` or `dataloader.batch_sampler` objects, and it has to return a natural number. "
                f"Your `dataloader.batch_size` is None and `dataloader.batch_sampler` (`{type(dataloader.batch_sampler)}`) does not have the `batch_size` attribute set."
            )
        if batch_size_for_check > 1 and batch_size_for_check % num_processes != 0:
            raise ValueError(
                f"To use a `DataLoader` in `split_batches` mode, the batch size ({dataloader.batch_size}) "
                f"needs to be a round multiple of the number of processes ({num_processes})."
            )

    new_dataset = dataloader.dataset
    new_batch_sampler = dataloader.batch_sampler if not isinstance(new_dataset, IterableDataset) else None
    sampler_is_batch_sampler = isinstance(dataloader.sampler, BatchSampler)
    synchronized_generator = None
    sampler = get_sampler(dataloader)
    if isinstance(sampler, RandomSampler) and use_seedable_sampler:
        sampler = SeedableRandomSampler(
            data_source=sampler.data_source, replacement=sampler.replacement, num_samples=sampler._num_samples,
            generator=getattr(sampler, "generator", torch.Generator(device=torch.get_default_device() if hasattr(torch, "get_default_device") else "cpu")),
            data_seed=data_seed,
        )
    if isinstance(dataloader.sampler, RandomSampler) and state.distributed_type == DistributedType.XLA:
        pass  # not reachable on this build

    if (num_processes != 1 or state.distributed_type == DistributedType.MEGATRON_LM) and not dispatch_batches:
        if isinstance(new_dataset, IterableDataset):
            if getattr(dataloader.dataset, "generator", None) is not None:
                synchronized_generator = dataloader.dataset.generator
            new_dataset = IterableDatasetShard(new_dataset, batch_size=dataloader.batch_size, drop_last=dataloader.drop_last,
                                               num_processes=num_processes, process_index=process_index, split_batches=split_batches)
        else:
            if not use_seedable_sampler and hasattr(sampler, "generator"):
                if sampler.generator is None:
                    sampler.generator = torch.Generator(device=torch.get_default_device() if hasattr(torch, "get_default_device") else "cpu")
                    seed = int(torch.empty((), dtype=torch.int64).random_().item())
                    sampler.generator.manual_seed(seed)
                synchronized_generator = sampler.generator
            batch_sampler = dataloader.sampler if sampler_is_batch_sampler else dataloader.batch_sampler
            new_batch_sampler = BatchSamplerShard(batch_sampler, num_processes=num_processes, process_index=process_index,
                                                  split_batches=split_batches, even_batches=even_batches)

    ignore_kwargs = ["batch_size", "shuffle", "sampler", "batch_sampler", "drop_last"]
    if rng_types is not None and synchronized_generator is None and "generator" in rng_types:
        rng_types = [r for r in rng_types if r != "generator"]
    kwargs = {k: getattr(dataloader, k, _PYTORCH_DATALOADER_KWARGS[k]) for k in _PYTORCH_DATALOADER_KWARGS if k not in ignore_kwargs}
    if new_batch_sampler is None:
        kwargs["drop_last"] = dataloader.drop_last
        kwargs["batch_size"] = dataloader.batch_size // num_processes if split_batches and not dispatch_batches else dataloader.batch_size

    if dispatch_batches:
        kwargs.pop("generator")
        dataloader = DataLoaderDispatcher(new_dataset, split_batches=split_batches, batch_sampler=new_batch_sampler,
                                          _drop_last=dataloader.drop_last, _non_blocking=non_blocking, slice_fn=slice_fn_for_dispatch,
                                          use_stateful_dataloader=use_stateful_dataloader, torch_device_mesh=torch_device_mesh, **kwargs)
    elif sampler_is_batch_sampler:
        dataloader = DataLoaderShard(new_dataset, device=device if put_on_device else None, sampler=new_batch_sampler,
                                     batch_size=dataloader.batch_size, rng_types=rng_types, _drop_last=dataloader.drop_last,
                                     _non_blocking=non_blocking, synchronized_generator=synchronized_generator,
                                     use_stateful_dataloader=use_stateful_dataloader, **kwargs)
    else:
        dataloader = DataLoaderShard(new_dataset, device=device if put_on_device else None, batch_sampler=new_batch_sampler,
                                     rng_types=rng_types, synchronized_generator=synchronized_generator, _drop_last=dataloader.drop_last,
                                     _non_blocking=non_blocking, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
    if isinstance(sampler, SeedableRandomSampler) and use_seedable_sampler:
        dataloader.set_sampler(sampler)
    return dataloader


class SkipBatchSampler(BatchSampler):
    """Batch sampler that drops the first ``skip_batches`` batches of another one."""

    def __init__(self, batch_sampler, skip_batches: int = 0):
        self.batch_sampler = batch_sampler
        self.skip_batches = skip_batches

    def __iter__(self):
        for index, samples in enumerate(self.batch_sampler):
            if index >= self.skip_batches:
                yield samples

    @property
    def total_length(self):
        return len(self.batch_sampler)

    def __len__(self):
        return len(self.batch_sampler) - self.skip_batches


class SkipDataLoader(DataLoaderAdapter, DataLoaderStateMixin):
    """Dataloader that drops its first ``skip_batches`` batches (for iterable datasets)."""

    def __init__(self, dataset, skip_batches: int = 0, use_stateful_dataloader: bool = False, **kwargs):
        super().__init__(dataset, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
        self.skip_batches = skip_batches
        self.gradient_state = GradientState()

    def __iter__(self):
        self.begin()
        for index, batch in enumerate(self.base_dataloader.__iter__()):
            if index >= self.skip_batches:
                self._update_state_dict()
                yield batch
        self.end()

    def __len__(self):
        return len(self.base_dataloader) - self.skip_batches

    def __reduce__(self):
        args = super().__reduce__()
        return (SkipDataLoader, *args[1:])


def skip_first_batches(dataloader, num_batches: int = 0):
    """A dataloader equal to ``dataloader`` minus its first ``num_batches`` batches (resume mid-epoch)."""
    dataset = dataloader.dataset
    sampler_is_batch_sampler = False
    if isinstance(dataset, IterableDataset):
        new_batch_sampler = None
    else:
        sampler_is_batch_sampler = isinstance(dataloader.sampler, BatchSampler)
        batch_sampler = dataloader.sampler if sampler_is_batch_sampler else dataloader.batch_sampler
        new_batch_sampler = SkipBatchSampler(batch_sampler, skip_batches=num_batches)
    ignore_kwargs = ["batch_size", "shuffle", "sampler", "batch_sampler", "drop_last"]
    kwargs = {k: getattr(dataloader, k, _PYTORCH_DATALOADER_KWARGS[k]) for k in _PYTORCH_DATALOADER_KWARGS if k not in ignore_kwargs}
    if new_batch_sampler is None:
        kwargs["drop_last"] = dataloader.drop_last
        kwargs["batch_size"] = dataloader.batch_size
    # the adapters fake ``__class__``; look at the real type
    is_dispatcher = issubclass(type(dataloader), DataLoaderDispatcher)
    is_shard = issubclass(type(dataloader), DataLoaderShard)
    if is_dispatcher:
        if new_batch_sampler is None:
            kwargs["skip_batches"] = num_batches
        kwargs.pop("generator", None)
        return DataLoaderDispatcher(dataset, split_batches=dataloader.split_batches, batch_sampler=new_batch_sampler,
                                    _drop_last=dataloader._drop_last, use_stateful_dataloader=dataloader.use_stateful_dataloader, **kwargs)
    if is_shard:
        if new_batch_sampler is None:
            kwargs["skip_batches"] = num_batches
        elif sampler_is_batch_sampler:
            kwargs["sampler"] = new_batch_sampler
            kwargs["batch_size"] = dataloader.batch_size
        else:
            kwargs["batch_sampler"] = new_batch_sampler
        return DataLoaderShard(dataset, device=dataloader.device, rng_types=dataloader.rng_types,
                               synchronized_generator=dataloader.synchronized_generator, _drop_last=dataloader._drop_last,
                               use_stateful_dataloader=dataloader.use_stateful_dataloader, **kwargs)
    if new_batch_sampler is None:
        return SkipDataLoader(dataset, skip_batches=num_batches, **kwargs)
    return DataLoader(dataset, batch_sampler=new_batch_sampler, **kwargs)
