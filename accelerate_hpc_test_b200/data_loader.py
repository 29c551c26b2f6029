"""Data-loading layer: what ``Accelerator.prepare`` turns a ``torch.utils.data.DataLoader`` into.

Three jobs:

* **who reads what** — ``BatchSamplerShard`` (map-style data) and ``IterableDatasetShard`` (streams) give each of ``P`` ranks its
  share of every *round* of ``P`` batches; a ragged last round is dropped, left ragged (``even_batches=False``) or completed with
  samples recycled from the start of the epoch (``_Refill``), so that every rank takes the same number of equally sized steps;
  ``DataLoaderDispatcher`` is the other policy: rank 0 reads, everybody receives a slice over the process group;
* **where the batch lives** — on a CUDA device ``DataLoaderShard`` issues the host→device copy of batch ``i+1`` from pinned memory on
  a copy stream while step ``i`` computes (``_DeviceStager``); the loader keeps one batch in hand anyway, because it has to know
  which batch is the last one (``end_of_dataloader`` / ``remainder`` drive ``gather_for_metrics`` and gradient accumulation);
* **reproducibility** — ``SeedableRandomSampler`` (order is a function of seed + epoch), RNG synchronisation at the start of an
  epoch, ``skip_first_batches`` for resuming in the middle of one, ``state_dict`` pass-through for torchdata's stateful loader.

Behavioural parity: reference ``src/accelerate/data_loader.py`` — sampler :73-107, batch-sampler shard :110-263, iterable shard
:266-362, state mixin :365-405, adapter :408-497, shard loader :500-648, dispatcher :704-959, ``prepare_data_loader`` :996-1309
(mesh arithmetic :1109-1145), skipping :1312-1451.  ``tests/test_data_loader.py`` pins the exact index lists per rank.
"""

from __future__ import annotations

import inspect
import math
import os
from typing import Callable, Iterable, Iterator, Optional

import torch
from torch.utils.data import BatchSampler, DataLoader, IterableDataset, RandomSampler

from .state import GradientState, PartialState
from .utils.dataclasses import DistributedType
from .utils.imports import is_torchdata_stateful_dataloader_available
from .utils.operations import (
    broadcast,
    broadcast_object_list,
    concatenate,
    find_batch_size,
    get_data_structure,
    initialize_tensors,
    recursively_apply,
    send_to_device,
    slice_tensors,
)
from .utils.random import synchronize_rng_states

# every keyword ``DataLoader`` accepts besides the dataset, with its default — read off the installed torch instead of a hand-kept table
_PYTORCH_DATALOADER_KWARGS = {
    name: (None if spec.default is inspect.Parameter.empty else spec.default)
    for name, spec in inspect.signature(DataLoader.__init__).parameters.items()
    if name not in ("self", "dataset")
}
# the ones a rebuilt loader decides for itself
_REBUILT = frozenset({"batch_size", "shuffle", "sampler", "batch_sampler", "drop_last"})


def _inherited_kwargs(dataloader) -> dict:
    """Construction arguments a rebuilt loader takes over from ``dataloader`` (workers, collate, pinning, ...)."""
    return {name: getattr(dataloader, name, default) for name, default in _PYTORCH_DATALOADER_KWARGS.items() if name not in _REBUILT}


def _cpu_generator() -> torch.Generator:
    where = torch.get_default_device() if hasattr(torch, "get_default_device") else "cpu"
    return torch.Generator(device=where)


# ------------------------------------------------------------------------------------------------- sampling
class SeedableRandomSampler(RandomSampler):
    """Random order that is a pure function of ``data_seed + epoch``: ranks agree, and a resumed run repeats it, without any
    RNG state being shipped around.  The epoch advances by itself after a full pass; ``set_epoch`` overrides it."""

    def __init__(self, *args, data_seed: Optional[int] = None, **kwargs):
        super().__init__(*args, **kwargs)
        self.initial_seed = torch.random.initial_seed() if data_seed is None else data_seed
        self.epoch = 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __iter__(self):
        if self.generator is None:
            self.generator = _cpu_generator()
        self.generator.manual_seed(self.initial_seed + self.epoch)
        yield from super().__iter__()
        self.epoch += 1


class _Refill:
    """Samples remembered from the start of an epoch, handed out again (cyclically) to complete a ragged last round."""

    def __init__(self, capacity: int):
        self.pool: list = []
        self.capacity = capacity  # how many more samples to remember
        self.cursor = 0

    def remember(self, samples: Iterable) -> None:
        if self.capacity > 0:
            kept = list(samples)[: self.capacity]
            self.pool.extend(kept)
            self.capacity -= len(kept)

    def take(self, count: int) -> list:
        out = []
        while len(out) < count:
            out.append(self.pool[self.cursor % len(self.pool)])
            self.cursor += 1
        return out


class BatchSamplerShard(BatchSampler):
    """What rank ``process_index`` of ``num_processes`` sees of ``batch_sampler``.

    Dealing mode (default): batches go round the ranks — rank r gets batches r, r+P, r+2P, ...  Slicing mode
    (``split_batches=True``): every batch is cut into P equal slices, rank r keeps slice r.  When the data does not fill the last
    round: ``drop_last`` of the wrapped sampler drops it; otherwise ``even_batches=True`` completes it with samples from the start
    of the epoch, and ``even_batches=False`` leaves the ranks with what there is (some get a shorter batch or none).
    """

    def __init__(self, batch_sampler: BatchSampler, num_processes: int = 1, process_index: int = 0, split_batches: bool = False,
                 even_batches: bool = True):
        self.batch_sampler = batch_sampler
        self.num_processes, self.process_index = num_processes, process_index
        self.split_batches, self.even_batches = split_batches, even_batches
        self.batch_size = getattr(batch_sampler, "batch_size", None)
        self.drop_last = getattr(batch_sampler, "drop_last", False)
        if split_batches and self.batch_size % num_processes:
            raise ValueError(
                f"To use `BatchSamplerShard` in `split_batches` mode, the batch size ({self.batch_size}) needs to be a round "
                f"multiple of the number of processes ({num_processes})."
            )
        if even_batches and self.batch_size is None:
            raise ValueError(
                "You need to use `even_batches=False` when the batch sampler has no batch size. If you are not calling this method "
                "directly, set `accelerator.even_batches=False` instead."
            )

    @property
    def total_length(self) -> int:
        return len(self.batch_sampler)

    def __len__(self) -> int:
        batches = len(self.batch_sampler)
        if self.split_batches:
            return batches
        rounds, extra = divmod(batches, self.num_processes)
        if extra == 0 or self.drop_last:
            return rounds
        if self.even_batches or self.process_index < extra:
            return rounds + 1
        return rounds

    def __iter__(self):
        return self._sliced() if self.split_batches else self._dealt()

    def _short(self, batch) -> bool:
        return self.batch_size is not None and len(batch) < self.batch_size

    def _sliced(self):
        width = self.batch_size // self.num_processes
        mine = slice(width * self.process_index, width * (self.process_index + 1))
        refill = _Refill(self.batch_size)
        tail = None
        for batch in self.batch_sampler:
            refill.remember(batch)
            if self._short(batch):
                tail = list(batch)  # only the last batch can be short
            else:
                yield batch[mine]
        if tail is None or self.drop_last:
            return
        if self.even_batches:
            yield (tail + refill.take(self.batch_size - len(tail)))[mine]
        elif len(tail) > mine.start:
            yield tail[mine]

    def _dealt(self):
        ranks, me = self.num_processes, self.process_index
        refill = _Refill(0 if self.drop_last or self.batch_size is None else ranks * self.batch_size)
        hand: list = []  # the round being dealt
        for batch in self.batch_sampler:
            refill.remember(batch)
            hand.append(batch)
            if len(hand) == ranks and not self._short(batch):
                yield hand[me]
                hand = []
        if not hand or self.drop_last:
            return
        if not self.even_batches:
            if me < len(hand) and len(hand[me]) > 0:
                yield hand[me]
            return
        # top the short batch up, then mint the batches the round still lacks, all from the start of the epoch
        hand = [list(batch) for batch in hand]
        hand[-1] += refill.take(self.batch_size - len(hand[-1]))
        while len(hand) < ranks:
            hand.append(refill.take(self.batch_size))
        yield hand[me]


class IterableDatasetShard(IterableDataset):
    """Rank ``process_index``'s view of a stream: buffer one *global* batch (``batch_size`` x ranks, or just ``batch_size`` when
    ``split_batches``), emit this rank's contiguous part of it.  A partial last global batch is dropped or completed from the
    first one."""

    def __init__(self, dataset: IterableDataset, batch_size: int = 1, drop_last: bool = False, num_processes: int = 1,
                 process_index: int = 0, split_batches: bool = False):
        if split_batches and batch_size > 1 and batch_size % num_processes:
            raise ValueError(
                f"To use `IterableDatasetShard` in `split_batches` mode, the batch size ({batch_size}) needs to be a round "
                f"multiple of the number of processes ({num_processes})."
            )
        self.dataset = dataset
        self.batch_size, self.drop_last = batch_size, drop_last
        self.num_processes, self.process_index = num_processes, process_index
        self.split_batches = split_batches

    def _geometry(self) -> tuple[int, int]:
        """(samples per global batch, samples of it this rank keeps)"""
        if self.split_batches:
            return self.batch_size, self.batch_size // self.num_processes
        return self.batch_size * self.num_processes, self.batch_size

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch
        if hasattr(self.dataset, "set_epoch"):
            self.dataset.set_epoch(epoch)

    def __len__(self) -> int:
        whole, part = self._geometry()
        count = len(self.dataset) / whole
        return (math.floor(count) if self.drop_last else math.ceil(count)) * part

    def __iter__(self):
        own_generator = getattr(self.dataset, "generator", None)
        if isinstance(own_generator, torch.Generator) and not hasattr(self.dataset, "set_epoch"):
            own_generator.manual_seed(getattr(self, "epoch", 0))
        whole, part = self._geometry()
        lo = self.process_index * part
        refill = _Refill(whole)
        pending: list = []
        for sample in self.dataset:
            pending.append(sample)
            if len(pending) == whole:
                refill.remember(pending)
                yield from pending[lo : lo + part]
                pending = []
        if pending and not self.drop_last:
            refill.remember(pending)
            pending += refill.take(whole - len(pending))
            yield from pending[lo : lo + part]


# ------------------------------------------------------------------------------------------------- sampler plumbing
def _index_batcher(dataloader):
    """The object that produces lists of indices: with ``DataLoader(sampler=<a BatchSampler>, batch_size=None)`` that is the sampler."""
    return dataloader.sampler if isinstance(dataloader.sampler, BatchSampler) else dataloader.batch_sampler


def _batcher_layers(dataloader) -> Iterator:
    """The batch sampler and whatever it wraps (shard / skip wrappers expose the wrapped one as ``.batch_sampler``), outside in."""
    layer, seen = _index_batcher(dataloader), set()
    while layer is not None and id(layer) not in seen:
        seen.add(id(layer))
        yield layer
        layer = getattr(layer, "batch_sampler", None)


def get_sampler(dataloader):
    """The index sampler under the batch-sampler wrappers of ``dataloader`` (None for iterable datasets)."""
    for layer in _batcher_layers(dataloader):
        found = getattr(layer, "sampler", None)
        if found is not None:
            return found
    return None


def _install_sampler(dataloader, sampler) -> None:
    layers = list(_batcher_layers(dataloader))
    for depth, layer in enumerate(layers):
        if depth == 0 or hasattr(layer, "sampler"):
            layer.sampler = sampler


def _announce_epoch(dataloader, epoch: int) -> None:
    """Tell everything that shuffles by epoch: batch samplers, their index samplers, else the dataset."""
    told = set()
    for layer in _batcher_layers(dataloader):
        for target in (layer, getattr(layer, "sampler", None)):
            if target is not None and id(target) not in told and hasattr(target, "set_epoch"):
                told.add(id(target))
                target.set_epoch(epoch)
    if not told and hasattr(dataloader.dataset, "set_epoch"):
        dataloader.dataset.set_epoch(epoch)


# ------------------------------------------------------------------------------------------------- loader building blocks
class DataLoaderStateMixin:
    """``end_of_dataloader`` (true while the last batch is being consumed) and ``remainder`` (how many real samples the last
    global batch holds; -1 when unknown), plus registration with ``GradientState`` for the duration of an iteration."""

    def __init_subclass__(cls, **kwargs):
        cls.end_of_dataloader = False
        cls.remainder = -1

    def reset(self) -> None:
        self.end_of_dataloader, self.remainder = False, -1

    def begin(self) -> None:
        self.reset()
        if not self._drop_last:
            try:
                samples = getattr(self.dataset, "total_dataset_length", None)
                self.remainder = (len(self.dataset) if samples is None else samples) % self.total_batch_size
            except Exception:  # streams without a length, batch samplers without a batch size
                pass
        self.gradient_state._add_dataloader(self)

    def end(self) -> None:
        self.gradient_state._remove_dataloader(self)


class DataLoaderAdapter:
    """Owns the real loader (``torch.utils.data.DataLoader`` or torchdata's ``StatefulDataLoader``) and forwards to it.  It also
    reports that loader's class as its own, so ``isinstance(prepared, DataLoader)`` keeps holding in user code."""

    def __init__(self, dataset, use_stateful_dataloader: bool = False, batch_sampler=None, **kwargs):
        self.use_stateful_dataloader = use_stateful_dataloader
        loader_cls = DataLoader
        if use_stateful_dataloader:
            if not is_torchdata_stateful_dataloader_available():
                raise ImportError("StatefulDataLoader is not available. Please install torchdata version 0.8.0 or higher to use it.")
            from torchdata.stateful_dataloader import StatefulDataLoader as loader_cls
        self.base_dataloader = loader_cls(dataset, batch_sampler=batch_sampler, **kwargs)
        self._snapshot(adjust=False)

    __class__ = property(lambda self: self.base_dataloader.__class__)

    def __getattr__(self, name):
        if name == "base_dataloader":  # not constructed yet (unpickling): do not recurse
            raise AttributeError(name)
        return getattr(self.base_dataloader, name)

    def __len__(self) -> int:
        return len(self.base_dataloader)

    def state_dict(self):
        return self.dl_state_dict

    def load_state_dict(self, state_dict) -> None:
        self.base_dataloader.load_state_dict(state_dict)

    def adjust_state_dict_for_prefetch(self) -> None:
        """The wrapped loader runs ahead of what has been handed out (one batch in hand per rank): take that back out of the counts."""
        if PartialState().distributed_type == DistributedType.NO:
            return
        ahead = PartialState().num_processes - 1
        state = self.dl_state_dict
        for counter in ("_sampler_iter_yielded", "_num_yielded"):
            if state[counter] > 0:
                state[counter] -= ahead
        inner = state["_index_sampler_state"]
        if inner is not None and inner.get("samples_yielded", 0) > 0:
            inner["samples_yielded"] -= self.batch_size * ahead

    def _snapshot(self, adjust: bool = True) -> None:
        if not hasattr(self.base_dataloader, "state_dict"):
            return
        self.dl_state_dict = self.base_dataloader.state_dict()
        if adjust:
            self.adjust_state_dict_for_prefetch()
            self.dl_state_dict["_iterator_finished"] = self.end_of_dataloader

    _update_state_dict = _snapshot  # reference-compatible name


def _flag_last(iterator: Iterator, before_fetch: Optional[Callable[[], None]] = None, on_fetch: Callable = lambda item: item):
    """``(item, is_last)`` for every item of ``iterator``, which takes reading one item ahead.  ``before_fetch`` runs before every
    read-ahead, ``on_fetch`` maps every item the moment it has been read (that is where the device copy is started)."""
    nothing = object()
    held = next(iterator, nothing)
    if held is nothing:
        return
    held = on_fetch(held)
    while True:
        if before_fetch is not None:
            before_fetch()
        ahead = next(iterator, nothing)
        if ahead is nothing:
            yield held, True
            return
        ahead = on_fetch(ahead)
        yield held, False
        held = ahead


class _DeviceStager:
    """Host→device copies one batch ahead of the consumer.

    ``stage`` pins what is not pinned yet and enqueues the copies on a private stream; ``claim`` makes the consumer's stream wait
    for exactly those copies (an event, no host synchronisation) and hands the tensors over to it (``record_stream``, so the
    caching allocator does not recycle them under a kernel still reading).  Off CUDA, or when the loader was not asked for
    non-blocking copies, ``stage`` is a plain synchronous ``send_to_device`` and ``claim`` does nothing.
    """

    def __init__(self, device, non_blocking: bool):
        self.device = None if device is None else torch.device(device)
        self.non_blocking = non_blocking
        self.overlapped = (
            non_blocking and self.device is not None and self.device.type == "cuda" and torch.cuda.is_available()
            and os.environ.get("ACCELERATE_B200_ASYNC_H2D", "1") != "0"
        )
        self.stream = torch.cuda.Stream(device=self.device) if self.overlapped else None

    @staticmethod
    def _pinned(t: torch.Tensor) -> torch.Tensor:
        return t if t.device.type != "cpu" or t.is_pinned() else t.pin_memory()

    def stage(self, batch):
        if self.device is None:
            return batch, None
        if not self.overlapped:
            return send_to_device(batch, self.device, non_blocking=self.non_blocking), None
        batch = recursively_apply(self._pinned, batch)
        with torch.cuda.stream(self.stream):
            moved = send_to_device(batch, self.device, non_blocking=True)
            landed = torch.cuda.Event()
            landed.record(self.stream)
        return moved, landed

    def claim(self, staged):
        moved, landed = staged
        if landed is not None:
            consumer = torch.cuda.current_stream(self.device)
            consumer.wait_event(landed)
            recursively_apply(lambda t: (t.record_stream(consumer), t)[1] if t.is_cuda else t, moved)
        return moved


class DataLoaderShard(DataLoaderAdapter, DataLoaderStateMixin):
    """The loader every rank iterates on its own share of the data.  At the start of a pass it synchronises the requested RNGs
    and announces the epoch; batches arrive on ``device``; the last batch is flagged while the loop body still runs on it."""

    def __init__(self, dataset, device=None, rng_types=None, synchronized_generator=None, skip_batches: int = 0,
                 use_stateful_dataloader: bool = False, _drop_last: bool = False, _non_blocking: bool = False,
                 torch_device_mesh=None, **kwargs):
        super().__init__(dataset, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
        self.device, self.rng_types, self.synchronized_generator = device, rng_types, synchronized_generator
        self.skip_batches = skip_batches
        self._drop_last, self._non_blocking = _drop_last, _non_blocking
        self.gradient_state = GradientState()
        self.iteration = 0

    def __iter__(self):
        if self.rng_types is not None:
            synchronize_rng_states(self.rng_types, self.synchronized_generator)
        self.begin()
        self.set_epoch(self.iteration)
        stager = _DeviceStager(self.device, self._non_blocking)
        stream = _flag_last(iter(self.base_dataloader), before_fetch=self._snapshot, on_fetch=stager.stage)
        for position, (staged, is_last) in enumerate(stream):
            if is_last:
                self.end_of_dataloader = True
                self._snapshot()
            if position >= self.skip_batches:
                yield stager.claim(staged)
        self.iteration += 1
        self.end()

    def __reduce__(self):
        return (DataLoaderShard, *super().__reduce__()[1:])

    def set_epoch(self, epoch: int) -> None:
        self.iteration = epoch
        _announce_epoch(self, epoch)

    @property
    def total_batch_size(self) -> int:
        batcher = _index_batcher(self)
        per_rank_factor = 1 if getattr(batcher, "split_batches", False) else getattr(batcher, "num_processes", 1)
        return batcher.batch_size * per_rank_factor

    @property
    def total_dataset_length(self) -> int:
        return getattr(self.dataset, "total_length", None) or len(self.dataset)

    def get_sampler(self):
        return get_sampler(self)

    def set_sampler(self, sampler) -> None:
        _install_sampler(self, sampler)


class DataLoaderDispatcher(DataLoaderAdapter, DataLoaderStateMixin):
    """Rank 0 is the only reader.  For every global batch it tells the group what is coming (the batch's structure — shapes and
    dtypes — and whether it is the last), then the tensors are broadcast and every rank keeps its slice.  Meant for streams that
    cannot be sharded; costs one object broadcast and one tensor broadcast per step."""

    def __init__(self, dataset, split_batches: bool = False, skip_batches: int = 0, use_stateful_dataloader: bool = False,
                 _drop_last: bool = False, _non_blocking: bool = False, slice_fn=None, torch_device_mesh=None, **kwargs):
        super().__init__(dataset, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
        self.split_batches, self.skip_batches = split_batches, skip_batches
        self._drop_last, self._non_blocking = _drop_last, _non_blocking
        self.gradient_state, self.state = GradientState(), PartialState()
        self.slice_fn = slice_fn or slice_tensors
        self.torch_device_mesh = torch_device_mesh
        self.iteration = 0
        axes = () if torch_device_mesh is None else tuple(torch_device_mesh.mesh_dim_names or ())
        self.submesh_tp, self.submesh_dp, self.submesh_fsdp = (torch_device_mesh[a] if a in axes else None for a in ("tp", "dp", "fsdp"))
        if self.submesh_tp is not None:
            # ranks of one tp group would need the same slice: that routing does not exist here (the sharded loader has it)
            raise ValueError("TP + (DP/FSDP) is not yet supported in dispatch mode")

    # -- rank 0 ---------------------------------------------------------------------------------------------
    def _global_batches(self):
        """Global batches as rank 0 reads them: one loader batch (``split_batches``) or ``num_processes`` of them concatenated; a
        last incomplete group is kept unless the loader drops its tail."""
        source = iter(self.base_dataloader)
        group_size = 1 if self.split_batches else self.state.num_processes
        while True:
            group = []
            for _ in range(group_size):
                self._snapshot()
                try:
                    group.append(next(source))
                except StopIteration:
                    break
            exhausted = len(group) < group_size
            if group and not (exhausted and self._drop_last):
                try:
                    yield group[0] if self.split_batches else concatenate(group, dim=0)
                except RuntimeError as err:
                    raise RuntimeError(
                        "You can't use batches of different size with `dispatch_batches=True` or when using an `IterableDataset`."
                        "either pass `dispatch_batches=False` and have each process fetch its own batch  or pass `split_batches=True`. "
                        "By doing so, the main process will fetch a full batch and slice it into `num_processes` batches for each process."
                    ) from err
            if exhausted:
                return

    def _headers(self):
        """``(batch or None, structure, is_last)`` on every rank, in lock step.  A closing ``structure=None`` header ends an empty pass."""
        if self.state.process_index == 0:
            sent = False
            for batch, is_last in _flag_last(self._global_batches()):
                broadcast_object_list([get_data_structure(batch), is_last])
                sent = True
                yield batch, None, is_last
            if not sent:
                broadcast_object_list([None, True])
            return
        while True:
            header = [None, True]
            broadcast_object_list(header)
            if header[0] is None:
                return
            yield None, header[0], header[1]
            if header[1]:
                return

    # -- every rank -----------------------------------------------------------------------------------------
    def __iter__(self):
        self.begin()
        self.set_epoch(self.iteration)
        ranks, me = self.state.num_processes, self.state.process_index
        opening = None  # first `ranks` samples of the epoch, to complete a last global batch that does not divide
        for position, (batch, structure, is_last) in enumerate(self._headers()):
            if batch is None:
                batch = initialize_tensors(structure)
            batch = broadcast(send_to_device(batch, self.state.device, non_blocking=self._non_blocking), from_process=0)
            if opening is None and not self._drop_last:
                opening = self.slice_fn(batch, slice(0, ranks), process_index=me, num_processes=ranks)
            observed = find_batch_size(batch)
            share = observed // ranks
            if is_last and observed % ranks and not self._drop_last:
                batch = concatenate([batch, opening], dim=0)
                share += 1
            batch = self.slice_fn(batch, slice(me * share, (me + 1) * share), process_index=me, num_processes=ranks)
            if is_last:
                self.end_of_dataloader = True
                self._snapshot()
                self.remainder = observed
            if position >= self.skip_batches:
                yield batch
        self.iteration += 1
        self.end()

    def set_epoch(self, epoch: int) -> None:
        self.iteration = epoch
        _announce_epoch(self, epoch)

    def __len__(self) -> int:
        batches = len(self.base_dataloader)
        if self.split_batches:
            return batches
        per_rank = batches / self.state.num_processes
        return math.floor(per_rank) if self._drop_last else math.ceil(per_rank)

    def __reduce__(self):
        return (DataLoaderDispatcher, *super().__reduce__()[1:])

    @property
    def total_batch_size(self) -> int:
        inner = self.dataset
        return inner.batch_size if self.split_batches else inner.batch_size * inner.num_processes

    @property
    def total_dataset_length(self) -> int:
        return len(self.dataset)

    def get_sampler(self):
        return get_sampler(self)

    def set_sampler(self, sampler) -> None:
        _install_sampler(self, sampler)


# ------------------------------------------------------------------------------------------------- prepare
def _mesh_data_ranks(torch_device_mesh, process_index: int, num_processes: int) -> tuple[int, int]:
    """(index, count) of this rank among the ranks that must read *different* data.

    Ranks that differ only in their tp / cp / sp coordinate work on the same batch; the dp_replicate x dp_shard grid reads
    different ones, replicate-major (reference data_loader.py:1109-1145).  A mesh with axes this does not know is left alone.
    """
    axes = tuple(torch_device_mesh.mesh_dim_names or ())

    def extent(axis):
        return torch_device_mesh[axis].size() if axis in axes else 1

    def coordinate(axis):
        return torch_device_mesh.get_local_rank(axis) if axis in axes else 0

    same_data = extent("tp") * extent("cp") * extent("sp")
    readers = extent("dp_replicate") * extent("dp_shard")
    if same_data * readers != torch_device_mesh.size():
        return process_index, num_processes
    return coordinate("dp_replicate") * extent("dp_shard") + coordinate("dp_shard"), readers


def _check_splittable(dataloader, num_processes: int) -> None:
    size = dataloader.batch_size
    if size is None:
        size = getattr(dataloader.batch_sampler, "batch_size", None)
    if size is None:
        raise ValueError(
            "In order to use `split_batches==True` you must have a `batch_size` attribute either in the passed `dataloader` or "
            "`dataloader.batch_sampler` objects, and it has to return a natural number. Your `dataloader.batch_size` is None and "
            f"`dataloader.batch_sampler` (`{type(dataloader.batch_sampler)}`) does not have the `batch_size` attribute set."
        )
    if size > 1 and size % num_processes:
        raise ValueError(
            f"To use a `DataLoader` in `split_batches` mode, the batch size ({dataloader.batch_size}) needs to be a round multiple "
            f"of the number of processes ({num_processes})."
        )


def _as_seedable(sampler: RandomSampler, data_seed) -> SeedableRandomSampler:
    generator = getattr(sampler, "generator", None) or _cpu_generator()
    return SeedableRandomSampler(data_source=sampler.data_source, replacement=sampler.replacement, num_samples=sampler._num_samples,
                                 generator=generator, data_seed=data_seed)


def _generator_to_synchronise(sampler):
    """The generator whose state rank 0 will broadcast at every epoch start, created (randomly seeded) if the sampler has none yet."""
    if not hasattr(sampler, "generator"):
        return None
    if sampler.generator is None:
        sampler.generator = _cpu_generator()
        sampler.generator.manual_seed(int(torch.empty((), dtype=torch.int64).random_().item()))
    return sampler.generator


def prepare_data_loader(dataloader: DataLoader, device=None, num_processes: Optional[int] = None, process_index: Optional[int] = None,
                        split_batches: bool = False, put_on_device: bool = False, rng_types=None, dispatch_batches: Optional[bool] = None,
                        even_batches: bool = True, slice_fn_for_dispatch: Optional[Callable] = None, use_seedable_sampler: bool = False,
                        data_seed: Optional[int] = None, non_blocking: bool = False, use_stateful_dataloader: bool = False,
                        torch_device_mesh=None) -> DataLoader:
    """Rebuild ``dataloader`` for this rank: its share of the data (or rank-0 dispatch), batches placed on ``device``.

    ``dispatch_batches=None`` means "dispatch exactly when the dataset is a stream and batches go to the device".  ``split_batches``
    keeps the global batch size equal to the loader's batch size (each rank gets a slice) instead of multiplying it by the number of
    ranks.  With a device mesh only the data-parallel axes count as readers.
    """
    streaming = isinstance(dataloader.dataset, IterableDataset)
    if dispatch_batches is None:
        dispatch_batches = put_on_device and streaming
    elif dispatch_batches and not put_on_device:
        raise ValueError("Using `dispatch_batches=True` requires `put_on_device=True`.")
    state = PartialState()
    num_processes = state.num_processes if num_processes is None else num_processes
    process_index = state.process_index if process_index is None else process_index
    if torch_device_mesh is not None:
        process_index, num_processes = _mesh_data_ranks(torch_device_mesh, process_index, num_processes)
    if split_batches:
        _check_splittable(dataloader, num_processes)

    sampler = get_sampler(dataloader)
    if use_seedable_sampler and isinstance(sampler, RandomSampler):
        sampler = _as_seedable(sampler, data_seed)

    dataset = dataloader.dataset
    batcher = None if streaming else dataloader.batch_sampler
    batcher_is_sampler = isinstance(dataloader.sampler, BatchSampler)
    synchronized_generator = None
    shard_here = not dispatch_batches and (num_processes != 1 or state.distributed_type == DistributedType.MEGATRON_LM)
    if shard_here and streaming:
        synchronized_generator = getattr(dataset, "generator", None)
        dataset = IterableDatasetShard(dataset, batch_size=dataloader.batch_size, drop_last=dataloader.drop_last,
                                       num_processes=num_processes, process_index=process_index, split_batches=split_batches)
    elif shard_here:
        if not use_seedable_sampler:
            synchronized_generator = _generator_to_synchronise(sampler)
        batcher = BatchSamplerShard(_index_batcher(dataloader), num_processes=num_processes, process_index=process_index,
                                    split_batches=split_batches, even_batches=even_batches)

    if rng_types is not None and synchronized_generator is None:
        rng_types = [kind for kind in rng_types if kind != "generator"]
    kwargs = _inherited_kwargs(dataloader)
    if batcher is None:  # streams: the loader itself batches
        kwargs["drop_last"] = dataloader.drop_last
        kwargs["batch_size"] = dataloader.batch_size // num_processes if split_batches and not dispatch_batches else dataloader.batch_size
    common = dict(_drop_last=dataloader.drop_last, _non_blocking=non_blocking, use_stateful_dataloader=use_stateful_dataloader)

    if dispatch_batches:
        kwargs.pop("generator", None)
        prepared = DataLoaderDispatcher(dataset, split_batches=split_batches, batch_sampler=batcher, slice_fn=slice_fn_for_dispatch,
                                        torch_device_mesh=torch_device_mesh, **common, **kwargs)
    else:
        if batcher_is_sampler:  # DataLoader(sampler=<BatchSampler>, batch_size=None): keep that calling convention
            placement = {"sampler": batcher, "batch_size": dataloader.batch_size}
        else:
            placement = {"batch_sampler": batcher}
        prepared = DataLoaderShard(dataset, device=device if put_on_device else None, rng_types=rng_types,
                                   synchronized_generator=synchronized_generator, **placement, **common, **kwargs)
    if isinstance(sampler, SeedableRandomSampler):
        prepared.set_sampler(sampler)
    return prepared


# ------------------------------------------------------------------------------------------------- resuming inside an epoch
class SkipBatchSampler(BatchSampler):
    """``batch_sampler`` without its first ``skip_batches`` batches."""

    def __init__(self, batch_sampler, skip_batches: int = 0):
        self.batch_sampler, self.skip_batches = batch_sampler, skip_batches

    @property
    def total_length(self) -> int:
        return len(self.batch_sampler)

    def __len__(self) -> int:
        return len(self.batch_sampler) - self.skip_batches

    def __iter__(self):
        produced = iter(self.batch_sampler)
        for _ in range(self.skip_batches):
            if next(produced, None) is None:
                return
        yield from produced


class SkipDataLoader(DataLoaderAdapter, DataLoaderStateMixin):
    """A loader that reads and discards its first ``skip_batches`` batches — the only way to skip in a stream."""

    def __init__(self, dataset, skip_batches: int = 0, use_stateful_dataloader: bool = False, **kwargs):
        super().__init__(dataset, use_stateful_dataloader=use_stateful_dataloader, **kwargs)
        self.skip_batches = skip_batches
        self.gradient_state = GradientState()
        self._drop_last = kwargs.get("drop_last", False)

    def __len__(self) -> int:
        return len(self.base_dataloader) - self.skip_batches

    def __iter__(self):
        self.begin()
        for position, batch in enumerate(iter(self.base_dataloader)):
            if position < self.skip_batches:
                continue
            self._snapshot()
            yield batch
        self.end()

    def __reduce__(self):
        return (SkipDataLoader, *super().__reduce__()[1:])


def skip_first_batches(dataloader, num_batches: int = 0):
    """A loader that continues where ``dataloader`` would be after ``num_batches`` batches, of the same kind (sharded, dispatching
    or plain).  Map-style data skips at the index level (nothing is read); streams have to be read through."""
    dataset = dataloader.dataset
    streaming = isinstance(dataset, IterableDataset)
    kwargs = _inherited_kwargs(dataloader)
    if streaming:
        kwargs.update(drop_last=dataloader.drop_last, batch_size=dataloader.batch_size)
        placement = {}
    else:
        skipping = SkipBatchSampler(_index_batcher(dataloader), skip_batches=num_batches)
        if isinstance(dataloader.sampler, BatchSampler):
            placement = {"sampler": skipping, "batch_size": dataloader.batch_size}
        else:
            placement = {"batch_sampler": skipping}
    kind = type(dataloader)  # the adapters report a borrowed __class__; type() sees the real one
    if issubclass(kind, DataLoaderDispatcher):
        kwargs.pop("generator", None)
        if streaming:
            kwargs["skip_batches"] = num_batches
        else:
            placement = {"batch_sampler": skipping}
        return DataLoaderDispatcher(dataset, split_batches=dataloader.split_batches, _drop_last=dataloader._drop_last,
                                    use_stateful_dataloader=dataloader.use_stateful_dataloader, **placement, **kwargs)
    if issubclass(kind, DataLoaderShard):
        if streaming:
            kwargs["skip_batches"] = num_batches
        return DataLoaderShard(dataset, device=dataloader.device, rng_types=dataloader.rng_types,
                               synchronized_generator=dataloader.synchronized_generator, _drop_last=dataloader._drop_last,
                               _non_blocking=dataloader._non_blocking, use_stateful_dataloader=dataloader.use_stateful_dataloader,
                               **placement, **kwargs)
    if streaming:
        return SkipDataLoader(dataset, skip_batches=num_batches, **kwargs)
    return DataLoader(dataset, batch_sampler=skipping, **kwargs)
