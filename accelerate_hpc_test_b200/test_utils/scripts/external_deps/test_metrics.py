"""``gather_for_metrics`` must hand back exactly the evaluation set — no sample twice, none missing — whatever the sharding mode
(per-rank shards, rank-0 dispatch, split batches), for tensors and for arbitrary objects, for map-style and iterable datasets, and
also when ``drop_last`` removed the ragged tail.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_metrics
"""

import math

import torch
from torch.utils.data import DataLoader, IterableDataset, TensorDataset

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.test_utils.training import RegressionDataset, RegressionModel
from accelerate_hpc_test_b200.utils import DataLoaderConfiguration, set_seed


class ListHandler:
    """A dataset whose items are plain Python objects (strings here): gathered with ``use_gather_object=True``."""

    def __init__(self, items):
        self.items = items

    def __len__(self):
        return len(self.items)

    def __getitem__(self, index):
        return self.items[index]


class Stream(IterableDataset):
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __iter__(self):
        yield from (torch.tensor(float(i)) for i in range(self.n))


def fresh(**loader_options):
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    return Accelerator(dataloader_config=DataLoaderConfiguration(**loader_options))


def check_predictions_match_one_process(dispatch_batches, split_batches, samples=82, batch_size=16):
    """The distributed evaluation of a model gives, after gathering, the predictions of a single-process evaluation in order."""
    accelerator = fresh(dispatch_batches=dispatch_batches, split_batches=split_batches)
    set_seed(42)
    model = RegressionModel()
    data = RegressionDataset(length=samples, seed=42)
    per_loader = batch_size * (accelerator.num_processes if split_batches else 1)
    with torch.no_grad():
        expected = torch.cat([model(b["x"]) for b in DataLoader(data, batch_size=batch_size)])
    ddp_model, loader = accelerator.prepare(RegressionModel(), DataLoader(data, batch_size=per_loader))
    accelerator.unwrap_model(ddp_model).load_state_dict(model.state_dict())
    got, targets = [], []
    for batch in loader:
        with torch.no_grad():
            out = ddp_model(batch["x"])
        out, y = accelerator.gather_for_metrics((out, batch["y"]))
        got.append(out.cpu())
        targets.append(y.cpu())
    got, targets = torch.cat(got), torch.cat(targets)
    assert len(got) == samples, f"{len(got)} predictions for {samples} samples (dispatch={dispatch_batches}, split={split_batches})"
    assert torch.allclose(got, expected, atol=1e-6) and torch.allclose(targets, torch.as_tensor(data.y), atol=1e-6)


def check_objects(iterable: bool):
    accelerator = fresh()
    count = 30 + accelerator.num_processes + 1  # never a multiple of the global batch
    if iterable:
        loader = accelerator.prepare(DataLoader(Stream(count), batch_size=4))
        seen = []
        for batch in loader:
            seen.extend(accelerator.gather_for_metrics(batch).tolist())
        assert seen == [float(i) for i in range(count)], seen
    else:
        words = [f"sample-{i}" for i in range(count)]
        loader = accelerator.prepare(DataLoader(ListHandler(words), batch_size=4, collate_fn=list))
        seen = []
        for batch in loader:
            seen.extend(accelerator.gather_for_metrics(batch, use_gather_object=True))
        assert sorted(seen) == sorted(words) and len(seen) == count, (len(seen), count)


def check_drop_last():
    accelerator = fresh()
    per_rank = 4
    count = accelerator.num_processes * per_rank * 3 + 3  # three full global batches and a tail that gets dropped
    loader = accelerator.prepare(DataLoader(TensorDataset(torch.arange(count)), batch_size=per_rank, drop_last=True))
    sizes = [len(accelerator.gather_for_metrics(batch[0])) for batch in loader]
    assert sizes == [accelerator.num_processes * per_rank] * 3, sizes  # nothing is trimmed: there is no padding to remove


def check_accuracy_metric(samples=103):
    """A metric computed from gathered predictions equals the metric of the whole set computed in one place."""
    accelerator = fresh()
    g = torch.Generator().manual_seed(5)
    logits, labels = torch.randn(samples, 3, generator=g), torch.randint(0, 3, (samples,), generator=g)
    loader = accelerator.prepare(DataLoader(TensorDataset(logits, labels), batch_size=8))
    correct = total = 0
    for batch_logits, batch_labels in loader:
        predictions, references = accelerator.gather_for_metrics((batch_logits.argmax(-1), batch_labels))
        correct += int((predictions == references).sum())
        total += len(references)
    assert total == samples and math.isclose(correct / total, float((logits.argmax(-1) == labels).float().mean()), rel_tol=1e-6)


def main():
    for dispatch in (False, True):
        for split in (False, True):
            check_predictions_match_one_process(dispatch, split)
            AcceleratorState().print(f"predictions match the single-process run (dispatch_batches={dispatch}, split_batches={split})")
    check_objects(iterable=False)
    check_objects(iterable=True)
    AcceleratorState().print("objects and iterable datasets are gathered without duplicates")
    check_drop_last()
    check_accuracy_metric()
    AcceleratorState().print("test_metrics: all passed")


if __name__ == "__main__":
    main()
