"""k-fold cross validation: train one model per fold, evaluate each on a held-out test set, and ensemble the folds by averaging
their logits — the distributed part is that every rank holds a slice of the test predictions, so they are gathered
(``gather_for_metrics``) before they are averaged, and the accelerator is reset between folds (``free_memory``)
(reference examples/by_feature/cross_validation.py).

    python examples/by_feature/cross_validation.py --cpu --model tiny --num_folds 3 --num_epochs 1
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch
from torch.optim import AdamW
from torch.utils.data import DataLoader, Subset

from _common import base_parser, default_config, linear_schedule_with_warmup

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification, pad_collate, synthetic_mrpc
from accelerate_hpc_test_b200.utils import set_seed


# New Code #
def fold_indices(num_samples: int, num_folds: int, seed: int):
    """``[(train indices, validation indices)]``: a seeded shuffle cut into ``num_folds`` validation parts."""
    order = torch.randperm(num_samples, generator=torch.Generator().manual_seed(seed)).tolist()
    bounds = [round(i * num_samples / num_folds) for i in range(num_folds + 1)]
    folds = []
    for i in range(num_folds):
        held_out = order[bounds[i] : bounds[i + 1]]
        folds.append((order[: bounds[i]] + order[bounds[i + 1] :], held_out))
    return folds


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    lr, num_epochs, seed, batch_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"])
    cfg = BertConfig.tiny() if args.model == "tiny" else BertConfig.base_cased()
    with accelerator.main_process_first():
        pool = synthetic_mrpc(args.train_size, vocab_size=cfg.vocab_size, seed=0)
        test = synthetic_mrpc(args.eval_size, vocab_size=cfg.vocab_size, seed=1)
    test_references = None
    # New Code #
    fold_logits = []
    for fold, (train_idx, valid_idx) in enumerate(fold_indices(len(pool), int(args.num_folds), seed)):
        set_seed(seed)
        train_dataloader = DataLoader(Subset(pool, train_idx), shuffle=True, collate_fn=pad_collate, batch_size=batch_size, drop_last=True)
        valid_dataloader = DataLoader(Subset(pool, valid_idx), shuffle=False, collate_fn=pad_collate, batch_size=32)
        test_dataloader = DataLoader(test, shuffle=False, collate_fn=pad_collate, batch_size=32)
        model = BertForSequenceClassification(cfg).to(accelerator.device)
        optimizer = AdamW(params=model.parameters(), lr=lr)
        lr_scheduler = linear_schedule_with_warmup(optimizer, 2, len(train_dataloader) * num_epochs)
        model, optimizer, train_dataloader, valid_dataloader, test_dataloader, lr_scheduler = accelerator.prepare(
            model, optimizer, train_dataloader, valid_dataloader, test_dataloader, lr_scheduler
        )
        for epoch in range(num_epochs):
            model.train()
            for batch in train_dataloader:
                accelerator.backward(model(**batch).loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
            model.eval()
            correct = total = 0
            for batch in valid_dataloader:
                with torch.no_grad():
                    predictions = model(**batch).logits.argmax(dim=-1)
                predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
                correct += int((predictions == references).sum())
                total += int(references.numel())
            accelerator.print(f"fold {fold} epoch {epoch}:", {"accuracy": correct / max(1, total), "eval_samples": total})
        # New Code #
        # this fold's logits on the whole test set, in dataset order on every rank
        logits, references = [], []
        for batch in test_dataloader:
            with torch.no_grad():
                out = model(**batch).logits
            out, refs = accelerator.gather_for_metrics((out, batch["labels"]))
            logits.append(out.float().cpu())
            references.append(refs.cpu())
        fold_logits.append(torch.cat(logits))
        test_references = torch.cat(references)
        # New Code #
        # drop the references the accelerator keeps to this fold's objects before the next fold is prepared
        model, optimizer, lr_scheduler = accelerator.free_memory(model, optimizer, lr_scheduler)
    # New Code #
    ensemble = torch.stack(fold_logits).mean(dim=0).argmax(dim=-1)
    metrics = {"accuracy": float((ensemble == test_references).float().mean()), "eval_samples": int(test_references.numel())}
    accelerator.print(f"average of {len(fold_logits)} folds on the test set:", metrics)
    accelerator.end_training()
    return metrics


def main():
    parser = base_parser()
    # New Code #
    parser.add_argument("--num_folds", type=int, default=3, help="The number of splits to perform across the dataset")
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
