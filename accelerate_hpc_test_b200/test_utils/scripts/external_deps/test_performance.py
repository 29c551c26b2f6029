"""Does distributed training still *learn*?  A small BERT is trained on a learnable synthetic pair-classification task; the best
validation accuracy has to clear ``--performance_lower_bound``, which catches silent breakage that parity tests on single steps miss
(gradients not averaged, a shard updated twice, a schedule that never steps).

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_performance \
        --performance_lower_bound 0.80
"""

import argparse
import json
import os

import torch
from torch.utils.data import DataLoader, Dataset

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification
from accelerate_hpc_test_b200.utils import set_seed


class PairTask(Dataset):
    """Two token segments; label 1 iff the marker token occurs in the second one.  Easy enough for a two-layer model to learn in a few
    epochs, impossible to get right without training (the classes are balanced)."""

    MARKER = 4

    def __init__(self, size, vocab, seed):
        g = torch.Generator().manual_seed(seed)
        self.rows = []
        for _ in range(size):
            a = torch.randint(5, vocab, (6,), generator=g)
            b = torch.randint(5, vocab, (6,), generator=g)
            label = int(torch.rand((), generator=g) < 0.5)
            if label:
                b[int(torch.randint(0, 6, (), generator=g))] = self.MARKER
            ids = torch.cat([torch.tensor([1]), a, torch.tensor([2]), b, torch.tensor([2])])
            types = torch.cat([torch.zeros(8, dtype=torch.long), torch.ones(7, dtype=torch.long)])
            self.rows.append({"input_ids": ids, "token_type_ids": types, "attention_mask": torch.ones_like(ids), "labels": torch.tensor(label)})

    def __len__(self):
        return len(self.rows)

    def __getitem__(self, index):
        return self.rows[index]


def collate(rows):
    return {key: torch.stack([row[key] for row in rows]) for key in rows[0]}


def training_function(args):
    accelerator = Accelerator()
    set_seed(42)
    cfg = BertConfig.tiny(vocab_size=64) if hasattr(BertConfig, "tiny") else BertConfig()
    train = DataLoader(PairTask(args.train_size, cfg.vocab_size, 0), shuffle=True, collate_fn=collate, batch_size=args.batch_size, drop_last=True)
    evald = DataLoader(PairTask(args.eval_size, cfg.vocab_size, 1), shuffle=False, collate_fn=collate, batch_size=32)
    model = BertForSequenceClassification(cfg)
    optimizer = torch.optim.AdamW(model.parameters(), lr=args.lr)
    steps = (len(train) // accelerator.num_processes) * args.num_epochs
    scheduler = torch.optim.lr_scheduler.OneCycleLR(optimizer, max_lr=args.lr, total_steps=max(2, steps))
    model, optimizer, train, evald, scheduler = accelerator.prepare(model, optimizer, train, evald, scheduler)
    best, history = 0.0, {}
    for epoch in range(args.num_epochs):
        model.train()
        for batch in train:
            accelerator.backward(model(**batch).loss)
            optimizer.step()
            scheduler.step()
            optimizer.zero_grad()
        model.eval()
        correct = total = 0
        for batch in evald:
            with torch.no_grad():
                predictions = model(**batch).logits.argmax(-1)
            predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
            correct += int((predictions == references).sum())
            total += len(references)
        history[f"epoch-{epoch}"] = correct / total
        best = max(best, correct / total)
        accelerator.print(f"epoch {epoch}: accuracy {correct / total:.3f}")
        if args.performance_lower_bound is not None and best >= args.performance_lower_bound:
            break
    if args.performance_lower_bound is not None:
        assert best >= args.performance_lower_bound, f"Best performance metric {best} is lower than the lower bound {args.performance_lower_bound}"
    accelerator.wait_for_everyone()
    if accelerator.is_main_process and args.output_dir is not None:
        os.makedirs(args.output_dir, exist_ok=True)
        with open(os.path.join(args.output_dir, "all_results.json"), "w") as f:
            json.dump(history, f)
    accelerator.print(f"test_performance: best accuracy {best:.3f}: all passed")
    accelerator.end_training()


def main():
    parser = argparse.ArgumentParser(description="Training must reach a validation accuracy.")
    parser.add_argument("--performance_lower_bound", type=float, default=None, help="Fail unless the best validation accuracy reaches this.")
    parser.add_argument("--num_epochs", type=int, default=12)
    parser.add_argument("--train_size", type=int, default=1024)
    parser.add_argument("--eval_size", type=int, default=200)
    parser.add_argument("--batch_size", type=int, default=16)
    parser.add_argument("--lr", type=float, default=3e-3)
    parser.add_argument("--output_dir", type=str, default=None)
    training_function(parser.parse_args())


if __name__ == "__main__":
    main()
