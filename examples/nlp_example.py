"""BERT sequence classification on MRPC-shaped data — the smallest complete training script.

Counterpart of the reference's examples/nlp_example.py (BERT-base-cased on GLUE/MRPC); there is no network in this
environment, so the sentence pairs are synthetic with the real task's shapes (3668 train / 408 validation pairs,
``[CLS] a [SEP] b [SEP]``, pad-to-longest collate).  Runs unchanged on one CPU process, on 2 CPU processes (gloo), or
on B200s:

    python examples/nlp_example.py --cpu --model tiny --num_epochs 1
    accelerate launch --num_processes 2 --cpu examples/nlp_example.py --model tiny
    accelerate launch --num_processes 8 --mixed_precision bf16 examples/nlp_example.py
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))  # run from a source checkout

import torch
from torch.optim import AdamW
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator, DistributedType
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification, pad_collate, synthetic_mrpc
from accelerate_hpc_test_b200.utils import set_seed

MAX_GPU_BATCH_SIZE = 16
EVAL_BATCH_SIZE = 32


def get_dataloaders(accelerator: Accelerator, batch_size: int = 16, n_train: int = 3668, n_eval: int = 408, vocab_size: int = 28996):
    """Train / validation loaders over synthetic MRPC pairs (built on the main process first, like a cached tokenised set)."""
    with accelerator.main_process_first():
        train = synthetic_mrpc(n_train, vocab_size=vocab_size, seed=0)
        evald = synthetic_mrpc(n_eval, vocab_size=vocab_size, seed=1)
    train_dataloader = DataLoader(train, shuffle=True, collate_fn=pad_collate, batch_size=batch_size, drop_last=True)
    eval_dataloader = DataLoader(evald, shuffle=False, collate_fn=pad_collate, batch_size=EVAL_BATCH_SIZE, drop_last=False)
    return train_dataloader, eval_dataloader


def linear_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps):
    def lr_lambda(step):
        if step < num_warmup_steps:
            return step / max(1, num_warmup_steps)
        return max(0.0, (num_training_steps - step) / max(1, num_training_steps - num_warmup_steps))

    return torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda)


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    lr, num_epochs, seed, batch_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"])

    # big batches are run as several accumulated micro-batches
    gradient_accumulation_steps = 1
    if batch_size > MAX_GPU_BATCH_SIZE and accelerator.distributed_type != DistributedType.XLA:
        gradient_accumulation_steps = batch_size // MAX_GPU_BATCH_SIZE
        batch_size = MAX_GPU_BATCH_SIZE

    set_seed(seed)
    cfg = BertConfig.tiny() if args.model == "tiny" else BertConfig.base_cased()
    train_dataloader, eval_dataloader = get_dataloaders(accelerator, batch_size, args.train_size, args.eval_size, cfg.vocab_size)
    model = BertForSequenceClassification(cfg).to(accelerator.device)
    optimizer = AdamW(params=model.parameters(), lr=lr)
    lr_scheduler = linear_schedule_with_warmup(
        optimizer, num_warmup_steps=100 if args.model != "tiny" else 2,
        num_training_steps=(len(train_dataloader) * num_epochs) // gradient_accumulation_steps,
    )
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )

    metrics = {}
    for epoch in range(num_epochs):
        model.train()
        for step, batch in enumerate(train_dataloader):
            outputs = model(**batch)
            loss = outputs.loss / gradient_accumulation_steps
            accelerator.backward(loss)
            if step % gradient_accumulation_steps == 0:
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()

        model.eval()
        correct = total = 0
        for step, batch in enumerate(eval_dataloader):
            with torch.no_grad():
                outputs = model(**batch)
            predictions = outputs.logits.argmax(dim=-1)
            predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
            correct += int((predictions == references).sum())
            total += int(references.numel())
        metrics = {"accuracy": correct / max(1, total), "eval_samples": total}
        accelerator.print(f"epoch {epoch}:", metrics)
    accelerator.end_training()
    return metrics


def main():
    parser = argparse.ArgumentParser(description="Simple example of training script.")
    parser.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16", "fp8"],
                        help="Whether to use mixed precision. Choose between fp16 and bf16 (bfloat16). Bf16 requires a B200.")
    parser.add_argument("--cpu", action="store_true", help="If passed, will train on the CPU.")
    parser.add_argument("--model", default="base", choices=["base", "tiny"], help="BERT-base-cased shape or a tiny test model.")
    parser.add_argument("--num_epochs", type=int, default=3)
    parser.add_argument("--train_size", type=int, default=3668)
    parser.add_argument("--eval_size", type=int, default=408)
    args = parser.parse_args()
    config = {"lr": 2e-5, "num_epochs": args.num_epochs, "seed": 42, "batch_size": 16}
    training_function(config, args)


if __name__ == "__main__":
    main()
