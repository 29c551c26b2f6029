"""Gradient accumulation with ``accelerator.accumulate``: ``optimizer.step()`` / ``zero_grad()`` are called on every
micro-batch and become no-ops until ``gradient_accumulation_steps`` batches have been seen (reference
examples/by_feature/gradient_accumulation.py).

    python examples/by_feature/gradient_accumulation.py --cpu --model tiny --gradient_accumulation_steps 4
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch
from torch.optim import AdamW

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification
from accelerate_hpc_test_b200.utils import set_seed
from nlp_example import get_dataloaders, linear_schedule_with_warmup


def training_function(config, args):
    # New Code #
    gradient_accumulation_steps = int(args.gradient_accumulation_steps)
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, gradient_accumulation_steps=gradient_accumulation_steps)
    lr, num_epochs, seed, batch_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"])
    set_seed(seed)
    cfg = BertConfig.tiny() if args.model == "tiny" else BertConfig.base_cased()
    train_dataloader, eval_dataloader = get_dataloaders(accelerator, batch_size, args.train_size, args.eval_size, cfg.vocab_size)
    model = BertForSequenceClassification(cfg).to(accelerator.device)
    optimizer = AdamW(params=model.parameters(), lr=lr)
    lr_scheduler = linear_schedule_with_warmup(optimizer, 2, len(train_dataloader) * num_epochs)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )
    for epoch in range(num_epochs):
        model.train()
        for step, batch in enumerate(train_dataloader):
            # New Code #
            # inside the context gradients are only synchronised (and the optimizer only steps) every N-th batch or at
            # the end of the dataloader; the loss is divided by N in accelerator.backward
            with accelerator.accumulate(model):
                output = model(**batch)
                accelerator.backward(output.loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
        model.eval()
        correct = total = 0
        for batch in eval_dataloader:
            with torch.no_grad():
                predictions = model(**batch).logits.argmax(dim=-1)
            predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
            correct += int((predictions == references).sum())
            total += int(references.numel())
        accelerator.print(f"epoch {epoch}:", {"accuracy": correct / max(1, total)})
    accelerator.end_training()


def main():
    parser = argparse.ArgumentParser(description="Simple example of training script.")
    parser.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16", "fp8"])
    parser.add_argument("--gradient_accumulation_steps", type=int, default=1, help="The number of minibatches to be ran before gradients are accumulated.")
    parser.add_argument("--cpu", action="store_true", help="If passed, will train on the CPU.")
    parser.add_argument("--model", default="base", choices=["base", "tiny"])
    parser.add_argument("--num_epochs", type=int, default=3)
    parser.add_argument("--train_size", type=int, default=3668)
    parser.add_argument("--eval_size", type=int, default=408)
    args = parser.parse_args()
    config = {"lr": 2e-5, "num_epochs": args.num_epochs, "seed": 42, "batch_size": 16}
    training_function(config, args)


if __name__ == "__main__":
    main()
