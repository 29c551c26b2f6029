"""Pieces shared by the by-feature examples: argument parser, model / data / schedule construction, evaluation loop.
Every script in this folder is nlp_example.py plus ONE feature, marked with ``# New Code #`` comments."""

import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch
from torch.optim import AdamW

from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification
from accelerate_hpc_test_b200.utils import set_seed
from nlp_example import get_dataloaders, linear_schedule_with_warmup  # noqa: F401


def base_parser(description="Simple example of training script."):
    parser = argparse.ArgumentParser(description=description)
    parser.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16", "fp8"])
    parser.add_argument("--cpu", action="store_true", help="If passed, will train on the CPU.")
    parser.add_argument("--model", default="base", choices=["base", "tiny"])
    parser.add_argument("--num_epochs", type=int, default=3)
    parser.add_argument("--train_size", type=int, default=3668)
    parser.add_argument("--eval_size", type=int, default=408)
    return parser


def default_config(args):
    return {"lr": 2e-5, "num_epochs": args.num_epochs, "seed": 42, "batch_size": 16}


def build(accelerator, config, args, batch_size=None):
    """Model, optimizer, loaders and schedule — not yet prepared."""
    set_seed(int(config["seed"]))
    cfg = BertConfig.tiny() if args.model == "tiny" else BertConfig.base_cased()
    train_dl, eval_dl = get_dataloaders(accelerator, batch_size or int(config["batch_size"]), args.train_size, args.eval_size, cfg.vocab_size)
    model = BertForSequenceClassification(cfg).to(accelerator.device)
    optimizer = AdamW(params=model.parameters(), lr=config["lr"])
    scheduler = linear_schedule_with_warmup(optimizer, 2 if args.model == "tiny" else 100, len(train_dl) * int(config["num_epochs"]))
    return model, optimizer, train_dl, eval_dl, scheduler


def evaluate(accelerator, model, eval_dataloader):
    model.eval()
    correct = total = 0
    for batch in eval_dataloader:
        with torch.no_grad():
            predictions = model(**batch).logits.argmax(dim=-1)
        predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
        correct += int((predictions == references).sum())
        total += int(references.numel())
    return {"accuracy": correct / max(1, total), "eval_samples": total}
