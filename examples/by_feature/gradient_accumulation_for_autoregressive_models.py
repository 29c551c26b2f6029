"""Gradient accumulation for a causal language model with variable-length samples.

Averaging the *per-micro-batch mean* losses weights every micro-batch equally, however many real (non-padding) tokens it holds; what
one wants is the mean over all tokens of the effective batch.  So each optimisation step first counts the target tokens of the
micro-batches it is going to accumulate — on every rank, summed with ``accelerator.gather`` — and the summed token losses are divided by
that count.  Two corrections undo what the framework does for the ordinary case: ``accelerator.backward`` divides by the number of
accumulation steps and the data-parallel reduction averages over ranks, so the loss is multiplied by both
(reference examples/by_feature/gradient_accumulation_for_autoregressive_models.py).

    python examples/by_feature/gradient_accumulation_for_autoregressive_models.py --cpu --model tiny --gradient_accumulation_steps 2
"""

import argparse
import contextlib
import math
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.utils import set_seed


def synthetic_corpus(n, vocab, max_len, seed):
    """Token sequences of very different lengths (8 .. max_len), so that per-batch means and the token mean differ."""
    g = torch.Generator().manual_seed(seed)
    lengths = torch.randint(8, max_len + 1, (n,), generator=g).tolist()
    return [torch.randint(1, vocab, (length,), generator=g) for length in lengths]


def collate(samples):
    width = max(len(s) for s in samples)
    ids = torch.zeros(len(samples), width, dtype=torch.long)
    labels = torch.full((len(samples), width), -100, dtype=torch.long)
    for i, s in enumerate(samples):
        ids[i, : len(s)] = s
        labels[i, : len(s)] = s
    return {"input_ids": ids, "labels": labels}


def training_function(args):
    gradient_accumulation_steps = int(args.gradient_accumulation_steps)
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, gradient_accumulation_steps=gradient_accumulation_steps)
    set_seed(42)
    cfg = (LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512, num_attention_heads=4, num_key_value_heads=2)
           if args.model == "tiny" else LlamaConfig.llama3_1b())
    dtype = torch.float32 if args.cpu else torch.bfloat16
    model = LlamaForCausalLM(cfg, dtype=dtype, device=accelerator.device)
    optimizer = torch.optim.AdamW(model.parameters(), lr=args.lr)
    train = synthetic_corpus(args.train_size, cfg.vocab_size, args.max_length, seed=0)
    evald = synthetic_corpus(args.eval_size, cfg.vocab_size, args.max_length, seed=1)
    train_dataloader = DataLoader(train, shuffle=True, collate_fn=collate, batch_size=args.batch_size, drop_last=True)
    eval_dataloader = DataLoader(evald, shuffle=False, collate_fn=collate, batch_size=args.batch_size)
    model, optimizer, train_dataloader, eval_dataloader = accelerator.prepare(model, optimizer, train_dataloader, eval_dataloader)

    # New Code #
    updates_per_epoch = math.ceil(len(train_dataloader) / gradient_accumulation_steps)
    leftover = len(train_dataloader) % gradient_accumulation_steps or gradient_accumulation_steps
    for epoch in range(args.num_epochs):
        model.train()
        batches = iter(train_dataloader)
        total_loss = 0.0
        for update in range(updates_per_epoch):
            # New Code #
            # pull the micro-batches of this optimisation step ahead of time and count their target tokens (labels are shifted
            # by one inside the model: position 0 is never a target)
            n_micro = gradient_accumulation_steps if update < updates_per_epoch - 1 else leftover
            micro_batches = [next(batches) for _ in range(n_micro)]
            local_tokens = sum(int((b["labels"][:, 1:] != -100).sum()) for b in micro_batches)
            num_items_in_batch = accelerator.gather(torch.tensor([local_tokens], device=accelerator.device)).sum().item()
            for i, batch in enumerate(micro_batches):
                # New Code #
                # the last micro-batch of an epoch may come before the accumulation count is reached: force the sync there
                if i < len(micro_batches) - 1 and accelerator.num_processes > 1:
                    context = accelerator.no_sync(model)
                else:
                    context = contextlib.nullcontext()
                with context:
                    logits = model(input_ids=batch["input_ids"]).logits
                    # sum (not mean) of the token losses, divided by the tokens of the whole effective batch
                    loss = torch.nn.functional.cross_entropy(logits[:, :-1].reshape(-1, logits.shape[-1]).float(),
                                                             batch["labels"][:, 1:].reshape(-1), ignore_index=-100, reduction="sum")
                    loss = loss / num_items_in_batch
                    # data-parallel reduction averages over ranks and backward() divides by the accumulation steps: undo both
                    loss = loss * accelerator.num_processes * gradient_accumulation_steps
                    accelerator.backward(loss)
                    total_loss += float(loss.detach()) / (accelerator.num_processes * gradient_accumulation_steps)
            optimizer.step()
            optimizer.zero_grad()
        model.eval()
        losses, tokens = 0.0, 0
        for batch in eval_dataloader:
            with torch.no_grad():
                logits = model(input_ids=batch["input_ids"]).logits
            loss = torch.nn.functional.cross_entropy(logits[:, :-1].reshape(-1, logits.shape[-1]).float(), batch["labels"][:, 1:].reshape(-1),
                                                     ignore_index=-100, reduction="sum")
            count = (batch["labels"][:, 1:] != -100).sum()
            loss, count = accelerator.gather_for_metrics((loss.reshape(1), count.reshape(1)))
            losses += float(loss.sum())
            tokens += int(count.sum())
        accelerator.print(f"epoch {epoch}:", {"train_loss": total_loss / updates_per_epoch, "eval_loss": losses / max(1, tokens),
                                              "perplexity": math.exp(min(20.0, losses / max(1, tokens)))})
    accelerator.end_training()


def main():
    parser = argparse.ArgumentParser(description="Gradient accumulation with a token-level loss for causal language models.")
    parser.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16", "fp8"])
    parser.add_argument("--gradient_accumulation_steps", type=int, default=2, help="The number of minibatches to be ran before gradients are accumulated.")
    parser.add_argument("--cpu", action="store_true", help="If passed, will train on the CPU.")
    parser.add_argument("--model", default="llama3-1b", choices=["llama3-1b", "tiny"])
    parser.add_argument("--num_epochs", type=int, default=1)
    parser.add_argument("--batch_size", type=int, default=4)
    parser.add_argument("--max_length", type=int, default=128)
    parser.add_argument("--train_size", type=int, default=256)
    parser.add_argument("--eval_size", type=int, default=32)
    parser.add_argument("--lr", type=float, default=2e-4)
    training_function(parser.parse_args())


if __name__ == "__main__":
    main()
