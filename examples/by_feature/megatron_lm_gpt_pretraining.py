"""Causal-LM pre-training in the Megatron-LM style: the training loop hands a whole *global* batch to an engine whose call is one
full optimisation step — it cuts the batch into micro-batches, runs forward / backward for each, steps the optimizer and the schedule
and returns the averaged loss.  Megatron-LM itself is not part of this framework; its tensor-parallel degree maps onto the native
tensor-parallel engine (``ParallelismConfig``) and the engine API is provided by ``MegatronEngine`` on the native Llama
(reference examples/by_feature/megatron_lm_gpt_pretraining.py).

    python examples/by_feature/megatron_lm_gpt_pretraining.py --cpu --model tiny --num_micro_batches 2 --max_train_steps 6
"""

import argparse
import math
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW

# New Code #
from accelerate_hpc_test_b200.utils import (MegatronEngine, MegatronLMOptimizerWrapper, MegatronLMSchedulerWrapper,
                                            avg_losses_across_data_parallel_group, set_seed)


def blocks_of_tokens(n, block_size, vocab, seed):
    """What ``group_texts`` produces from a tokenised corpus: fixed-length blocks, labels = inputs."""
    g = torch.Generator().manual_seed(seed)
    return [{"input_ids": torch.randint(0, vocab, (block_size,), generator=g)} for _ in range(n)]


def collate(samples):
    ids = torch.stack([s["input_ids"] for s in samples])
    return {"input_ids": ids, "labels": ids.clone()}


def main():
    parser = argparse.ArgumentParser(description="GPT-style pre-training through the Megatron-LM engine API")
    parser.add_argument("--cpu", action="store_true")
    parser.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16"])
    parser.add_argument("--model", default="llama3-1b", choices=["llama3-1b", "tiny"])
    parser.add_argument("--block_size", type=int, default=128)
    parser.add_argument("--per_device_train_batch_size", type=int, default=4, help="sequences per rank and optimisation step")
    parser.add_argument("--num_micro_batches", type=int, default=2)
    parser.add_argument("--learning_rate", type=float, default=5e-4)
    parser.add_argument("--weight_decay", type=float, default=0.0)
    parser.add_argument("--num_warmup_steps", type=int, default=2)
    parser.add_argument("--max_train_steps", type=int, default=20)
    parser.add_argument("--train_size", type=int, default=256)
    parser.add_argument("--eval_size", type=int, default=32)
    parser.add_argument("--output_dir", type=str, default=None)
    args = parser.parse_args()

    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    set_seed(42)
    cfg = (LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512, num_attention_heads=4, num_key_value_heads=2)
           if args.model == "tiny" else LlamaConfig.llama3_1b())
    model = LlamaForCausalLM(cfg, dtype=torch.float32 if args.cpu else torch.bfloat16, device=accelerator.device)
    optimizer = FusedAdamW(model.parameters(), lr=args.learning_rate, weight_decay=args.weight_decay)
    warm, total = args.num_warmup_steps, args.max_train_steps
    lr_scheduler = torch.optim.lr_scheduler.LambdaLR(
        optimizer, lambda step: step / max(1, warm) if step < warm else max(0.0, (total - step) / max(1, total - warm)))
    train_dataloader = DataLoader(blocks_of_tokens(args.train_size, args.block_size, cfg.vocab_size, 0), shuffle=True, collate_fn=collate,
                                  batch_size=args.per_device_train_batch_size, drop_last=True)
    eval_dataloader = DataLoader(blocks_of_tokens(args.eval_size, args.block_size, cfg.vocab_size, 1), collate_fn=collate,
                                 batch_size=args.per_device_train_batch_size, drop_last=True)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )

    # New Code #
    # the engine owns the step: micro-batching, backward, optimizer and schedule all happen inside `engine(**batch)`
    engine = MegatronEngine(accelerator, model, MegatronLMOptimizerWrapper(optimizer.optimizer),
                            MegatronLMSchedulerWrapper(lr_scheduler.scheduler, optimizer.optimizer))
    engine.num_micro_batches = args.num_micro_batches
    total_batch_size = args.per_device_train_batch_size * accelerator.num_processes
    accelerator.print(f"  Total train batch size (w. parallel, distributed & micro-batches) = {total_batch_size}")
    accelerator.print(f"  Micro-batches per step = {args.num_micro_batches}, total optimization steps = {args.max_train_steps}")

    completed, epoch = 0, 0
    while completed < args.max_train_steps:
        engine.train()
        for batch in train_dataloader:
            # New Code #
            # no backward(), no optimizer.step(), no scheduler.step() here
            loss = engine(**batch)
            completed += 1
            if completed >= args.max_train_steps:
                break
        engine.eval()
        losses = []
        for batch in eval_dataloader:
            with torch.no_grad():
                loss = engine(**batch)
            # New Code #
            # losses are reported per data-parallel group: average them over the groups
            losses.append(avg_losses_across_data_parallel_group([loss]))
        eval_loss = torch.stack([x.reshape(()) for x in losses]).mean().item()
        accelerator.print(f"epoch {epoch}: perplexity: {math.exp(min(20.0, eval_loss)):.2f} eval_loss: {eval_loss:.4f} "
                          f"lr: {optimizer.param_groups[0]['lr']:.2e} steps: {completed}")
        epoch += 1
    if args.output_dir is not None:
        # New Code #
        engine.save_checkpoint(args.output_dir)
    accelerator.end_training()


if __name__ == "__main__":
    main()
