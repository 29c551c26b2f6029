"""nlp_example.py plus everything a real run needs: experiment tracking, checkpoints every N steps / every epoch, and
resuming in the middle of an epoch (counterpart of the reference's examples/complete_nlp_example.py).

    python examples/complete_nlp_example.py --cpu --model tiny --num_epochs 2 --checkpointing_steps epoch --output_dir /tmp/run
    python examples/complete_nlp_example.py --cpu --model tiny --num_epochs 2 --resume_from_checkpoint /tmp/run/epoch_0
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))  # run from a source checkout

import torch
from torch.optim import AdamW

from accelerate_hpc_test_b200 import Accelerator, DistributedType
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification
from accelerate_hpc_test_b200.utils import set_seed
from nlp_example import MAX_GPU_BATCH_SIZE, get_dataloaders, linear_schedule_with_warmup


def training_function(config, args):
    if args.with_tracking:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, log_with=args.log_with, project_dir=args.project_dir)
    else:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)

    # "epoch", a number of steps, or nothing
    checkpointing_steps = args.checkpointing_steps
    if checkpointing_steps is not None and checkpointing_steps != "epoch":
        if not str(checkpointing_steps).isdigit():
            raise ValueError(f"Argument `checkpointing_steps` must be either a number or `epoch`. `{checkpointing_steps}` passed.")
        checkpointing_steps = int(checkpointing_steps)

    lr, num_epochs, seed, batch_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"])
    if args.with_tracking:
        run = os.path.split(__file__)[-1].split(".")[0]
        accelerator.init_trackers(run, config)

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

    overall_step = 0
    starting_epoch = 0
    resume_step = None
    if args.resume_from_checkpoint:
        accelerator.print(f"Resumed from checkpoint: {args.resume_from_checkpoint}")
        accelerator.load_state(args.resume_from_checkpoint)
        training_difference = os.path.splitext(os.path.basename(os.path.normpath(args.resume_from_checkpoint)))[0]
        if "epoch" in training_difference:
            starting_epoch = int(training_difference.replace("epoch_", "")) + 1
        else:
            resume_step = int(training_difference.replace("step_", ""))
            starting_epoch = resume_step // len(train_dataloader)
            resume_step -= starting_epoch * len(train_dataloader)
            overall_step = starting_epoch * len(train_dataloader) + resume_step

    metrics = {}
    for epoch in range(starting_epoch, num_epochs):
        model.train()
        total_loss = 0.0
        if args.resume_from_checkpoint and epoch == starting_epoch and resume_step is not None:
            # skip the batches this epoch already consumed before the checkpoint
            active_dataloader = accelerator.skip_first_batches(train_dataloader, resume_step)
        else:
            active_dataloader = train_dataloader
        for step, batch in enumerate(active_dataloader):
            outputs = model(**batch)
            loss = outputs.loss / gradient_accumulation_steps
            if args.with_tracking:
                total_loss += loss.detach().float()
            accelerator.backward(loss)
            if step % gradient_accumulation_steps == 0:
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
            overall_step += 1
            if isinstance(checkpointing_steps, int) and overall_step % checkpointing_steps == 0:
                output_dir = f"step_{overall_step}"
                if args.output_dir is not None:
                    output_dir = os.path.join(args.output_dir, output_dir)
                accelerator.save_state(output_dir)

        model.eval()
        correct = total = 0
        for step, batch in enumerate(eval_dataloader):
            with torch.no_grad():
                outputs = model(**batch)
            predictions = outputs.logits.argmax(dim=-1)
            predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
            correct += int((predictions == references).sum())
            total += int(references.numel())
        metrics = {"accuracy": correct / max(1, total)}
        accelerator.print(f"epoch {epoch}:", metrics)
        if args.with_tracking:
            accelerator.log(
                {"accuracy": metrics["accuracy"], "train_loss": float(total_loss) / max(1, len(train_dataloader)), "epoch": epoch},
                step=epoch,
            )
        if checkpointing_steps == "epoch":
            output_dir = f"epoch_{epoch}"
            if args.output_dir is not None:
                output_dir = os.path.join(args.output_dir, output_dir)
            accelerator.save_state(output_dir)
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
    parser.add_argument("--checkpointing_steps", type=str, default=None,
                        help="Whether the various states should be saved at the end of every n steps, or 'epoch' for each epoch.")
    parser.add_argument("--resume_from_checkpoint", type=str, default=None, help="If the training should continue from a checkpoint folder.")
    parser.add_argument("--with_tracking", action="store_true", help="Whether to load in all available experiment trackers and log to them.")
    parser.add_argument("--log_with", type=str, default="jsonl",
                        help="Tracker(s) to use: `all` for every installed one (wandb etc. need their own login), or a name such as "
                             "`tensorboard` / `jsonl` (the built-in dependency-free tracker).")
    parser.add_argument("--output_dir", type=str, default=".", help="Where to store the checkpoints.")
    parser.add_argument("--project_dir", type=str, default="logs", help="Location of logs and project directories.")
    args = parser.parse_args()
    config = {"lr": 2e-5, "num_epochs": args.num_epochs, "seed": 42, "batch_size": 16}
    training_function(config, args)


if __name__ == "__main__":
    main()
