"""cv_example.py with everything the by-feature scripts add that makes sense for it: experiment tracking, checkpoints every N steps
or every epoch, resuming in the middle of an epoch (reference examples/complete_cv_example.py).

    python examples/complete_cv_example.py --cpu --num_epochs 2 --checkpointing_steps epoch --with_tracking --output_dir out
    python examples/complete_cv_example.py --cpu --num_epochs 2 --resume_from_checkpoint out/epoch_0
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch
from torch.optim.lr_scheduler import OneCycleLR
from torch.utils.data import DataLoader

from cv_example import SmallResNet, SyntheticPets

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.utils import set_seed


def _checkpoint_every(spec):
    """``--checkpointing_steps``: a number of steps, the word ``epoch``, or nothing."""
    if spec is None or spec == "epoch":
        return spec
    if not str(spec).isdigit():
        raise ValueError(f"Argument `checkpointing_steps` must be either a number or `epoch`. `{spec}` passed.")
    return int(spec)


def _where_to_resume(checkpoint_dir, steps_per_epoch):
    """``(first epoch to run, batches of it already consumed, global step)`` from a folder named ``epoch_N`` or ``step_N``."""
    name = os.path.basename(os.path.normpath(checkpoint_dir))
    if name.startswith("epoch_"):
        epoch = int(name[len("epoch_"):]) + 1
        return epoch, None, epoch * steps_per_epoch
    done = int(name[len("step_"):])
    epoch, within = divmod(done, steps_per_epoch)
    return epoch, within, done


def training_function(config, args):
    if args.with_tracking:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, log_with=args.log_with, project_dir=args.project_dir)
    else:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    every = _checkpoint_every(args.checkpointing_steps)
    lr, num_epochs, seed, batch_size, image_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"]), config["image_size"]
    if args.with_tracking:
        accelerator.init_trackers(os.path.splitext(os.path.basename(__file__))[0], config)
    set_seed(seed)
    train_dataloader = DataLoader(SyntheticPets(args.train_size, image_size, args.num_classes, seed=0), shuffle=True, batch_size=batch_size)
    eval_dataloader = DataLoader(SyntheticPets(args.eval_size, image_size, args.num_classes, seed=1), shuffle=False, batch_size=batch_size)
    model = SmallResNet(args.num_classes).to(accelerator.device)
    mean = torch.tensor([0.5, 0.5, 0.5])[None, :, None, None].to(accelerator.device)
    std = torch.tensor([0.6, 0.6, 0.6])[None, :, None, None].to(accelerator.device)
    optimizer = torch.optim.Adam(params=model.parameters(), lr=lr / 25)
    lr_scheduler = OneCycleLR(optimizer=optimizer, max_lr=lr, epochs=num_epochs, steps_per_epoch=len(train_dataloader))
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )

    overall_step, starting_epoch, resume_step = 0, 0, None
    if args.resume_from_checkpoint:
        accelerator.print(f"Resumed from checkpoint: {args.resume_from_checkpoint}")
        accelerator.load_state(args.resume_from_checkpoint)
        starting_epoch, resume_step, overall_step = _where_to_resume(args.resume_from_checkpoint, len(train_dataloader))

    def checkpoint(name):
        accelerator.save_state(os.path.join(args.output_dir, name) if args.output_dir is not None else name)

    metrics = {}
    for epoch in range(starting_epoch, num_epochs):
        model.train()
        total_loss = 0.0
        active_dataloader = train_dataloader
        if resume_step is not None and epoch == starting_epoch:
            # the batches of this epoch that were consumed before the checkpoint are skipped at the index level (nothing is loaded)
            active_dataloader = accelerator.skip_first_batches(train_dataloader, resume_step)
        for batch in active_dataloader:
            inputs = (batch["image"] - mean) / std
            loss = torch.nn.functional.cross_entropy(model(inputs), batch["label"])
            if args.with_tracking:
                total_loss += loss.detach().float()
            accelerator.backward(loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
            overall_step += 1
            if isinstance(every, int) and overall_step % every == 0:
                checkpoint(f"step_{overall_step}")
        model.eval()
        accurate = num_elems = 0
        for batch in eval_dataloader:
            with torch.no_grad():
                predictions = model((batch["image"] - mean) / std).argmax(dim=-1)
            predictions, references = accelerator.gather_for_metrics((predictions, batch["label"]))
            accurate += int((predictions == references).sum())
            num_elems += int(references.numel())
        metrics = {"accuracy": accurate / max(1, num_elems), "eval_samples": num_elems}
        accelerator.print(f"epoch {epoch}: {100 * metrics['accuracy']:.2f}", metrics)
        if args.with_tracking:
            accelerator.log({"accuracy": 100 * metrics["accuracy"], "train_loss": float(total_loss) / max(1, len(train_dataloader)), "epoch": epoch},
                            step=overall_step)
        if every == "epoch":
            checkpoint(f"epoch_{epoch}")
    accelerator.end_training()
    return metrics


def main():
    parser = argparse.ArgumentParser(description="Simple example of training script.")
    parser.add_argument("--mixed_precision", type=str, default=None, choices=["no", "fp16", "bf16", "fp8"])
    parser.add_argument("--cpu", action="store_true", help="If passed, will train on the CPU.")
    parser.add_argument("--num_epochs", type=int, default=3)
    parser.add_argument("--train_size", type=int, default=2048)
    parser.add_argument("--eval_size", type=int, default=256)
    parser.add_argument("--num_classes", type=int, default=8)
    parser.add_argument("--image_size", type=int, default=64)
    parser.add_argument("--checkpointing_steps", type=str, default=None,
                        help="Whether the various states should be saved at the end of every n steps, or 'epoch' for each epoch.")
    parser.add_argument("--output_dir", type=str, default=".", help="Optional save directory where all checkpoint folders will be stored.")
    parser.add_argument("--resume_from_checkpoint", type=str, default=None, help="If the training should continue from a checkpoint folder.")
    parser.add_argument("--with_tracking", action="store_true", help="Whether to load in all available experiment trackers from the environment and use them for logging.")
    parser.add_argument("--log_with", type=str, default="jsonl", help="Tracker(s) to log to: `all`, or a name such as `tensorboard` / `jsonl`.")
    parser.add_argument("--project_dir", type=str, default="logs", help="Location on where to store experiment tracking logs and relevent project information")
    args = parser.parse_args()
    config = {"lr": 3e-2, "num_epochs": args.num_epochs, "seed": 42, "batch_size": 64, "image_size": args.image_size}
    training_function(config, args)


if __name__ == "__main__":
    main()
