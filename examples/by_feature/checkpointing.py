"""Saving and resuming the whole training state (model, optimizer, scheduler, RNG, dataloader position) with
``accelerator.save_state`` / ``load_state`` / ``skip_first_batches`` (reference examples/by_feature/checkpointing.py).

    python examples/by_feature/checkpointing.py --cpu --model tiny --checkpointing_steps epoch --output_dir /tmp/ckpt
    python examples/by_feature/checkpointing.py --cpu --model tiny --resume_from_checkpoint /tmp/ckpt/epoch_0
"""

import os

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    # New Code #
    checkpointing_steps = args.checkpointing_steps
    if checkpointing_steps is not None and checkpointing_steps != "epoch":
        checkpointing_steps = int(checkpointing_steps)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args))

    # New Code #
    overall_step, starting_epoch, resume_step = 0, 0, None
    if args.resume_from_checkpoint:
        accelerator.load_state(args.resume_from_checkpoint)
        name = os.path.basename(os.path.normpath(args.resume_from_checkpoint))
        if name.startswith("epoch_"):
            starting_epoch = int(name[len("epoch_"):]) + 1
        else:
            resume_step = int(name[len("step_"):])
            starting_epoch, resume_step = divmod(resume_step, len(train_dataloader))
            overall_step = starting_epoch * len(train_dataloader) + resume_step

    for epoch in range(starting_epoch, int(config["num_epochs"])):
        model.train()
        # New Code #
        loader = train_dataloader
        if resume_step is not None and epoch == starting_epoch:
            loader = accelerator.skip_first_batches(train_dataloader, resume_step)
        for batch in loader:
            accelerator.backward(model(**batch).loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
            # New Code #
            overall_step += 1
            if isinstance(checkpointing_steps, int) and overall_step % checkpointing_steps == 0:
                accelerator.save_state(os.path.join(args.output_dir, f"step_{overall_step}"))
        accelerator.print(f"epoch {epoch}:", evaluate(accelerator, model, eval_dataloader))
        # New Code #
        if checkpointing_steps == "epoch":
            accelerator.save_state(os.path.join(args.output_dir, f"epoch_{epoch}"))
    accelerator.end_training()


def main():
    parser = base_parser()
    parser.add_argument("--checkpointing_steps", type=str, default=None, help="Save every n steps, or 'epoch' for each epoch.")
    parser.add_argument("--output_dir", type=str, default=".", help="Where to store the checkpoints.")
    parser.add_argument("--resume_from_checkpoint", type=str, default=None, help="Checkpoint folder to continue from.")
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
