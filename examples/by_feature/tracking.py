"""Experiment tracking: ``Accelerator(log_with=...)``, ``init_trackers``, ``log``, ``end_training`` (reference
examples/by_feature/tracking.py).  ``jsonl`` is the built-in dependency-free tracker; ``all`` uses every installed one.

    python examples/by_feature/tracking.py --cpu --model tiny --with_tracking --project_dir /tmp/logs
"""

import os

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator


def training_function(config, args):
    # New Code #
    if args.with_tracking:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, log_with=args.log_with, project_dir=args.project_dir)
        accelerator.init_trackers(os.path.splitext(os.path.basename(__file__))[0], config)
    else:
        accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args))
    for epoch in range(int(config["num_epochs"])):
        model.train()
        total_loss = 0.0
        for batch in train_dataloader:
            loss = model(**batch).loss
            # New Code #
            if args.with_tracking:
                total_loss += loss.detach().float()
            accelerator.backward(loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
        metrics = evaluate(accelerator, model, eval_dataloader)
        accelerator.print(f"epoch {epoch}:", metrics)
        # New Code #
        if args.with_tracking:
            accelerator.log({"accuracy": metrics["accuracy"], "train_loss": float(total_loss) / len(train_dataloader), "epoch": epoch}, step=epoch)
    # New Code #: flushes and closes every tracker
    accelerator.end_training()


def main():
    parser = base_parser()
    parser.add_argument("--with_tracking", action="store_true", help="Whether to log to experiment trackers.")
    parser.add_argument("--log_with", type=str, default="jsonl", help="`all`, or a tracker name (tensorboard, wandb, jsonl, ...).")
    parser.add_argument("--project_dir", type=str, default="logs", help="Location of logs and project directories.")
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
