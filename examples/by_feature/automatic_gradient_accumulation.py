"""Batch-size search and gradient accumulation together: when the batch size has to shrink to fit memory, the number
of accumulation steps grows so that the *observed* batch size stays what was asked for (reference
examples/by_feature/automatic_gradient_accumulation.py).

    python examples/by_feature/automatic_gradient_accumulation.py --cpu --model tiny
"""

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.utils import find_executable_batch_size


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    observed_batch_size = int(config["batch_size"])

    # New Code #
    @find_executable_batch_size(starting_batch_size=observed_batch_size)
    def inner_training_loop(batch_size):
        nonlocal accelerator
        # New Code #: keep batch_size * accumulation_steps == observed_batch_size
        accelerator.gradient_accumulation_steps = max(1, observed_batch_size // batch_size)
        accelerator.free_memory()
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args, batch_size))
        for epoch in range(int(config["num_epochs"])):
            model.train()
            for batch in train_dataloader:
                with accelerator.accumulate(model):
                    accelerator.backward(model(**batch).loss)
                    optimizer.step()
                    lr_scheduler.step()
                    optimizer.zero_grad()
            accelerator.print(f"epoch {epoch} (batch {batch_size} x {accelerator.gradient_accumulation_steps}):",
                              evaluate(accelerator, model, eval_dataloader))

    inner_training_loop()
    accelerator.end_training()


def main():
    args = base_parser().parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
