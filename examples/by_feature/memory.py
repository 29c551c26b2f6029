"""Automatic batch-size search: ``find_executable_batch_size`` re-runs the inner function with a smaller batch size
whenever it dies with an out-of-memory error (reference examples/by_feature/memory.py).

    python examples/by_feature/memory.py --cpu --model tiny
"""

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator
# New Code #
from accelerate_hpc_test_b200.utils import find_executable_batch_size


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)

    # New Code #
    # everything that depends on the batch size lives inside the decorated function; its first argument is the batch
    # size the decorator is currently trying (multiplied by 0.9 after every OOM)
    @find_executable_batch_size(starting_batch_size=int(config["batch_size"]))
    def inner_training_loop(batch_size):
        nonlocal accelerator
        accelerator.free_memory()  # drop references to the objects of a failed attempt
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args, batch_size))
        for epoch in range(int(config["num_epochs"])):
            model.train()
            for batch in train_dataloader:
                accelerator.backward(model(**batch).loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
            accelerator.print(f"epoch {epoch} (batch size {batch_size}):", evaluate(accelerator, model, eval_dataloader))

    # New Code #
    inner_training_loop()
    accelerator.end_training()


def main():
    args = base_parser().parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
