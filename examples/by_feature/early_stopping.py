"""Stopping all processes together: any rank may ``set_trigger()``, every rank sees it through ``check_trigger()`` (an
all-reduce of the flag) at the same step (reference examples/by_feature/early_stopping.py).

    accelerate launch --num_processes 2 --cpu examples/by_feature/early_stopping.py --model tiny
"""

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator


class EarlyStoppingCallback:
    """Trips once the loss has not improved by ``min_delta`` for ``patience`` consecutive checks."""

    def __init__(self, min_delta=0.0, patience=5):
        self.min_delta, self.patience = min_delta, patience
        self.counter, self.lowest_loss = 0, float("inf")

    def check_early_stopping(self, loss):
        if self.lowest_loss - loss > self.min_delta:
            self.lowest_loss, self.counter = loss, 0
            return False
        self.counter += 1
        return self.counter >= self.patience


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args))
    # New Code #
    callback = EarlyStoppingCallback(patience=args.patience)
    stopped = False
    for epoch in range(int(config["num_epochs"])):
        model.train()
        for batch in train_dataloader:
            loss = model(**batch).loss
            # New Code #
            # the ranks see different data, so only some of them may want to stop: raise the flag locally ...
            if callback.check_early_stopping(loss.item()):
                accelerator.set_trigger()
            accelerator.backward(loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
            # ... and let every rank leave the loop at the same step
            if accelerator.check_trigger():
                stopped = True
                break
        accelerator.print(f"epoch {epoch}{' (stopped early)' if stopped else ''}:", evaluate(accelerator, model, eval_dataloader))
        if stopped:
            break
    accelerator.end_training()


def main():
    parser = base_parser()
    parser.add_argument("--patience", type=int, default=5)
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
