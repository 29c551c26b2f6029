"""Exact metrics with several processes: the last batch of a sharded evaluation loader is padded with repeated
samples; ``gather_for_metrics`` drops them, the manual version below shows what it does (reference
examples/by_feature/multi_process_metrics.py).

    accelerate launch --num_processes 2 --cpu examples/by_feature/multi_process_metrics.py --model tiny --eval_size 41
"""

import torch
from _common import base_parser, build, default_config

from accelerate_hpc_test_b200 import Accelerator


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args))
    for epoch in range(int(config["num_epochs"])):
        model.train()
        for batch in train_dataloader:
            accelerator.backward(model(**batch).loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
        model.eval()
        # New Code #
        samples_seen = correct = 0
        for step, batch in enumerate(eval_dataloader):
            with torch.no_grad():
                predictions = model(**batch).logits.argmax(dim=-1)
            predictions, references = accelerator.gather((predictions, batch["labels"]))
            if accelerator.use_distributed and step == len(eval_dataloader) - 1:
                # the last gathered batch contains duplicates: keep only what is left of the dataset
                keep = len(eval_dataloader.dataset) - samples_seen
                predictions, references = predictions[:keep], references[:keep]
            samples_seen += references.shape[0]
            correct += int((predictions == references).sum())
        assert samples_seen == len(eval_dataloader.dataset), (samples_seen, len(eval_dataloader.dataset))
        accelerator.print(f"epoch {epoch}:", {"accuracy": correct / samples_seen, "eval_samples": samples_seen})
    accelerator.end_training()


def main():
    args = base_parser().parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
