"""Interrupt-and-resume must be invisible: training N epochs in one go and training with a stop after every epoch — new process state,
``load_state`` from the last ``epoch_k`` folder — must end with the same weights, optimizer state, learning rate and metrics.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_checkpointing
"""

import argparse
import json
import os
import shutil
import tempfile

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification, pad_collate, synthetic_mrpc
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.utils import DataLoaderConfiguration, set_seed
from accelerate_hpc_test_b200.utils.operations import broadcast_object_list


def build(accelerator, epochs):
    set_seed(42)
    cfg = BertConfig.tiny()
    train = DataLoader(synthetic_mrpc(64, vocab_size=cfg.vocab_size, seed=0), shuffle=True, collate_fn=pad_collate, batch_size=8, drop_last=True)
    evald = DataLoader(synthetic_mrpc(41, vocab_size=cfg.vocab_size, seed=1), shuffle=False, collate_fn=pad_collate, batch_size=8)
    model = BertForSequenceClassification(cfg)
    optimizer = torch.optim.AdamW(model.parameters(), lr=1e-3)
    total = (len(train) // accelerator.num_processes) * epochs
    scheduler = torch.optim.lr_scheduler.LambdaLR(optimizer, lambda step: max(0.0, 1 - step / max(1, total)))
    return accelerator.prepare(model, optimizer, train, evald, scheduler)


def evaluate(accelerator, model, evald):
    model.eval()
    correct = total = 0
    for batch in evald:
        with torch.no_grad():
            predictions = model(**batch).logits.argmax(-1)
        predictions, references = accelerator.gather_for_metrics((predictions, batch["labels"]))
        correct += int((predictions == references).sum())
        total += len(references)
    return correct / total


def run(epochs, output_dir, resume_from=None, stop_after=None):
    """Train epochs ``[resumed epoch + 1, stop_after]``; returns a summary of where training stands."""
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    accelerator = Accelerator(dataloader_config=DataLoaderConfiguration(use_seedable_sampler=True))
    model, optimizer, train, evald, scheduler = build(accelerator, epochs)
    first = 0
    if resume_from is not None:
        accelerator.load_state(resume_from)
        first = int(os.path.basename(resume_from).split("_")[1]) + 1
        with open(os.path.join(resume_from, "summary.json")) as f:
            before = json.load(f)
        # the loaded state reproduces what was recorded when it was saved
        assert abs(before["lr"] - optimizer.param_groups[0]["lr"]) < 1e-12, (before["lr"], optimizer.param_groups[0]["lr"])
        assert abs(before["accuracy"] - evaluate(accelerator, model, evald)) < 1e-9
    last = epochs - 1 if stop_after is None else stop_after
    summary = {}
    for epoch in range(first, last + 1):
        model.train()
        train.set_epoch(epoch)
        for batch in train:
            accelerator.backward(model(**batch).loss)
            optimizer.step()
            scheduler.step()
            optimizer.zero_grad()
        folder = os.path.join(output_dir, f"epoch_{epoch}")
        accelerator.save_state(folder)
        summary = {"epoch": epoch, "lr": optimizer.param_groups[0]["lr"], "accuracy": evaluate(accelerator, model, evald),
                   "step": scheduler.scheduler.last_epoch}
        if accelerator.is_main_process:
            with open(os.path.join(folder, "summary.json"), "w") as f:
                json.dump(summary, f)
        accelerator.wait_for_everyone()
    weights = {k: v.detach().float().cpu() for k, v in accelerator.unwrap_model(model).state_dict().items()}
    return summary, weights


def main():
    parser = argparse.ArgumentParser()
    parser.add_argument("--num_epochs", type=int, default=3)
    args = parser.parse_args()
    from accelerate_hpc_test_b200 import PartialState

    holder = [tempfile.mkdtemp() if PartialState().is_main_process else None]
    broadcast_object_list(holder)
    root = holder[0]
    straight, straight_weights = run(args.num_epochs, os.path.join(root, "straight"))
    stop_go, stop_go_weights = None, None
    for epoch in range(args.num_epochs):
        resume = os.path.join(root, "stop_go", f"epoch_{epoch - 1}") if epoch > 0 else None
        stop_go, stop_go_weights = run(args.num_epochs, os.path.join(root, "stop_go"), resume_from=resume, stop_after=epoch)
    assert stop_go["epoch"] == straight["epoch"] and stop_go["step"] == straight["step"], (stop_go, straight)
    assert abs(stop_go["lr"] - straight["lr"]) < 1e-12 and abs(stop_go["accuracy"] - straight["accuracy"]) < 1e-9, (stop_go, straight)
    worst = max((straight_weights[k] - stop_go_weights[k]).abs().max().item() for k in straight_weights)
    assert worst < 1e-6, f"resumed training diverged from uninterrupted training: max weight difference {worst}"
    state = PartialState()
    state.wait_for_everyone()
    if state.is_main_process:
        shutil.rmtree(root, ignore_errors=True)
        print(f"stop-and-resume after every epoch reproduces uninterrupted training (max weight difference {worst:.1e})")
        print("test_checkpointing: all passed")


if __name__ == "__main__":
    main()
