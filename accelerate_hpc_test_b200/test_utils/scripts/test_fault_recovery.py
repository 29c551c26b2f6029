"""Elastic recovery drill: train N steps with a checkpoint after every step; a rank is killed at a given step by
``ACCELERATE_B200_FAULT_INJECT``; torchelastic (``accelerate launch --max_restarts 1``) restarts the group, the script finds the
newest checkpoint, resumes and must end with exactly the parameters of an uninterrupted run (the reference
leaves resume to the user and only has a toy ``tough_nut_function``, test_utils/scripts/test_notebook.py:34-43)."""

import json
import os
import sys

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.test_utils.training import RegressionDataset, RegressionModel
from accelerate_hpc_test_b200.utils import set_seed


def main():
    out_dir = sys.argv[1]
    total_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    accelerator = Accelerator(cpu=True)
    set_seed(0)
    model = RegressionModel(a=0.0, b=0.0)
    optimizer = torch.optim.SGD(model.parameters(), lr=0.1)
    data = RegressionDataset(length=total_steps * 4 * accelerator.num_processes, seed=1)
    loader = torch.utils.data.DataLoader(data, batch_size=4)
    model, optimizer, loader = accelerator.prepare(model, optimizer, loader)

    done = 0
    # a checkpoint counts only once every rank has finished writing it (marker file written after the barrier)
    ckpts = sorted((d for d in os.listdir(out_dir) if d.startswith("step_") and os.path.exists(os.path.join(out_dir, d, "COMPLETE"))),
                   key=lambda d: int(d.split("_")[1])) if os.path.isdir(out_dir) else []
    if ckpts:
        accelerator.load_state(os.path.join(out_dir, ckpts[-1]))
        done = int(ckpts[-1].split("_")[1])
        accelerator.print(f"attempt {os.environ.get('TORCHELASTIC_RESTART_COUNT', '0')}: resumed after step {done}")
    active = accelerator.skip_first_batches(loader, done) if done else loader
    for step, batch in enumerate(active, start=done + 1):
        loss = torch.nn.functional.mse_loss(model(batch["x"]), batch["y"])
        accelerator.backward(loss)
        optimizer.step()  # <- the injected fault fires here (optimizer-step counter is per attempt, see below)
        optimizer.zero_grad()
        accelerator.save_state(os.path.join(out_dir, f"step_{step}"))
        accelerator.wait_for_everyone()
        if accelerator.is_main_process:
            open(os.path.join(out_dir, f"step_{step}", "COMPLETE"), "w").close()
    if accelerator.is_main_process:
        m = accelerator.unwrap_model(model)
        with open(os.path.join(out_dir, "final.json"), "w") as f:
            json.dump({"a": m.a.item(), "b": m.b.item(), "attempt": int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0"))}, f)
    accelerator.end_training()


if __name__ == "__main__":
    main()
