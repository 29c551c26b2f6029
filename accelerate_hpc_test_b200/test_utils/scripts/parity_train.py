"""One training script, two frameworks: ``python parity_train.py <package> <out.json> [num_processes]`` trains a small
classifier on fixed synthetic data through ``<package>.Accelerator`` (``accelerate`` = the unmodified reference installed under
baseline/_ref, ``accelerate_hpc_test_b200`` = this framework) and dumps the per-step losses of rank 0 plus the final weights'
checksum.  tests/test_reference_parity.py runs both and expects identical numbers."""

import importlib
import json
import sys


def train(pkg_name, out_path):
    import torch
    from torch.utils.data import DataLoader, TensorDataset

    pkg = importlib.import_module(pkg_name)
    accelerator = pkg.Accelerator(cpu=True, gradient_accumulation_steps=2)
    importlib.import_module(pkg_name + ".utils").set_seed(1234)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(96, 16, generator=g)
    y = (x[:, :4].sum(-1) > 0).long()
    model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 2))
    optimizer = torch.optim.AdamW(model.parameters(), lr=1e-2)
    scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=2, gamma=0.9)
    loader = DataLoader(TensorDataset(x, y), batch_size=8, shuffle=True)
    model, optimizer, loader, scheduler = accelerator.prepare(model, optimizer, loader, scheduler)
    losses = []
    for epoch in range(2):
        for xb, yb in loader:
            with accelerator.accumulate(model):
                loss = torch.nn.functional.cross_entropy(model(xb), yb)
                accelerator.backward(loss)
                accelerator.clip_grad_norm_(model.parameters(), 1.0)
                optimizer.step()
                scheduler.step()
                optimizer.zero_grad()
            losses.append(accelerator.gather(loss.detach().reshape(1)).mean().item())
    state = accelerator.unwrap_model(model).state_dict()
    checksum = float(sum(v.double().abs().sum() for v in state.values()))
    if accelerator.is_main_process:
        with open(out_path, "w") as f:
            json.dump({"losses": losses, "checksum": checksum, "lr": optimizer.param_groups[0]["lr"],
                       "num_processes": accelerator.num_processes}, f)
    accelerator.wait_for_everyone()


def main():
    pkg_name, out_path = sys.argv[1], sys.argv[2]
    nproc = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    if nproc == 1:
        train(pkg_name, out_path)
    else:
        launcher = importlib.import_module(pkg_name + ".launchers").debug_launcher
        launcher(train, args=(pkg_name, out_path), num_processes=nproc)


if __name__ == "__main__":
    main()
