"""Image classification with a small ResNet — the computer-vision counterpart of nlp_example.py, and the example of an *arbitrary*
``nn.Module`` (convolutions, batch norm) going through the same ``Accelerator`` as the native model families.

Counterpart of the reference's examples/cv_example.py (ResNet-50 on the Oxford-IIIT Pets images); there is no network or image
library dependency here, so the images are synthetic: every class has its own colour / stripe pattern under noise, which a convnet
learns in an epoch.  Runs on one CPU process, several (gloo), or on GPUs:

    python examples/cv_example.py --cpu --num_epochs 1 --train_size 256
    accelerate launch --num_processes 2 --cpu examples/cv_example.py --num_epochs 1
    accelerate launch --num_processes 8 --mixed_precision bf16 examples/cv_example.py
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))  # run from a source checkout

import torch
import torch.nn as nn
from torch.optim.lr_scheduler import OneCycleLR
from torch.utils.data import DataLoader, Dataset

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.utils import set_seed


class SyntheticPets(Dataset):
    """``num_classes`` kinds of images: a class fixes the stripe period, orientation and tint; samples add noise and a shift."""

    def __init__(self, size: int, image_size: int = 64, num_classes: int = 8, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(0, num_classes, (size,), generator=g)
        self.shifts = torch.randint(0, image_size, (size,), generator=g)
        self.noise_seeds = torch.randint(0, 2**31 - 1, (size,), generator=g)
        self.image_size, self.num_classes = image_size, num_classes

    def __len__(self):
        return len(self.labels)

    def __getitem__(self, idx):
        label = int(self.labels[idx])
        n = self.image_size
        axis = torch.arange(n, dtype=torch.float32) + float(self.shifts[idx])
        stripes = torch.sin(axis * (2 * torch.pi / (4 + 3 * (label // 2))))
        plane = stripes[None, :].expand(n, n) if label % 2 == 0 else stripes[:, None].expand(n, n)
        tint = torch.tensor([((label * 37) % 11) / 11.0, ((label * 53) % 7) / 7.0, ((label * 17) % 5) / 5.0])
        image = plane[None] * 0.5 + tint[:, None, None]
        noise = torch.randn(3, n, n, generator=torch.Generator().manual_seed(int(self.noise_seeds[idx])))
        return {"image": image + 0.3 * noise, "label": label}


def _block(cin, cout, stride):
    return nn.Sequential(nn.Conv2d(cin, cout, 3, stride, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True),
                         nn.Conv2d(cout, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout))


class SmallResNet(nn.Module):
    def __init__(self, num_classes: int, width: int = 32):
        super().__init__()
        self.stem = nn.Sequential(nn.Conv2d(3, width, 3, 1, 1, bias=False), nn.BatchNorm2d(width), nn.ReLU(inplace=True))
        self.stages = nn.ModuleList([_block(width, width, 1), _block(width, 2 * width, 2), _block(2 * width, 4 * width, 2)])
        self.shortcuts = nn.ModuleList([nn.Identity(), nn.Conv2d(width, 2 * width, 1, 2, bias=False), nn.Conv2d(2 * width, 4 * width, 1, 2, bias=False)])
        self.head = nn.Linear(4 * width, num_classes)

    def get_classifier(self):
        return self.head

    def forward(self, x):
        x = self.stem(x)
        for stage, shortcut in zip(self.stages, self.shortcuts):
            x = torch.relu(stage(x) + shortcut(x))
        return self.head(x.mean(dim=(2, 3)))


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    lr, num_epochs, seed, batch_size, image_size = config["lr"], int(config["num_epochs"]), int(config["seed"]), int(config["batch_size"]), config["image_size"]
    set_seed(seed)
    train_dataset = SyntheticPets(args.train_size, image_size, args.num_classes, seed=0)
    eval_dataset = SyntheticPets(args.eval_size, image_size, args.num_classes, seed=1)
    train_dataloader = DataLoader(train_dataset, shuffle=True, batch_size=batch_size, num_workers=0)
    eval_dataloader = DataLoader(eval_dataset, shuffle=False, batch_size=batch_size, num_workers=0)
    model = SmallResNet(args.num_classes).to(accelerator.device)
    # fine-tuning recipe of the original example: freeze the body, train the classifier
    if args.freeze_body:
        for param in model.parameters():
            param.requires_grad = False
        for param in model.get_classifier().parameters():
            param.requires_grad = True
    # inputs are normalised on the device, after the (asynchronous) host-to-device copy
    mean = torch.tensor([0.5, 0.5, 0.5])[None, :, None, None].to(accelerator.device)
    std = torch.tensor([0.6, 0.6, 0.6])[None, :, None, None].to(accelerator.device)
    optimizer = torch.optim.Adam(params=[p for p in model.parameters() if p.requires_grad], lr=lr / 25)
    lr_scheduler = OneCycleLR(optimizer=optimizer, max_lr=lr, epochs=num_epochs, steps_per_epoch=len(train_dataloader))
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )
    metrics = {}
    for epoch in range(num_epochs):
        model.train()
        for batch in train_dataloader:
            inputs = (batch["image"] - mean) / std
            loss = torch.nn.functional.cross_entropy(model(inputs), batch["label"])
            accelerator.backward(loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
        model.eval()
        accurate = num_elems = 0
        for batch in eval_dataloader:
            inputs = (batch["image"] - mean) / std
            with torch.no_grad():
                predictions = model(inputs).argmax(dim=-1)
            predictions, references = accelerator.gather_for_metrics((predictions, batch["label"]))
            accurate += int((predictions == references).sum())
            num_elems += int(references.numel())
        metrics = {"accuracy": accurate / max(1, num_elems), "eval_samples": num_elems}
        accelerator.print(f"epoch {epoch}: {100 * metrics['accuracy']:.2f}", metrics)
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
    parser.add_argument("--freeze_body", action="store_true", help="Train the classifier only.")
    args = parser.parse_args()
    config = {"lr": 3e-2, "num_epochs": args.num_epochs, "seed": 42, "batch_size": 64, "image_size": args.image_size}
    training_function(config, args)


if __name__ == "__main__":
    main()
