"""Schedule-free optimisers (Defazio et al.) need no learning-rate schedule, but they keep two sets of iterates and must be told which
one the model should hold: ``optimizer.train()`` before training, ``optimizer.eval()`` before evaluating or saving.  The prepared
optimizer forwards both calls (reference examples/by_feature/schedule_free.py).  The `schedulefree` package is used when it is
installed; otherwise a small AdamW-schedule-free implementation from this file stands in.

    python examples/by_feature/schedule_free.py --cpu --model tiny --num_epochs 1
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch

from _common import base_parser, default_config, evaluate, get_dataloaders

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification
from accelerate_hpc_test_b200.utils import set_seed


# New Code #
class AdamWScheduleFree(torch.optim.Optimizer):
    """Schedule-free AdamW: ``z`` takes the Adam steps, ``x`` is their running average, the gradient is taken at ``y`` — an
    interpolation of the two, which is what the parameters hold in train mode; eval mode swaps ``x`` in."""

    def __init__(self, params, lr=2e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, warmup_steps=0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, warmup_steps=warmup_steps,
                                      k=0, train_mode=False, weight_sum=0.0))

    def _swap(self, to_train: bool):
        for group in self.param_groups:
            if group["train_mode"] == to_train:
                continue
            beta1 = group["betas"][0]
            for p in group["params"]:
                z = self.state[p].get("z")
                if z is not None:
                    # y = (1 - beta1) z + beta1 x  <->  x = (y - (1 - beta1) z) / beta1
                    if to_train:
                        p.data.lerp_(z, 1 - beta1)
                    else:
                        p.data.lerp_(z, 1 - 1 / beta1)
            group["train_mode"] = to_train

    def train(self):
        self._swap(True)

    def eval(self):
        self._swap(False)

    @torch.no_grad()
    def step(self, closure=None):
        for group in self.param_groups:
            if not group["train_mode"]:
                raise RuntimeError("call optimizer.train() before stepping a schedule-free optimizer")
            beta1, beta2 = group["betas"]
            k = group["k"]
            warm = min(1.0, (k + 1) / max(1, group["warmup_steps"])) if group["warmup_steps"] else 1.0
            lr = group["lr"] * warm * (1 - beta2 ** (k + 1)) ** 0.5
            group["weight_sum"] += lr**2
            ckp1 = lr**2 / group["weight_sum"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                state = self.state[p]
                if not state:
                    state["z"] = p.detach().clone()
                    state["exp_avg_sq"] = torch.zeros_like(p)
                state["exp_avg_sq"].mul_(beta2).addcmul_(p.grad, p.grad, value=1 - beta2)
                direction = p.grad / state["exp_avg_sq"].sqrt().add_(group["eps"])
                if group["weight_decay"]:
                    direction = direction.add(p, alpha=group["weight_decay"])
                # y moves towards z (averaging) and along the step, z takes the full step
                p.lerp_(state["z"], ckp1).add_(direction, alpha=lr * (beta1 * (1 - ckp1) - 1))
                state["z"].sub_(direction, alpha=lr)
            group["k"] = k + 1


def make_optimizer(params, lr):
    try:
        import schedulefree

        return schedulefree.AdamWScheduleFree(params, lr=lr, warmup_steps=2)
    except ImportError:
        return AdamWScheduleFree(params, lr=lr, warmup_steps=2)


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision)
    set_seed(int(config["seed"]))
    cfg = BertConfig.tiny() if args.model == "tiny" else BertConfig.base_cased()
    train_dataloader, eval_dataloader = get_dataloaders(accelerator, int(config["batch_size"]), args.train_size, args.eval_size, cfg.vocab_size)
    model = BertForSequenceClassification(cfg).to(accelerator.device)
    # New Code #
    # no scheduler is created, prepared or stepped anywhere in this script
    optimizer = make_optimizer(model.parameters(), config["lr"])
    model, optimizer, train_dataloader, eval_dataloader = accelerator.prepare(model, optimizer, train_dataloader, eval_dataloader)
    for epoch in range(int(config["num_epochs"])):
        model.train()
        # New Code #
        optimizer.train()
        for batch in train_dataloader:
            accelerator.backward(model(**batch).loss)
            optimizer.step()
            optimizer.zero_grad()
        # New Code #
        # evaluation (and checkpoints) must see the averaged iterate
        optimizer.eval()
        accelerator.print(f"epoch {epoch}:", evaluate(accelerator, model, eval_dataloader))
    accelerator.end_training()


def main():
    args = base_parser().parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
