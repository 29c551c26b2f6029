"""Gradient-synchronisation semantics of ``no_sync`` / ``accumulate`` under the data-parallel wrapper of the active
back-end (torch DDP on gloo, the native engine on CUDA); run by ``accelerate test`` and the grad-sync tests.

Every check trains three copies of the two-parameter regression model side by side:

* ``wrapped``  — prepared by the Accelerator, fed this rank's slice of the batch;
* ``local``    — plain copy fed the same slice (what the gradients look like while nothing is synchronised);
* ``whole``    — plain copy fed the batch of all ranks (what they look like after the all-reduce).

While gradients are being accumulated ``wrapped`` must match ``local`` exactly; on a synchronising step it must match
``whole``.  Comparing against both expectations (rather than asserting "differs from the other ranks") keeps the script
deterministic.  Behaviour covered: reference ``test_utils/scripts/test_sync.py:29-152``.
"""

from __future__ import annotations

from copy import deepcopy

import torch
import torch.nn.functional as F
from torch.optim import AdamW
from torch.optim.lr_scheduler import LambdaLR
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.test_utils.training import RegressionDataset, RegressionModel
from accelerate_hpc_test_b200.utils import DataLoaderConfiguration, DistributedType, GradientAccumulationPlugin, set_seed

TOL = dict(rtol=1e-4, atol=1e-6)


def _reset():
    GradientState._reset_state()
    AcceleratorState._reset_state(True)


def _grads(model):
    return [p.grad.detach().clone() for p in model.parameters() if p.requires_grad]


def _same(xs, ys):
    return all(torch.allclose(x.cpu(), y.cpu(), **TOL) for x, y in zip(xs, ys))


def _expect(model, wanted, what: str, step: int):
    got = _grads(model)
    assert _same(got, wanted), f"step {step}: wrapped gradients {got} are not the {what} gradients {wanted}"


def _plain_backward(model, x, y, divide_by: int = 1):
    loss = F.mse_loss(model(x), y.to(x.device)) / divide_by
    loss.backward()


class Trio:
    """wrapped / local / whole models on one data stream."""

    def __init__(self, accelerator: Accelerator, with_optimizer: bool = False):
        set_seed(42)
        self.acc = accelerator
        base = RegressionModel()
        self.wrapped, self.local, self.whole = deepcopy(base), deepcopy(base).to(accelerator.device), deepcopy(base).to(accelerator.device)
        loader = DataLoader(RegressionDataset(length=80, seed=7), batch_size=16)
        self.opts = self.scheds = None
        if with_optimizer:
            ramp = lambda epoch: epoch**0.65  # noqa: E731
            opts = [AdamW(m.parameters(), lr=1e-3) for m in (self.wrapped, self.whole)]
            scheds = [LambdaLR(o, lr_lambda=ramp) for o in opts]
            self.wrapped, opts[0], scheds[0], self.loader = accelerator.prepare(self.wrapped, opts[0], scheds[0], loader)
            self.opts, self.scheds = opts, scheds
        else:
            self.wrapped, self.loader = accelerator.prepare(self.wrapped, loader)

    def whole_batch(self, x, y):
        gx, gy = self.acc.gather((x, y))
        return gx.to(self.acc.device), gy.to(self.acc.device)

    def zero_plain(self):
        for m in (self.local, self.whole):
            m.zero_grad(set_to_none=True)


def check_no_sync(accelerator: Accelerator):
    """Every other step runs under ``no_sync``.  One process: always equal to ``whole``.  Several: a ``no_sync`` step
    leaves the local sum in place, the next synchronising backward averages everything accumulated since."""
    t = Trio(accelerator)
    x, y = next(iter(t.loader)).values()
    distributed = accelerator.num_processes > 1
    carried = None  # local gradient left behind by the previous no_sync step
    for step in range(4):
        t.zero_plain()
        t.wrapped.zero_grad(set_to_none=True) if carried is None else None
        gx, gy = t.whole_batch(x, y)
        _plain_backward(t.whole, gx, gy)
        _plain_backward(t.local, x, y)
        skip = step % 2 == 0
        if skip:
            with accelerator.no_sync(t.wrapped):
                accelerator.backward(F.mse_loss(t.wrapped(x), y))
        else:
            accelerator.backward(F.mse_loss(t.wrapped(x), y))
        if not distributed:
            _expect(t.wrapped, _grads(t.whole), "single-process", step)
            carried = None
        elif skip:
            _expect(t.wrapped, _grads(t.local), "local (unsynchronised)", step)
            carried = True
        else:
            # both micro-steps used the same weights and the loss is a mean over the global batch of each step
            want = [2 * g for g in _grads(t.whole)]
            _expect(t.wrapped, want, "averaged sum of two steps", step)
            carried = None
        torch.manual_seed(1337 + step)
        perm = torch.randperm(len(x))
        x, y = x[perm], y[perm]
    _reset()


def _accumulating_accelerator(split_batches, dispatch_batches, sync_each_batch):
    return Accelerator(
        dataloader_config=DataLoaderConfiguration(split_batches=split_batches, dispatch_batches=dispatch_batches),
        gradient_accumulation_plugin=GradientAccumulationPlugin(num_steps=2, sync_each_batch=sync_each_batch),
    )


def check_accumulate(split_batches=False, dispatch_batches=False, sync_each_batch=False):
    """``accumulate`` with 2 steps: gradients are global after every second batch, after the last batch of the loader, or
    always with ``sync_each_batch``; otherwise they are this rank's own."""
    accelerator = _accumulating_accelerator(split_batches, dispatch_batches, sync_each_batch)
    t = Trio(accelerator)
    n = len(t.loader)
    for step, batch in enumerate(t.loader):
        x, y = batch.values()
        gx, gy = t.whole_batch(x, y)
        # the plain copies accumulate without ever being cleared, like the wrapped model between optimizer steps
        _plain_backward(t.whole, gx, gy, divide_by=2)
        _plain_backward(t.local, x, y, divide_by=2)
        with accelerator.accumulate(t.wrapped):
            accelerator.backward(F.mse_loss(t.wrapped(x), y))
        boundary = (step + 1) % 2 == 0 or step == n - 1
        if accelerator.num_processes == 1 or boundary or sync_each_batch:
            _expect(t.wrapped, _grads(t.whole), "global", step)
        else:
            assert not accelerator.sync_gradients
            _expect(t.wrapped, _grads(t.local), "local (unsynchronised)", step)
        if boundary:
            t.zero_plain()
            t.wrapped.zero_grad(set_to_none=True)
    _reset()


def check_accumulate_with_optimizer_and_scheduler(split_batches=False, dispatch_batches=False):
    """Prepared optimizer + scheduler inside ``accumulate``: they act only on boundary steps, where the learning rate and the
    parameters must equal those of a plain loop fed the global batch."""
    accelerator = _accumulating_accelerator(split_batches, dispatch_batches, False)
    t = Trio(accelerator, with_optimizer=True)
    (w_opt, p_opt), (w_sched, p_sched) = t.opts, t.scheds
    n = len(t.loader)
    for step, batch in enumerate(t.loader):
        x, y = batch.values()
        gx, gy = t.whole_batch(x, y)
        t.whole.train()
        t.wrapped.train()
        _plain_backward(t.whole, gx, gy, divide_by=2)
        boundary = (step + 1) % 2 == 0 or step == n - 1
        if boundary:
            p_opt.step()
            for _ in range(1 if split_batches else accelerator.num_processes):
                p_sched.step()
            p_opt.zero_grad()
        with accelerator.accumulate(t.wrapped):
            accelerator.backward(F.mse_loss(t.wrapped(x), y))
            w_opt.step()
            w_sched.step()
            w_opt.zero_grad()
        lr_plain, lr_wrapped = p_opt.param_groups[0]["lr"], w_opt.param_groups[0]["lr"]
        assert lr_plain == lr_wrapped, f"step {step}: learning rates diverged ({lr_plain} vs {lr_wrapped})"
        if boundary:
            got = [p.detach().cpu() for p in accelerator.unwrap_model(t.wrapped).parameters()]
            want = [p.detach().cpu() for p in t.whole.parameters()]
            assert all(torch.allclose(a, b, rtol=1e-3, atol=1e-5) for a, b in zip(got, want)), f"step {step}: {got} vs {want}"
    _reset()


def check_nested_dataloaders():
    """``GradientState`` tracks the innermost active prepared dataloader and its ``end_of_dataloader`` flag."""
    accelerator = Accelerator()
    outer = DataLoader(RegressionDataset(length=80), batch_size=16)
    inner = DataLoader(RegressionDataset(length=96), batch_size=16)
    outer, inner = accelerator.prepare(outer, inner)
    gs = accelerator.gradient_state
    assert gs.active_dataloader is None
    last_outer = len(outer) - 1
    for i, _ in enumerate(outer):
        assert gs.active_dataloader is outer
        assert gs.end_of_dataloader == (i == last_outer)
        if i == 1:
            last_inner = len(inner) - 1
            for j, _ in enumerate(inner):
                assert gs.active_dataloader is inner
                assert gs.end_of_dataloader == (j == last_inner)
            assert gs.active_dataloader is outer
    assert gs.active_dataloader is None
    _reset()


def main():
    accelerator = Accelerator()
    chatty = accelerator.state.local_process_index == 0
    say = print if chatty else (lambda *a, **k: None)
    say("**Test `accumulate` gradient accumulation with dataloader break**")
    check_nested_dataloaders()  # resets the shared state: build a fresh Accelerator for the next checks
    accelerator = Accelerator()
    if accelerator.state.distributed_type in (DistributedType.NO, DistributedType.MULTI_GPU, DistributedType.MULTI_CPU):
        say("**Test `no_sync` context manager**")
        check_no_sync(accelerator)
    else:
        _reset()
    say("**Test gradient accumulation**")
    for split in (True, False):
        for dispatch in (True, False):
            for each in (True, False):
                check_accumulate(split, dispatch, each)
    say("**Test gradient accumulation with optimizer and scheduler**")
    for split in (True, False):
        for dispatch in (True, False):
            check_accumulate_with_optimizer_and_scheduler(split, dispatch)
    say("test_sync: all passed")


def _mp_fn(index):
    main()


if __name__ == "__main__":
    main()
