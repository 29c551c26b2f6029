"""Gradient-communication hooks under data parallelism: whatever goes over the wire (bf16 / fp16 compressed gradients, low-rank PowerSGD
factors with or without a compression wrapper), every rank must end a backward pass holding the same gradients, and — for the
lossless-enough hooks — the gradients of the un-hooked run.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.test_ddp_comm_hook
"""

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
from accelerate_hpc_test_b200.utils import DDPCommunicationHookType, DistributedDataParallelKwargs, set_seed
from accelerate_hpc_test_b200.utils.operations import gather


class TwoLayers(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.first = torch.nn.Linear(16, 32)
        self.second = torch.nn.Linear(32, 4)

    def forward(self, x):
        return self.second(torch.tanh(self.first(x)))


def gradients_with(hook, wrapper=DDPCommunicationHookType.NO, state_option=None):
    AcceleratorState._reset_state(True)
    GradientState._reset_state()
    handler = DistributedDataParallelKwargs(comm_hook=hook, comm_wrapper=wrapper, comm_state_option=state_option or {})
    accelerator = Accelerator(kwargs_handlers=[handler])
    set_seed(11)
    model = accelerator.prepare(TwoLayers().to(accelerator.device))
    # every rank its own inputs: the reduced gradient is the mean over ranks
    inputs = torch.randn(8, 16, generator=torch.Generator().manual_seed(100 + accelerator.process_index)).to(accelerator.device)
    model(inputs).pow(2).mean().backward()
    flat = torch.cat([p.grad.reshape(-1).float() for p in model.parameters()])
    everyone = gather(flat[None])
    assert torch.allclose(everyone, everyone[:1].expand_as(everyone), atol=1e-6), f"ranks disagree after the {hook} hook"
    return accelerator, flat.cpu()


def main():
    accelerator, plain = gradients_with(DDPCommunicationHookType.NO)
    if accelerator.num_processes < 2:
        raise SystemExit("hooks act on the gradient all-reduce: launch with at least two processes")
    cases = [
        (DDPCommunicationHookType.FP16, DDPCommunicationHookType.NO, None, 2e-3),
        (DDPCommunicationHookType.BF16, DDPCommunicationHookType.NO, None, 2e-2),
        (DDPCommunicationHookType.POWER_SGD, DDPCommunicationHookType.NO, {"matrix_approximation_rank": 2, "start_powerSGD_iter": 2}, 1e-5),
        (DDPCommunicationHookType.POWER_SGD, DDPCommunicationHookType.FP16, {"matrix_approximation_rank": 2, "start_powerSGD_iter": 2}, 2e-3),
        (DDPCommunicationHookType.BATCHED_POWER_SGD, DDPCommunicationHookType.BF16, {"matrix_approximation_rank": 1, "start_powerSGD_iter": 2}, 2e-2),
    ]
    for hook, wrapper, option, tolerance in cases:
        accelerator, hooked = gradients_with(hook, wrapper, option)
        # (PowerSGD runs plain all-reduce until `start_powerSGD_iter`, so its first backward is exact up to the wrapper's rounding)
        worst = (hooked - plain).abs().max().item() / plain.abs().max().item()
        assert worst < tolerance, (str(hook), str(wrapper), worst)
        accelerator.print(f"{hook} (wrapper {wrapper}): ranks agree, relative deviation from the un-hooked gradients {worst:.1e}")
    accelerator.print("test_ddp_comm_hook: all passed")


if __name__ == "__main__":
    main()
