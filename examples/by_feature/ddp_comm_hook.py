"""Communication hooks for data parallelism: ``DistributedDataParallelKwargs(comm_hook=...)`` changes what goes over the wire when
gradients are all-reduced.  ``bf16`` / ``fp16`` halve the bytes of fp32 gradients; the PowerSGD hooks (low-rank compression) belong to
torch DDP — the native engine's all-reduce runs at NVLink speed inside a node and refuses them, so they need ``ACCELERATE_B200_DDP=0``
(reference examples/by_feature/ddp_comm_hook.py).

    accelerate launch --num_processes 2 --cpu examples/by_feature/ddp_comm_hook.py --model tiny --ddp_comm_hook bf16
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator

# New Code #
from accelerate_hpc_test_b200.utils import DDPCommunicationHookType, DistributedDataParallelKwargs


def training_function(config, args):
    # New Code #
    hook = {"no": DDPCommunicationHookType.NO, "fp16": DDPCommunicationHookType.FP16, "bf16": DDPCommunicationHookType.BF16,
            "power_sgd": DDPCommunicationHookType.POWER_SGD, "batched_power_sgd": DDPCommunicationHookType.BATCHED_POWER_SGD}[args.ddp_comm_hook]
    ddp_kwargs = DistributedDataParallelKwargs(comm_hook=hook)
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, kwargs_handlers=[ddp_kwargs])
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = build(accelerator, config, args)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )
    for epoch in range(int(config["num_epochs"])):
        model.train()
        for batch in train_dataloader:
            # the hook acts inside backward, when a gradient bucket is complete and its all-reduce is launched
            accelerator.backward(model(**batch).loss)
            optimizer.step()
            lr_scheduler.step()
            optimizer.zero_grad()
        accelerator.print(f"epoch {epoch}:", evaluate(accelerator, model, eval_dataloader))
    accelerator.end_training()


def main():
    parser = base_parser()
    # New Code #
    parser.add_argument("--ddp_comm_hook", type=str, default="no", choices=["no", "fp16", "bf16", "power_sgd", "batched_power_sgd"],
                        help="DDP Communication hook to use. Choose between `no`, `fp16`, `bf16`, `power_sgd`, and `batched_power_sgd`.")
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
