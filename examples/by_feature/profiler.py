"""Profiling a few steps with ``accelerator.profile`` (``ProfileKwargs`` -> torch.profiler; one chrome trace per rank in
``output_trace_dir``; reference examples/by_feature/profiler.py).

    python examples/by_feature/profiler.py --cpu --model tiny --record_shapes --output_trace_dir /tmp/trace
"""

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator
# New Code #
from accelerate_hpc_test_b200.utils import ProfileKwargs


def training_function(config, args):
    # New Code #
    profile_kwargs = ProfileKwargs(
        activities=["cpu"] if args.cpu else ["cpu", "cuda"], record_shapes=args.record_shapes, profile_memory=args.profile_memory,
        with_flops=args.with_flops, output_trace_dir=args.output_trace_dir,
    )
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, kwargs_handlers=[profile_kwargs])
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args))
    for epoch in range(int(config["num_epochs"])):
        model.train()
        # New Code #
        with accelerator.profile() as prof:
            for step, batch in enumerate(train_dataloader):
                accelerator.backward(model(**batch).loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
                if step >= args.profile_steps:
                    break
        # New Code #
        accelerator.print(prof.key_averages().table(sort_by="self_cpu_time_total" if args.cpu else "self_cuda_time_total", row_limit=10))
        accelerator.print(f"epoch {epoch}:", evaluate(accelerator, model, eval_dataloader))
    accelerator.end_training()


def main():
    parser = base_parser()
    parser.add_argument("--record_shapes", action="store_true", help="Record the shapes of the operator inputs.")
    parser.add_argument("--profile_memory", action="store_true", help="Track tensor memory allocation / deallocation.")
    parser.add_argument("--with_flops", action="store_true", help="Estimate the FLOPS of the profiled operators.")
    parser.add_argument("--output_trace_dir", type=str, default=None, help="Write one chrome trace per rank into this folder.")
    parser.add_argument("--profile_steps", type=int, default=3)
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
