"""Local SGD: every rank takes ``local_sgd_steps`` optimizer steps on its own, then the parameters are averaged
(``LocalSGD`` context, reference examples/by_feature/local_sgd.py).  Cuts the gradient traffic by that factor.

    accelerate launch --num_processes 2 --cpu examples/by_feature/local_sgd.py --model tiny --local_sgd_steps 4
"""

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator
# New Code #
from accelerate_hpc_test_b200.local_sgd import LocalSGD


def training_function(config, args):
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, gradient_accumulation_steps=args.gradient_accumulation_steps)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(*build(accelerator, config, args))
    for epoch in range(int(config["num_epochs"])):
        model.train()
        # New Code #
        with LocalSGD(accelerator=accelerator, model=model, local_sgd_steps=args.local_sgd_steps, enabled=args.local_sgd_steps is not None) as local_sgd:
            for batch in train_dataloader:
                with accelerator.accumulate(model):
                    accelerator.backward(model(**batch).loss)
                    optimizer.step()
                    lr_scheduler.step()
                    optimizer.zero_grad()
                    # New Code #: counts the step and averages the parameters every `local_sgd_steps`
                    local_sgd.step()
        accelerator.print(f"epoch {epoch}:", evaluate(accelerator, model, eval_dataloader))
    accelerator.end_training()


def main():
    parser = base_parser()
    parser.add_argument("--gradient_accumulation_steps", type=int, default=1)
    parser.add_argument("--local_sgd_steps", type=int, default=8, help="Number of local SGD steps or None to disable local SGD")
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
