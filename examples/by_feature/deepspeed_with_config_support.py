"""Training driven by a DeepSpeed configuration file.  Scripts written this way describe the optimizer and the schedule in the JSON
config and pass placeholders (``DummyOptim`` / ``DummyScheduler``) to ``prepare``.  DeepSpeed itself is not part of this framework:
the same configuration is executed by the native engines — ZeRO stage 3 is the full-shard engine, stage 1/2 shard gradients and
optimizer state, ``"auto"`` entries are filled from the script's arguments — so such scripts run unchanged
(reference examples/by_feature/deepspeed_with_config_support.py; templates in examples/deepspeed_config_templates/).

    accelerate launch --num_processes 2 --cpu examples/by_feature/deepspeed_with_config_support.py --model tiny \
        --deepspeed_config examples/deepspeed_config_templates/zero_stage3_config.json
"""

import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch

from _common import base_parser, default_config, evaluate, get_dataloaders

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification

# New Code #
from accelerate_hpc_test_b200.utils import DeepSpeedPlugin, DummyOptim, DummyScheduler, set_seed


def training_function(config, args):
    # New Code #
    plugin = DeepSpeedPlugin(hf_ds_config=args.deepspeed_config, gradient_accumulation_steps=args.gradient_accumulation_steps)
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, deepspeed_plugin=plugin,
                              gradient_accumulation_steps=args.gradient_accumulation_steps)
    set_seed(int(config["seed"]))
    cfg = BertConfig.tiny() if args.model == "tiny" else BertConfig.base_cased()
    train_dataloader, eval_dataloader = get_dataloaders(accelerator, int(config["batch_size"]), args.train_size, args.eval_size, cfg.vocab_size)
    model = BertForSequenceClassification(cfg).to(accelerator.device)
    num_update_steps = math.ceil(len(train_dataloader) / args.gradient_accumulation_steps) * int(config["num_epochs"])

    # New Code #
    # the config file owns the optimizer / scheduler when it has those sections: hand `prepare` placeholders carrying the values
    # its "auto" entries should take; without the sections, ordinary objects are created here
    ds_config = plugin.deepspeed_config
    if "optimizer" in ds_config:
        optimizer = DummyOptim(model.parameters(), lr=config["lr"], weight_decay=args.weight_decay)
    else:
        optimizer = torch.optim.AdamW(model.parameters(), lr=config["lr"], weight_decay=args.weight_decay)
    if "scheduler" in ds_config:
        lr_scheduler = DummyScheduler(optimizer, total_num_steps=num_update_steps, warmup_num_steps=args.num_warmup_steps)
    else:
        lr_scheduler = torch.optim.lr_scheduler.LambdaLR(optimizer, lambda step: max(0.0, 1 - step / max(1, num_update_steps)))

    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )
    accelerator.print(f"engine: {type(model).__name__}, optimizer: {type(optimizer.optimizer).__name__}, "
                      f"ZeRO stage {ds_config.get('zero_optimization', {}).get('stage', plugin.zero_stage)}")
    for epoch in range(int(config["num_epochs"])):
        model.train()
        for batch in train_dataloader:
            with accelerator.accumulate(model):
                accelerator.backward(model(**batch).loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
        accelerator.print(f"epoch {epoch}:", evaluate(accelerator, model, eval_dataloader))
    accelerator.end_training()


def main():
    parser = base_parser("Training with a DeepSpeed configuration file on the native engines.")
    here = os.path.dirname(os.path.abspath(__file__))
    parser.add_argument("--deepspeed_config", type=str, default=os.path.join(here, "..", "deepspeed_config_templates", "zero_stage2_config.json"))
    parser.add_argument("--gradient_accumulation_steps", type=int, default=1)
    parser.add_argument("--weight_decay", type=float, default=0.0)
    parser.add_argument("--num_warmup_steps", type=int, default=2)
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
