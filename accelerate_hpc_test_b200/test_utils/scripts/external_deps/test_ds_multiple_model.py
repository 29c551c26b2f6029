"""Several DeepSpeed-style configurations in one program (``deepspeed_plugins={"student": ..., "teacher": ...}``): the plugin that is
*selected* when a model is prepared decides how that model is run.  The classic use: a frozen teacher for inference and a trained
student, each with its own ZeRO stage.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_ds_multiple_model
"""

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.state import AcceleratorState
from accelerate_hpc_test_b200.utils import DeepSpeedPlugin, get_active_deepspeed_plugin, set_seed


def tiny_llama(device, seed):
    set_seed(seed)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    return cfg, LlamaForCausalLM(cfg, dtype=torch.bfloat16 if device.type == "cuda" else torch.float32, device=device)


def main():
    plugins = {"student": DeepSpeedPlugin(zero_stage=2, gradient_accumulation_steps=1),
               "teacher": DeepSpeedPlugin(zero_stage=3, gradient_accumulation_steps=1)}
    for plugin in plugins.values():  # no dataloader is prepared below, so the batch size is stated instead of left on "auto"
        plugin.deepspeed_config["train_micro_batch_size_per_gpu"] = 2
    accelerator = Accelerator(deepspeed_plugins=plugins)
    state = AcceleratorState()
    if accelerator.num_processes < 2:
        raise SystemExit("plugin selection only matters for a distributed job: launch with at least two processes")
    # the first plugin is the active one until another is selected
    assert get_active_deepspeed_plugin(state) is plugins["student"] and state.deepspeed_plugin.zero_stage == 2
    assert state.get_deepspeed_plugin("teacher") is plugins["teacher"]

    cfg, student = tiny_llama(accelerator.device, 0)
    optimizer = torch.optim.AdamW(student.parameters(), lr=1e-2)
    student, optimizer = accelerator.prepare(student, optimizer)

    state.select_deepspeed_plugin("teacher")
    assert get_active_deepspeed_plugin(state) is plugins["teacher"] and not plugins["student"].selected
    _, teacher = tiny_llama(accelerator.device, 1)
    teacher.requires_grad_(False)
    teacher.eval()

    state.select_deepspeed_plugin("student")
    assert get_active_deepspeed_plugin(state) is plugins["student"]
    losses = []
    for step in range(4):
        ids = torch.randint(0, cfg.vocab_size, (2, 32), generator=torch.Generator().manual_seed(step * 5 + accelerator.process_index)).to(accelerator.device)
        with torch.no_grad():
            target = teacher(input_ids=ids).logits.float().softmax(-1)
        logits = student(input_ids=ids).logits.float()
        loss = torch.nn.functional.kl_div(logits.log_softmax(-1), target, reduction="batchmean")  # distillation
        accelerator.backward(loss)
        optimizer.step()
        optimizer.zero_grad()
        losses.append(float(loss))
    assert losses[-1] < losses[0], losses
    try:
        state.select_deepspeed_plugin("nobody")
    except KeyError:
        pass
    else:
        raise AssertionError("selecting an unknown plugin must fail")
    accelerator.print(f"student (ZeRO-2 plugin) distilled from a frozen teacher (ZeRO-3 plugin): KL {losses[0]:.4f} -> {losses[-1]:.4f}")
    accelerator.print("test_ds_multiple_model: all passed")
    accelerator.end_training()


if __name__ == "__main__":
    main()
