"""A ZeRO stage-3 configuration: the model is partitioned (here: wrapped by the native full-shard engine), trains, and comes back
whole through ``get_state_dict`` — the 16-bit-weights-on-save path of ZeRO-3.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_zero3_integration
"""

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.utils import DeepSpeedPlugin, DistributedType, set_seed


def main():
    config = {
        "zero_optimization": {"stage": 3, "stage3_gather_16bit_weights_on_model_save": True},
        "train_batch_size": "auto", "train_micro_batch_size_per_gpu": "auto", "gradient_accumulation_steps": 1,
    }
    plugin = DeepSpeedPlugin(hf_ds_config=config, zero3_init_flag=True)
    accelerator = Accelerator(deepspeed_plugin=plugin)
    assert accelerator.state.deepspeed_plugin is plugin and plugin.is_zero3_init_enabled()
    if accelerator.num_processes > 1:
        assert accelerator.distributed_type == DistributedType.DEEPSPEED
    set_seed(0)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    on_gpu = accelerator.device.type == "cuda"
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16 if on_gpu else torch.float32, device=accelerator.device)
    names = sorted(model.state_dict())
    before = {k: v.detach().float().cpu().clone() for k, v in model.state_dict().items()}
    optimizer = torch.optim.AdamW(model.parameters(), lr=1e-2)
    if accelerator.distributed_type == DistributedType.DEEPSPEED:
        try:
            accelerator.prepare(model, optimizer)
        except ValueError as error:  # `auto` batch sizes need a dataloader to be read from, as under DeepSpeed
            assert "requires you to pass at least one of training or evaluation dataloaders" in str(error)
        else:
            raise AssertionError("`auto` batch sizes were accepted without a dataloader")
    g = torch.Generator().manual_seed(accelerator.process_index)
    data = [torch.randint(0, cfg.vocab_size, (32,), generator=g) for _ in range(6)]
    loader = torch.utils.data.DataLoader(data, batch_size=2)
    model, optimizer, loader = accelerator.prepare(model, optimizer, loader)
    resolved = plugin.deepspeed_config
    if accelerator.distributed_type == DistributedType.DEEPSPEED:
        assert resolved["train_micro_batch_size_per_gpu"] == 2 and resolved["train_batch_size"] == 2 * accelerator.num_processes
    if accelerator.num_processes > 1:
        assert type(model).__name__ == "B200ShardedModule", type(model).__name__  # stage 3 == parameters sharded
    losses = []
    for ids in loader:
        loss = model(input_ids=ids, labels=ids).loss
        accelerator.backward(loss)
        optimizer.step()
        optimizer.zero_grad()
        losses.append(float(loss))
    assert all(torch.isfinite(torch.tensor(losses)))
    whole = accelerator.get_state_dict(model)
    assert sorted(whole) == names, (sorted(whole)[:3], names[:3])
    assert all(tuple(whole[k].shape) == tuple(before[k].shape) for k in names)
    assert any(not torch.allclose(whole[k].float().cpu(), before[k], atol=1e-6) for k in names), "the weights did not move"
    accelerator.print(f"ZeRO-3 configuration: sharded, trained ({losses[0]:.3f} -> {losses[-1]:.3f}), gathered whole again")
    accelerator.print("test_zero3_integration: all passed")
    accelerator.end_training()


if __name__ == "__main__":
    main()
