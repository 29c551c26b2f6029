"""Ulysses sequence parallelism (what DeepSpeed's ALST provides in the reference) must be a pure re-arrangement of work: training
with every sequence cut over ``sp`` ranks gives the losses of training on one process with whole sequences.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_ds_alst_ulysses_sp
"""

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig
from accelerate_hpc_test_b200.utils import DeepSpeedSequenceParallelConfig, FullyShardedDataParallelPlugin, set_seed

SEQ, STEPS = 64, 4


def sequences(vocab):
    g = torch.Generator().manual_seed(9)
    return [{"input_ids": torch.randint(0, vocab, (SEQ,), generator=g)} for _ in range(STEPS)]


def collate(rows):
    ids = torch.stack([r["input_ids"] for r in rows])
    return {"input_ids": ids, "labels": ids.clone()}


def main():
    from accelerate_hpc_test_b200 import PartialState

    world = PartialState().num_processes
    if world < 2:
        raise SystemExit("sequence parallelism needs a group: launch with at least two processes")
    pc = ParallelismConfig(sp_size=world, sp_backend="b200",
                           sp_handler=DeepSpeedSequenceParallelConfig(sp_seq_length=SEQ, sp_seq_length_is_variable=False))
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    accelerator = Accelerator(parallelism_config=pc, fsdp_plugin=plugin)
    on_gpu = accelerator.device.type == "cuda"
    dtype = torch.bfloat16 if on_gpu else torch.float32
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=256, num_attention_heads=4,
                           num_key_value_heads=2, num_hidden_layers=2)
    # the one-process run of the same thing, on every rank
    set_seed(3)
    reference = LlamaForCausalLM(cfg, dtype=dtype, device=accelerator.device)
    reference_optimizer = torch.optim.SGD(reference.parameters(), lr=0.05)
    expected = []
    for batch in DataLoader(sequences(cfg.vocab_size), batch_size=1, collate_fn=collate):
        batch = {k: v.to(accelerator.device) for k, v in batch.items()}
        loss = reference(**batch).loss
        loss.backward()
        reference_optimizer.step()
        reference_optimizer.zero_grad()
        expected.append(float(loss))

    set_seed(3)
    model = LlamaForCausalLM(cfg, dtype=dtype, device=accelerator.device)
    optimizer = torch.optim.SGD(model.parameters(), lr=0.05)
    model, optimizer, loader = accelerator.prepare(model, optimizer, DataLoader(sequences(cfg.vocab_size), batch_size=1, collate_fn=collate))
    got = []
    for batch in loader:
        assert batch["input_ids"].shape[1] == SEQ // world and "shift_labels" in batch and "position_ids" in batch
        loss = model(**batch).loss
        accelerator.backward(loss)
        optimizer.step()
        optimizer.zero_grad()
        got.append(float(loss))
    tolerance = 3e-2 if on_gpu else 2e-4
    assert len(got) == STEPS and all(abs(a - b) < tolerance * max(1.0, abs(b)) for a, b in zip(got, expected)), (got, expected)
    accelerator.print(f"sp={world}: losses {[round(x, 4) for x in got]} equal the one-process losses {[round(x, 4) for x in expected]}")
    accelerator.print("test_ds_alst_ulysses_sp: all passed")
    accelerator.end_training()


if __name__ == "__main__":
    main()
