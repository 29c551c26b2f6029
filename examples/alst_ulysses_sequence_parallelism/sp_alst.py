"""Long-sequence training with Ulysses sequence parallelism (reference examples/alst_ulysses_sequence_parallelism/sp-alst.py,
which needs DeepSpeed >= 0.18.2).  Every rank of the ``sp`` group holds ``seq_len / sp`` tokens of each sequence; inside
attention the QKV projection is fused with the all-to-all that re-shards by heads, flash attention runs over the whole
sequence for ``heads / sp`` heads, and the output projection's input comes back sequence-sharded.  The prepared dataloader
hands out sequence slices with pre-shifted labels and global position ids; the loss is the global token mean.

    accelerate launch --num_processes 8 examples/alst_ulysses_sequence_parallelism/sp_alst.py --sp-size 8 --sequence-length 65536
    accelerate launch --num_processes 2 --cpu examples/alst_ulysses_sequence_parallelism/sp_alst.py --cpu --model tiny --sp-size 2 --sequence-length 128
"""

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig
from accelerate_hpc_test_b200.utils import DeepSpeedSequenceParallelConfig, FullyShardedDataParallelPlugin, set_seed


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--sp-size", type=int, default=2)
    p.add_argument("--dp-shard-size", type=int, default=1)
    p.add_argument("--sequence-length", type=int, default=16384)
    p.add_argument("--num-steps", type=int, default=10)
    p.add_argument("--model", default="llama3-8b", choices=["llama3-8b", "llama3-1b", "tiny"])
    p.add_argument("--cpu", action="store_true")
    args = p.parse_args()
    set_seed(42)
    pc = ParallelismConfig(
        sp_size=args.sp_size, dp_shard_size=args.dp_shard_size, sp_backend="b200",
        sp_handler=DeepSpeedSequenceParallelConfig(sp_seq_length=args.sequence_length, sp_seq_length_is_variable=False,
                                                   sp_attn_implementation="flash_attention_2"),
    )
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=None if args.cpu else "bf16", parallelism_config=pc, fsdp_plugin=plugin)
    cfg = {"llama3-8b": LlamaConfig.llama3_8b, "llama3-1b": LlamaConfig.llama3_1b,
           "tiny": lambda: LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512,
                                            num_attention_heads=4, num_key_value_heads=2)}[args.model]()
    model = LlamaForCausalLM(cfg, dtype=torch.float32 if args.cpu else torch.bfloat16, device=accelerator.device)
    optimizer = torch.optim.AdamW(model.parameters(), lr=1e-5)
    g = torch.Generator().manual_seed(0)
    data = [{"input_ids": torch.randint(0, cfg.vocab_size, (args.sequence_length,), generator=g)} for _ in range(args.num_steps * args.dp_shard_size)]
    collate = lambda ex: {"input_ids": torch.stack([e["input_ids"] for e in ex]), "labels": torch.stack([e["input_ids"] for e in ex])}  # noqa: E731
    dataloader = DataLoader(data, batch_size=1, collate_fn=collate)
    # prepare() wraps the loader in the Ulysses adapter: batches arrive as [1, seq_len / sp] slices with `shift_labels`
    model, optimizer, dataloader = accelerator.prepare(model, optimizer, dataloader)

    t0 = time.perf_counter()
    for step, batch in enumerate(dataloader):
        assert batch["input_ids"].shape[1] == args.sequence_length // args.sp_size and "shift_labels" in batch
        loss = model(**batch).loss  # already the mean over ALL tokens of the sequence (identical on the sp ranks)
        accelerator.backward(loss)
        optimizer.step()
        optimizer.zero_grad()
        if step % 2 == 0:
            accelerator.print(f"step {step}: loss {loss.item():.4f}")
    if accelerator.device.type == "cuda":
        torch.cuda.synchronize()
    n = len(dataloader)
    accelerator.print(f"{n * args.sequence_length * args.dp_shard_size / (time.perf_counter() - t0):.0f} tokens/s "
                      f"(sp={args.sp_size}, sequence length {args.sequence_length})")
    accelerator.end_training()


if __name__ == "__main__":
    main()
