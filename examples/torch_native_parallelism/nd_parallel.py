"""N-D parallel Llama training: any mix of dp_replicate x dp_shard x cp x tp from one ``ParallelismConfig`` (reference
examples/torch_native_parallelism/nd_parallel.py).  The native engines do the work: sharded (FSDP2-semantics) engine over
dp_shard(+cp), tensor-parallel linears with fused all-gather->GEMM / GEMM->reduce-scatter, context-parallel attention.

    accelerate launch --num_processes 8 examples/torch_native_parallelism/nd_parallel.py --dp-shard-size 4 --tp-size 2
    accelerate launch --num_processes 4 --cpu examples/torch_native_parallelism/nd_parallel.py --model tiny --tp-size 2 --dp-shard-size 2
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
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin, set_seed


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--dp-replicate-size", type=int, default=1)
    p.add_argument("--dp-shard-size", type=int, default=1)
    p.add_argument("--tp-size", type=int, default=1)
    p.add_argument("--cp-size", type=int, default=1)
    p.add_argument("--sequence-length", type=int, default=1024)
    p.add_argument("--num-steps", type=int, default=20)
    p.add_argument("--batch-size", type=int, default=1, help="sequences per data-parallel rank and step")
    p.add_argument("--model", default="llama3-8b", choices=["llama3-8b", "llama3-1b", "tiny"])
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--save-dir", type=str, default=None)
    return p.parse_args()


def synthetic_tokens(n, seq_len, vocab, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [{"input_ids": torch.randint(0, vocab, (seq_len,), generator=g)} for _ in range(n)]


def collate(examples):
    ids = torch.stack([e["input_ids"] for e in examples])
    return {"input_ids": ids, "labels": ids.clone()}


def main():
    args = parse_args()
    set_seed(42)
    pc = ParallelismConfig(dp_replicate_size=args.dp_replicate_size, dp_shard_size=args.dp_shard_size, tp_size=args.tp_size, cp_size=args.cp_size)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"], state_dict_type="SHARDED_STATE_DICT")
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=None if args.cpu else "bf16", parallelism_config=pc, fsdp_plugin=plugin)
    cfg = {"llama3-8b": LlamaConfig.llama3_8b, "llama3-1b": LlamaConfig.llama3_1b,
           "tiny": lambda: LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512,
                                            num_attention_heads=4, num_key_value_heads=2)}[args.model]()
    dtype = torch.float32 if args.cpu else torch.bfloat16
    model = LlamaForCausalLM(cfg, dtype=dtype, device=accelerator.device)
    optimizer = torch.optim.AdamW(model.parameters(), lr=1e-5)
    data = synthetic_tokens(args.num_steps * args.batch_size * max(1, accelerator.num_processes), args.sequence_length, cfg.vocab_size)
    dataloader = DataLoader(data, batch_size=args.batch_size, collate_fn=collate)
    # tp / cp ranks of one group receive the same batch; dp_replicate x dp_shard ranks different ones
    model, optimizer, dataloader = accelerator.prepare(model, optimizer, dataloader)

    tokens, t0 = 0, time.perf_counter()
    for step, batch in enumerate(dataloader):
        if step >= args.num_steps:
            break
        # with cp_size > 1 the buffers are sharded along the sequence for the duration of the step (no-op otherwise)
        shift = torch.full_like(batch["labels"], -100)
        shift[:, :-1] = batch["labels"][:, 1:]
        buffers = [batch["input_ids"], shift]
        with accelerator.maybe_context_parallel(buffers=buffers, buffer_seq_dims=[1, 1], no_restore_buffers=set(buffers)):
            loss = model(input_ids=batch["input_ids"], shift_labels=shift).loss
            accelerator.backward(loss)
        optimizer.step()
        optimizer.zero_grad()
        tokens += args.batch_size * args.sequence_length
        if step % 5 == 0:
            avg = accelerator.reduce(loss.detach().float(), reduction="mean")
            accelerator.print(f"step {step}: loss {avg.item():.4f}")
    if accelerator.device.type == "cuda":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    dp = pc.dp_replicate_size * pc.dp_shard_size
    accelerator.print(f"{tokens * dp / dt:.0f} tokens/s over {accelerator.num_processes} processes ({pc})")
    if args.save_dir:
        accelerator.save_state(args.save_dir)
    accelerator.end_training()


if __name__ == "__main__":
    main()
