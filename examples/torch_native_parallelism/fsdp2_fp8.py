"""FSDP2 + fp8 (MXFP8) training of a Llama on the native kernels (reference examples/torch_native_parallelism/fsdp2_fp8.py,
which needs torchao): ``mixed_precision="fp8"`` routes the linear layers through the block-scaled tcgen05 GEMM (forward and
dgrad in e4m3 with one UE8M0 scale per 32 elements, wgrad in bf16); ``--precision bf16`` is the baseline to compare against.

    accelerate launch --num_processes 8 examples/torch_native_parallelism/fsdp2_fp8.py --precision fp8
"""

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin, set_seed


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--precision", choices=["fp8", "bf16"], default="fp8")
    p.add_argument("--sequence-length", type=int, default=2048)
    p.add_argument("--batch-size", type=int, default=2)
    p.add_argument("--num-steps", type=int, default=20)
    p.add_argument("--model", default="llama3-8b", choices=["llama3-8b", "llama3-1b"])
    args = p.parse_args()
    set_seed(42)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    accelerator = Accelerator(mixed_precision=args.precision, fsdp_plugin=plugin)
    cfg = LlamaConfig.llama3_8b() if args.model == "llama3-8b" else LlamaConfig.llama3_1b()
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=accelerator.device)
    optimizer = FusedAdamW(model.parameters(), lr=1e-5)
    model, optimizer = accelerator.prepare(model, optimizer)
    accelerator.print(f"fp8 back-end: {accelerator.fp8_backend}, recipe: {getattr(accelerator.unwrap_model(model), '_b200_fp8_recipe', None)}")
    g = torch.Generator(device=accelerator.device).manual_seed(accelerator.process_index)
    t0 = None
    for step in range(args.num_steps + 3):
        if step == 3:  # warm-up done
            torch.cuda.synchronize()
            t0 = time.perf_counter()
        ids = torch.randint(0, cfg.vocab_size, (args.batch_size, args.sequence_length), generator=g, device=accelerator.device)
        loss = model(input_ids=ids, labels=ids).loss
        accelerator.backward(loss)
        optimizer.step()
        optimizer.zero_grad()
        if step % 5 == 0:
            accelerator.print(f"step {step}: loss {loss.item():.4f}")
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tps = args.num_steps * args.batch_size * args.sequence_length * accelerator.num_processes / dt
    accelerator.print(f"{args.precision}: {tps:.0f} tokens/s, peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB per GPU")
    accelerator.end_training()


if __name__ == "__main__":
    main()
