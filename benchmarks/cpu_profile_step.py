"""Diagnostic: where does the HOST time of one sharded training step go (cProfile on rank 0, 3 steps)."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.getcwd())
import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin


def main():
    plugin = None
    if int(os.environ.get("WORLD_SIZE", 1)) > 1:
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    pc = None
    tp = int(os.environ.get("TP", 1))
    if tp > 1:  # TP=2 torchrun ... : tp x dp_shard mesh, to see what tensor parallelism adds on the host side
        from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig

        pc = ParallelismConfig(tp_size=tp, dp_shard_size=int(os.environ["WORLD_SIZE"]) // tp)
    acc = Accelerator(mixed_precision="bf16", fsdp_plugin=plugin, parallelism_config=pc)
    cfg = LlamaConfig.llama3_8b(num_hidden_layers=int(os.environ.get("LAYERS", 32)))
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=acc.device)
    opt = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    ids = torch.randint(0, cfg.vocab_size, (2 * tp, 2048), device=acc.device)  # same tokens per GPU as the dp runs

    def step():
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()

    for _ in range(4):
        step()
    torch.cuda.synchronize()
    import time

    t0 = time.perf_counter()
    for _ in range(3):
        step()
    host = (time.perf_counter() - t0) / 3 * 1e3
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 3 * 1e3
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        step()
    pr.disable()
    torch.cuda.synchronize()
    if acc.is_main_process:
        print(f"host ms/step to enqueue: {host:.1f}   wall ms/step incl. drain: {wall:.1f}")
        pstats.Stats(pr).sort_stats("tottime").print_stats(45)


if __name__ == "__main__":
    main()
