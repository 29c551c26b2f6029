"""Diagnostic: per-kernel device time of one sharded (FSDP2) training step, from torch.profiler on rank 0.
Launch with torchrun on N GPUs; LAYERS (default 8) decoder layers of the Llama-3-8B shape.  Numbers taken under a profiler
are for finding out WHERE time goes, never bench values."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin


def main():
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(mixed_precision="bf16", fsdp_plugin=plugin)
    cfg = LlamaConfig.llama3_8b(num_hidden_layers=int(os.environ.get("LAYERS", 8)))
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=acc.device)
    opt = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    ids = torch.randint(0, cfg.vocab_size, (2, 2048), device=acc.device)

    def step():
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()

    for _ in range(4):
        step()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.distributed.barrier()
    s.record()
    for _ in range(3):
        step()
    e.record()
    torch.cuda.synchronize()
    if acc.is_main_process:
        print(f"step ms (no profiler): {s.elapsed_time(e) / 3:.2f}")
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    if acc.is_main_process:
        print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=30, max_name_column_width=70))


if __name__ == "__main__":
    main()
