"""Diagnostic: device time of forward / backward / optimizer.step of the Llama-3-8B FSDP2 step (CUDA events, max over ranks)."""
import os
import sys

sys.path.insert(0, os.getcwd())
import torch
import torch.distributed as dist

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin


def main():
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(mixed_precision="bf16", fsdp_plugin=plugin)
    cfg = LlamaConfig.llama3_8b(num_hidden_layers=int(os.environ.get("LAYERS", 32)))
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=acc.device)
    opt = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    ids = torch.randint(0, cfg.vocab_size, (2, 2048), device=acc.device)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    rows = []
    for it in range(8):
        a, b, c, d = ev(), ev(), ev(), ev()
        opt.zero_grad()
        a.record()
        out = model(input_ids=ids, labels=ids)
        b.record()
        acc.backward(out.loss)
        c.record()
        opt.step()
        d.record()
        torch.cuda.synchronize()
        if it >= 3:
            rows.append((a.elapsed_time(b), b.elapsed_time(c), c.elapsed_time(d)))
    t = torch.tensor(rows, device="cuda").mean(0)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if acc.is_main_process:
        print(f"forward {t[0]:.1f} ms  backward {t[1]:.1f} ms  optimizer.step {t[2]:.1f} ms  total {t.sum():.1f} ms")


if __name__ == "__main__":
    main()
