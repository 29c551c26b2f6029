"""Find which Python call sites launch the non-native (ATen) kernels of a training step: torch.profiler with stacks
on a 4-layer Llama-3-8B-shaped model.  Diagnostic only (numbers under a profiler are never bench values)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW


def main():
    acc = Accelerator(mixed_precision="bf16")
    cfg = LlamaConfig.llama3_8b(num_hidden_layers=int(os.environ.get("LAYERS", 4)))
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=acc.device)
    opt = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    ids = torch.randint(0, cfg.vocab_size, (2, 2048), device=acc.device)

    def step():
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
        step()
        torch.cuda.synchronize()
    for e in prof.key_averages(group_by_stack_n=10):
        if any(k in e.key for k in ("copy", "fill", "zero", "mul", "add", "contiguous", "clone", "aten::to")):
            t = getattr(e, "device_time_total", None) or getattr(e, "cuda_time_total", 0)
            if t > 50:
                print(f"### {e.key}  device_us={t:.0f} calls={e.count}")
                for fr in e.stack[:10]:
                    print("      ", fr)
    print(prof.key_averages(group_by_stack_n=8).table(sort_by="self_cuda_time_total", row_limit=25, max_name_column_width=60,
                                                       max_src_column_width=110))


if __name__ == "__main__":
    main()
