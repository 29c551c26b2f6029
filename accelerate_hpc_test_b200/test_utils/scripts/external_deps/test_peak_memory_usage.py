"""Memory must stay within a budget.  A model is trained for a few steps while a tracker records the peak device memory (from the
caching allocator) or, on CPU, the peak resident set; the run fails when the peak exceeds ``--peak_memory_upper_bound`` MB.  With a
sharded engine the bound can be set below what replicated training needs — which is what proves that sharding shards.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_peak_memory_usage \
        --peak_memory_upper_bound 4000
"""

import argparse
import gc
import json
import os
import threading

import psutil
import torch
from torch.utils.data import DataLoader

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin, set_seed


class PeakMemory:
    """``with PeakMemory() as m`` → ``m.device_mb`` (allocator high-water mark above the entry level), ``m.host_mb`` (polled RSS)."""

    def __enter__(self):
        gc.collect()
        self.cuda = torch.cuda.is_available()
        if self.cuda:
            torch.cuda.empty_cache()
            torch.cuda.reset_peak_memory_stats()
            self.base = torch.cuda.memory_allocated()
        self.process = psutil.Process()
        self.host_base = self.host_peak = self.process.memory_info().rss
        self.done = threading.Event()
        self.thread = threading.Thread(target=self.poll, daemon=True)
        self.thread.start()
        return self

    def poll(self):
        while not self.done.wait(0.005):
            self.host_peak = max(self.host_peak, self.process.memory_info().rss)

    def __exit__(self, *exc):
        self.done.set()
        self.thread.join()
        self.device_mb = (torch.cuda.max_memory_allocated() - self.base) / 2**20 if self.cuda else 0.0
        self.host_mb = (max(self.host_peak, self.process.memory_info().rss) - self.host_base) / 2**20


def main():
    parser = argparse.ArgumentParser(description="Peak memory of a short training run must stay under a bound.")
    parser.add_argument("--peak_memory_upper_bound", type=float, default=None, help="MB; the run fails above it.")
    parser.add_argument("--use_fsdp", action="store_true")
    parser.add_argument("--hidden_size", type=int, default=256)
    parser.add_argument("--num_layers", type=int, default=4)
    parser.add_argument("--steps", type=int, default=4)
    parser.add_argument("--output_dir", type=str, default=None)
    args = parser.parse_args()

    plugin = None
    if args.use_fsdp:
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    accelerator = Accelerator(fsdp_plugin=plugin)
    set_seed(0)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=args.hidden_size, intermediate_size=4 * args.hidden_size, vocab_size=2048,
                           num_attention_heads=args.hidden_size // 32, num_key_value_heads=args.hidden_size // 64, num_hidden_layers=args.num_layers)
    on_gpu = accelerator.device.type == "cuda"
    with PeakMemory() as tracked:
        model = LlamaForCausalLM(cfg, dtype=torch.bfloat16 if on_gpu else torch.float32, device=accelerator.device)
        n_params = sum(p.numel() for p in model.parameters())
        optimizer = torch.optim.AdamW(model.parameters(), lr=1e-4)
        data = [torch.randint(0, cfg.vocab_size, (128,), generator=torch.Generator().manual_seed(i)) for i in range(4 * args.steps * accelerator.num_processes)]
        loader = DataLoader(data, batch_size=4)
        model, optimizer, loader = accelerator.prepare(model, optimizer, loader)
        for step, ids in enumerate(loader):
            accelerator.backward(model(input_ids=ids, labels=ids).loss)
            optimizer.step()
            optimizer.zero_grad()
            if step + 1 == args.steps:
                break
    peak = tracked.device_mb if on_gpu else tracked.host_mb
    params_mb = n_params * (2 if on_gpu else 4) / 2**20
    accelerator.print(f"peak {'device' if on_gpu else 'host'} memory {peak:.0f} MB on rank 0 ({params_mb:.0f} MB of parameters, "
                      f"{'sharded over ' + str(accelerator.num_processes) if args.use_fsdp else 'replicated'})")
    if args.peak_memory_upper_bound is not None:
        assert peak <= args.peak_memory_upper_bound, f"Peak memory usage exceeded the upper bound of {args.peak_memory_upper_bound} MB: {peak:.0f} MB"
    accelerator.wait_for_everyone()
    if accelerator.is_main_process and args.output_dir is not None:
        os.makedirs(args.output_dir, exist_ok=True)
        with open(os.path.join(args.output_dir, "peak_memory_utilization.json"), "w") as f:
            json.dump({"peak_mb": peak, "parameters_mb": params_mb}, f)
    accelerator.print("test_peak_memory_usage: all passed")
    accelerator.end_training()


if __name__ == "__main__":
    main()
