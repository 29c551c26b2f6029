"""Diagnostic: which kernels one training step launches (count and device time per kernel name, rank 0) and how the host time
compares with the device time.  torch.profiler (CUPTI activity records), two profiled steps after warm-up; the timing lines come
from un-profiled steps.

    PAR=tp2 LAYERS=32 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/kernel_census.py
"""
import collections
import json
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin


def main():
    world = int(os.environ.get("WORLD_SIZE", 1))
    par = os.environ.get("PAR", "fsdp" if world > 1 else "single")
    plugin, pc, mp = None, None, 1
    if par[:2] in ("tp", "sp", "cp"):
        from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig

        mp = int(par[2:])
        kw = {f"{par[:2]}_size": mp, "dp_shard_size": world // mp}
        if par[:2] == "sp":
            kw["sp_backend"] = "b200"
        pc = ParallelismConfig(**kw)
    if par == "fsdp" or pc is not None:
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(mixed_precision="bf16", fsdp_plugin=plugin, parallelism_config=pc)
    cfg = LlamaConfig.llama3_8b(num_hidden_layers=int(os.environ.get("LAYERS", 32)))
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=acc.device)
    opt = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    ids = torch.randint(0, cfg.vocab_size, (int(os.environ.get("MICRO_BATCH", 2)), 2048), device=acc.device)

    def step():
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()

    for _ in range(4):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        step()
    host = (time.perf_counter() - t0) / 3 * 1e3
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 3 * 1e3
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            step()
        torch.cuda.synchronize()
    per = collections.defaultdict(lambda: [0, 0.0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            per[ev.name][0] += 1
            per[ev.name][1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
    if acc.is_main_process:
        rows = sorted(per.items(), key=lambda kv: -kv[1][1])
        total_n = sum(v[0] for v in per.values()) / 2
        total_us = sum(v[1] for v in per.values()) / 2
        print(f"parallelism {par} world {world}: host enqueue {host:.1f} ms/step, wall {wall:.1f} ms/step, "
              f"{total_n:.0f} device activities/step, {total_us / 1e3:.1f} ms summed device time/step")
        for name, (n, us) in rows[:45]:
            print(f"{n / 2:8.0f} x {us / n:9.1f} us = {us / 2e3:8.2f} ms   {name[:110]}")
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump({"parallelism": par, "world": world, "host_ms": host, "wall_ms": wall,
                   "kernels": [{"name": k, "per_step": v[0] / 2, "ms_per_step": v[1] / 2e3} for k, v in rows]},
                  open(f"gpurun_out/kernel_census_{par}_n{world}.json", "w"), indent=1)


if __name__ == "__main__":
    main()
