#!/usr/bin/env python
"""Headline benchmark: Llama-3-8B training tokens/s on N B200s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference]

* ``--impl ours``      : this framework — native sm_100a kernels, FusedAdamW, own DDP/FSDP engines,
                         driven through the public ``Accelerator`` API.
* ``--impl reference`` : the UNMODIFIED reference installed in ``baseline/_ref`` (``pip install
                         --no-deps --target baseline/_ref``), transformers' LlamaForCausalLM + torch AdamW
                         through the reference ``Accelerator`` (FSDP2 plugin for N > 1), same config.

Both arms: synthetic token batches of the named shape, random-init weights, bf16 compute, AdamW,
seq 2048, fixed per-GPU micro-batch (weak scaling).  Timing: W warm-up steps, then exactly K steps
bracketed by barrier + ``torch.cuda.synchronize()``, CUDA events, max over ranks.  The working set of
one step (>100 GB of weights, optimizer state and activations) is far larger than the 126 MB L2, so
no explicit L2 flush is needed between steps.  A second timed loop (``e2e``) includes, every step,
the host->device copy of that step's batch from pinned memory and a device->host read of the loss.
Rank 0 prints one JSON line.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

# The only throughput the reference publishes is Llama-3.1-8B FSDP2 bf16 at seq 8192 with regional
# torch.compile on 8xH100: other hardware, other shape.  It is reported as `vs_published_h100` for context only;
# `vs_baseline` stays null because BASELINE.json.published is empty (the binding baseline is `--impl reference`).
PUBLISHED_H100_TOKENS_PER_S_PER_DEVICE = 8000.0

MODELS = {
    "llama3-8b": dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                      num_attention_heads=32, num_key_value_heads=8, head_dim=128, rope_theta=500000.0,
                      rms_norm_eps=1e-5, max_position_embeddings=8192),
    "llama3-1b": dict(vocab_size=128256, hidden_size=2048, intermediate_size=8192, num_hidden_layers=16,
                      num_attention_heads=16, num_key_value_heads=8, head_dim=128, rope_theta=500000.0,
                      rms_norm_eps=1e-5, max_position_embeddings=8192),
    "tiny": dict(vocab_size=2048, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                 num_attention_heads=4, num_key_value_heads=2, head_dim=128, rope_theta=500000.0,
                 rms_norm_eps=1e-5, max_position_embeddings=8192),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--model", default=os.environ.get("BENCH_MODEL", "llama3-8b"), choices=list(MODELS))
    ap.add_argument("--seq-len", type=int, default=2048)
    ap.add_argument("--micro-batch", type=int, default=int(os.environ.get("BENCH_MICRO_BATCH", 2)))
    ap.add_argument("--sweep", default="", help="diagnostic: engine_attribute=v1,v2,... re-times the step with each value (e.g. rs_ctas=32,64)")
    ap.add_argument("--parallelism", default="auto",
                    help="auto | ddp | fsdp | tpN | spN | cpN  (tp/sp/cp: N ranks per model-parallel group, the rest is dp_shard)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp8"], help="fp8: native e4m3 linear layers (ours arm)")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# --------------------------------------------------------------------------- clocks sampling
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of the local GPU while the timed region runs."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- shared harness
def dist_setup(n_gpus: int):
    import torch

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    return rank, world, local


def barrier_sync(world):
    import torch

    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def timed_loop(step_fn, steps: int, world: int) -> float:
    """Device time (ms) of `steps` calls of step_fn, max over ranks."""
    import torch

    barrier_sync(world)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(steps):
        step_fn(i)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e)
    barrier_sync(world)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def make_batches(n: int, micro_batch: int, seq_len: int, vocab: int, seed: int):
    """n pinned host batches of random token ids (synthetic data of the named shape)."""
    import torch

    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        ids = torch.randint(0, vocab, (micro_batch, seq_len), generator=g, dtype=torch.int64).pin_memory()
        out.append(ids)
    return out


def run(args, build_fn):
    import torch

    rank, world, local = dist_setup(args.gpus)
    cfg = MODELS[args.model]
    step_device, step_e2e, info = build_fn(args, cfg, rank, world, local)
    n_host = 4
    # ranks of one tp / sp / cp group work on the same sequences; the model-parallel axis is the fastest mesh axis
    mp = int(info.get("model_parallel_size", 1))
    dp_world = world // mp
    host_batches = make_batches(n_host, args.micro_batch, args.seq_len, cfg["vocab_size"], seed=1234 + rank // mp)
    if info.get("shard_batch") is not None:
        host_batches = [info["shard_batch"](b).contiguous().pin_memory() for b in host_batches]
    dev_batches = [b.cuda(non_blocking=True) for b in host_batches]
    torch.cuda.synchronize()

    first_loss = None
    for i in range(args.warmup):
        loss = step_device(dev_batches[i % n_host])
        if i == 0:
            first_loss = float(loss.item())  # random-init model on random tokens: ~ln(vocab)
    launches0 = info["launch_count"]() if info.get("launch_count") else 0
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed_loop(lambda i: step_device(dev_batches[i % n_host]), args.steps, world)
    clocks = sampler.stop() if rank == 0 else {}
    launches = (info["launch_count"]() - launches0) if info.get("launch_count") else None

    tokens_per_step = args.micro_batch * args.seq_len * dp_world
    value = tokens_per_step * args.steps / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        step_e2e(host_batches[0])  # warm the H2D path
        losses = []
        ms_e2e = timed_loop(lambda i: losses.append(step_e2e(host_batches[i % n_host])), args.steps, world)
        e2e = {"value": tokens_per_step * args.steps / (ms_e2e / 1e3), "unit": "tokens/s",
               "h2d_bytes_per_step": int(host_batches[0].numel() * host_batches[0].element_size()),
               "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps, "last_loss": losses[-1] if losses else None}

    exposed = None
    if info.get("exposed_comm") is not None and world > 1:
        exposed = info["exposed_comm"](step_device, dev_batches)

    sweep = None
    if args.sweep and info.get("engine") is not None:
        # diagnostic (not part of the headline): re-time the same step with another CTA budget for the collective that runs
        # beside backward, on this very box and model instance — "attr=v1,v2,..."
        attr, _, values = args.sweep.partition("=")
        engine, sweep = info["engine"], {"attribute": attr, "ms_per_step": {}}
        original = getattr(engine, attr)
        for v in [int(x) for x in values.split(",") if x]:
            setattr(engine, attr, v)
            for i in range(2):
                step_device(dev_batches[i % n_host])
            sweep["ms_per_step"][str(v)] = round(timed_loop(lambda i: step_device(dev_batches[i % n_host]), args.steps, world) / args.steps, 2)
        setattr(engine, attr, original)

    if rank == 0:
        out = {
            "metric": "Llama-3-8B training tokens/s (whole job)" if args.model == "llama3-8b" else f"{args.model} training tokens/s",
            "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None,
            "vs_published_h100": value / (PUBLISHED_H100_TOKENS_PER_S_PER_DEVICE * world) if args.model == "llama3-8b" else None,
            "dtype": args.precision if args.impl == "ours" else "bf16", "data": "synthetic random token ids; random-init weights", "impl": args.impl,
            "config": {"model": args.model, "global_batch": args.micro_batch * dp_world, "micro_batch_per_gpu": args.micro_batch,
                       "seq_len": args.seq_len, "parallelism": info["parallelism"], "optimizer": "AdamW",
                       "l2_policy": "working set per step (>100 GB) exceeds the 126 MB L2; no explicit flush",
                       "baseline_note": "BASELINE.json publishes no number (vs_baseline null); vs_published_h100 = value / (8000 tokens/s/device x N), "
                                        "the reference's 8xH100 seq-8192 torch.compile figure, a different shape and machine"},
            "first_loss": first_loss,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "exposed_comm_ms_per_step": exposed,
            "tflops_per_gpu": info.get("flops_per_token", 0) * value / world / 1e12 if info.get("flops_per_token") else None,
            "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30,
        }
        if sweep is not None:
            out["cta_sweep"] = sweep
        print(json.dumps(out), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


# --------------------------------------------------------------------------- our arm
def build_ours(args, cfg, rank, world, local):
    import torch

    sys.path.insert(0, ROOT)
    import accelerate_hpc_test_b200 as acc_pkg
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.ops import _lib
    from accelerate_hpc_test_b200.optim import FusedAdamW
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    par = args.parallelism
    if par == "auto":
        par = "fsdp" if world > 1 else "single"
    plugin = None
    pc = None
    mp_kind, mp = None, 1
    if par[:2] in ("tp", "sp", "cp") and par[2:].isdigit():
        from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig

        mp_kind, mp = par[:2], int(par[2:])
        if world % mp:
            raise SystemExit(f"--parallelism {par} needs a world size divisible by {mp}")
        kw = {f"{mp_kind}_size": mp, "dp_shard_size": world // mp}
        if mp_kind == "sp":
            kw["sp_backend"] = "b200"
        pc = ParallelismConfig(**kw)
    if par == "fsdp" or pc is not None:
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    accelerator = Accelerator(mixed_precision=args.precision, fsdp_plugin=plugin, parallelism_config=pc)
    lcfg = LlamaConfig(**cfg)
    torch.manual_seed(0)
    model = LlamaForCausalLM(lcfg, dtype=torch.bfloat16, device=accelerator.device)
    optimizer = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, optimizer = accelerator.prepare(model, optimizer)
    _lib.get_lib()

    def step_device(ids):
        optimizer.zero_grad()
        out = model(input_ids=ids, labels=ids)
        accelerator.backward(out.loss)
        optimizer.step()
        return out.loss

    def step_e2e(host_ids):
        ids = host_ids.to(accelerator.device, non_blocking=True)
        loss = step_device(ids)
        return float(loss.item())

    names = {"single": "single-gpu", "fsdp": f"fsdp2 full-shard dp_shard={world}", "ddp": f"ddp dp={world}"}
    shard_batch = None
    if mp_kind == "sp":
        c = args.seq_len // mp
        r = accelerator.torch_device_mesh.get_local_rank("sp")
        shard_batch = lambda b: b[:, r * c : (r + 1) * c]  # noqa: E731 - contiguous sequence slice of this sp rank
    elif mp_kind == "cp":
        from accelerate_hpc_test_b200.parallel.context_parallel import shard_sequence

        r = accelerator.torch_device_mesh.get_local_rank("cp")
        shard_batch = lambda b: shard_sequence(b, 1, r, mp)  # noqa: E731 - head + tail chunk of this cp rank
    def exposed_comm(step_fn, batches, steps=3):
        """ms/step the compute stream waited for un-hidden all-gathers / reduce-scatters (max over ranks); outside the timed region"""
        trace = getattr(model, "trace", None)
        if trace is None:
            return None
        trace.enabled = True
        trace.exposed_ms()  # reset
        for i in range(steps):
            step_fn(batches[i % len(batches)])
        per = trace.exposed_ms(by_tag=True)
        trace.enabled = False
        t = torch.tensor([sum(per.values()) / steps] + [per.get(k, 0.0) / steps for k in ("all_gather", "grad_buffer_reuse", "reduce_scatter_tail")],
                         device=accelerator.device)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        v = [round(float(x), 3) for x in t.tolist()]
        return {"total": v[0], "all_gather": v[1], "grad_buffer_reuse": v[2], "reduce_scatter_tail": v[3]}

    info = {"parallelism": names.get(par, f"{mp_kind}={mp} x fsdp2 dp_shard={world // mp}"), "exposed_comm": exposed_comm,
            "launch_count": _lib.launch_count, "flops_per_token": lcfg.flops_per_token(args.seq_len),
            "model_parallel_size": mp, "shard_batch": shard_batch, "engine": model}
    return step_device, step_e2e, info


# --------------------------------------------------------------------------- reference arm
def build_reference(args, cfg, rank, world, local):
    import torch

    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    sys.path.insert(0, ref_dir)
    import accelerate  # the unmodified reference
    from accelerate import Accelerator
    from accelerate.utils import FullyShardedDataParallelPlugin
    from transformers import LlamaConfig as HFLlamaConfig
    from transformers import LlamaForCausalLM as HFLlama

    assert os.path.realpath(accelerate.__file__).startswith(os.path.realpath(ref_dir)), accelerate.__file__
    plugin = None
    if world > 1:
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    accelerator = Accelerator(mixed_precision="bf16", fsdp_plugin=plugin)
    hcfg = HFLlamaConfig(vocab_size=cfg["vocab_size"], hidden_size=cfg["hidden_size"], intermediate_size=cfg["intermediate_size"],
                         num_hidden_layers=cfg["num_hidden_layers"], num_attention_heads=cfg["num_attention_heads"],
                         num_key_value_heads=cfg["num_key_value_heads"], head_dim=cfg["head_dim"], rope_theta=cfg["rope_theta"],
                         rms_norm_eps=cfg["rms_norm_eps"], max_position_embeddings=cfg["max_position_embeddings"],
                         tie_word_embeddings=False, use_cache=False, attention_bias=False)
    torch.manual_seed(0)
    with torch.device(accelerator.device):
        model = HFLlama(hcfg)  # fp32 weights, as in the reference's benchmarks/fsdp2/utils.py:get_model
    optimizer = torch.optim.AdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
    model, optimizer = accelerator.prepare(model, optimizer)

    def step_device(ids):
        optimizer.zero_grad()
        out = model(input_ids=ids, labels=ids, use_cache=False)
        accelerator.backward(out.loss)
        optimizer.step()
        return out.loss

    def step_e2e(host_ids):
        ids = host_ids.to(accelerator.device, non_blocking=True)
        loss = step_device(ids)
        return float(loss.item())

    info = {"parallelism": f"fsdp2 full-shard dp_shard={world}" if world > 1 else "single-gpu", "launch_count": None}
    return step_device, step_e2e, info


def ensure_reference() -> str:
    """Make ``baseline/_ref/accelerate`` exist (rank 0 installs it from the reference tree when that is visible; the
    other ranks wait for it).  Returns '' on success, else a one-line reason."""
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import install_reference as ir

    if ir.installed():
        return ""
    if int(os.environ.get("LOCAL_RANK", 0)) == 0:
        how = ir.install()
        return "" if ir.installed() else how
    t0 = time.time()
    while time.time() - t0 < 180:
        if ir.installed():
            return ""
        time.sleep(1.0)
    return "unavailable: rank 0 did not finish installing baseline/_ref"


def main():
    args = parse_args()
    is_rank0 = int(os.environ.get("RANK", 0)) == 0
    if args.impl == "reference":
        why = ensure_reference()
        if why:
            if is_rank0:
                print(json.dumps({"impl": "reference", "unavailable": why.replace("unavailable: ", "")}), flush=True)
            return
        try:
            run(args, build_reference)
        except Exception as exc:  # noqa: BLE001 - report instead of crashing the driver
            if is_rank0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(exc).__name__}: {str(exc)[:300]}"}), flush=True)
        return
    run(args, build_ours)


if __name__ == "__main__":
    main()
