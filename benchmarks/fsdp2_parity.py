"""Raw ``torch.distributed.fsdp.fully_shard`` loop vs the same loop through ``Accelerator``: the loss curves must agree, and each
arm's step time and peak memory are printed.  The counterpart of the reference's ``benchmarks/fsdp2/main.py`` (which asserts that
Accelerate adds nothing to a hand-written FSDP2 loop); here the second arm runs the *native* sharded engine on GPUs, so the same
script also answers "what does the native engine buy over torch's FSDP2 on this model".

    # CPU, two gloo processes (what tests/test_benchmarks.py runs)
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/fsdp2_parity.py --cpu --model tiny --steps 6
    # one B200 node
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/fsdp2_parity.py --model llama3-1b --steps 20

Every arm starts from the same seed, sees the same per-rank batches and uses AdamW with the same hyper-parameters; the optimizer of the
Accelerate arm is created *before* ``prepare`` (the order FSDP2 makes possible and the reference's benchmark stresses).  On GPUs both
arms keep fp32 sharded parameters and optimizer state and compute in bf16 (torch: fp32 model + ``MixedPrecisionPolicy(bf16)``; native
engine: bf16 model + fp32 master shards).
"""

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist


def build(args, device, dtype):
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM

    cfg = {"tiny": lambda: LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512, num_attention_heads=4,
                                            num_key_value_heads=2, num_hidden_layers=args.layers or 2),
           "llama3-1b": LlamaConfig.llama3_1b, "llama3-8b": LlamaConfig.llama3_8b}[args.model]()
    if args.layers and args.model != "tiny":
        cfg.num_hidden_layers = args.layers
    torch.manual_seed(args.seed)
    return LlamaForCausalLM(cfg, dtype=dtype, device=device), cfg


def batches(args, cfg, rank, device):
    """Two distinct batches per rank, repeated: random tokens can only be memorised, and two batches are memorised within a few steps."""
    g = torch.Generator().manual_seed(1000 + rank)
    distinct = [torch.randint(0, cfg.vocab_size, (args.batch, args.seq_len), generator=g).to(device) for _ in range(2)]
    return [distinct[i % 2] for i in range(args.steps)]


class Meter:
    """Step time (device events on GPU, perf_counter on CPU; first ``skip`` steps dropped) and peak memory of one arm."""

    def __init__(self, device, skip):
        self.cuda, self.skip, self.times = device.type == "cuda", skip, []
        if self.cuda:
            torch.cuda.synchronize()
            torch.cuda.reset_peak_memory_stats()

    def __enter__(self):
        if self.cuda:
            self.start, self.end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.start.record()
        else:
            self.t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self.end.record()
            self.end.synchronize()
            self.times.append(self.start.elapsed_time(self.end))
        else:
            self.times.append((time.perf_counter() - self.t0) * 1e3)

    def summary(self):
        kept = self.times[self.skip:] or self.times
        out = {"ms_per_step": sum(kept) / len(kept)}
        if self.cuda:
            out["peak_gb"] = torch.cuda.max_memory_allocated() / 2**30
        return out


def arm_torch(args, device, dtype):
    """The hand-written loop: fully_shard per decoder layer, then the root; optimizer after sharding."""
    from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard

    model, cfg = build(args, device, dtype)
    # the usual FSDP2 mixed-precision set-up: sharded parameters, optimizer state and update in fp32, compute and reduction in bf16.
    # It is also what the native engine does with a bf16 model (fp32 master shards + moments), so both arms round at the same places;
    # starting from the bf16-initialised weights keeps the two arms' initial values identical.
    model = model.float()
    policy = MixedPrecisionPolicy(param_dtype=dtype, reduce_dtype=dtype) if dtype != torch.float32 else MixedPrecisionPolicy()
    for layer in model.model.layers:
        fully_shard(layer, mp_policy=policy)
    fully_shard(model, mp_policy=policy)
    optimizer = torch.optim.AdamW(model.parameters(), lr=args.lr, weight_decay=0.01)
    losses, meter = [], Meter(device, args.skip)
    for ids in batches(args, cfg, dist.get_rank(), device):
        with meter:
            loss = model(input_ids=ids, labels=ids).loss
            loss.backward()
            optimizer.step()
            optimizer.zero_grad()
        losses.append(loss.detach().item())
    return losses, meter.summary()


def arm_accelerate(args, device, dtype):
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.state import AcceleratorState
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    AcceleratorState._reset_state(reset_partial_state=False)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"], reshard_after_forward=True)
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=None if dtype == torch.float32 else "bf16", fsdp_plugin=plugin)
    model, cfg = build(args, device, dtype)
    optimizer = torch.optim.AdamW(model.parameters(), lr=args.lr, weight_decay=0.01)      # before prepare, on purpose
    model, optimizer = accelerator.prepare(model, optimizer)
    losses, meter = [], Meter(device, args.skip)
    for ids in batches(args, cfg, dist.get_rank(), device):
        with meter:
            loss = model(input_ids=ids, labels=ids).loss
            accelerator.backward(loss)
            optimizer.step()
            optimizer.zero_grad()
        losses.append(loss.detach().item())
    return losses, {**meter.summary(), "engine": type(model).__name__}


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--model", default="tiny", choices=["tiny", "llama3-1b", "llama3-8b"])
    p.add_argument("--layers", type=int, default=0, help="cut the model to this many decoder layers (0: keep)")
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--skip", type=int, default=2, help="steps left out of the timing")
    p.add_argument("--batch", type=int, default=2)
    p.add_argument("--seq-len", type=int, default=64)
    p.add_argument("--lr", type=float, default=1e-3)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--rtol", type=float, default=None)
    p.add_argument("--out", default=None)
    args = p.parse_args(argv)

    from accelerate_hpc_test_b200 import PartialState

    state = PartialState(cpu=args.cpu)
    device = state.device
    dtype = torch.float32 if device.type == "cpu" else torch.bfloat16
    ours, ours_stats = arm_accelerate(args, device, dtype)
    theirs, theirs_stats = arm_torch(args, device, dtype)
    # fp32 on CPU: the two arms run the same math in a different order of reductions; bf16 on GPU: different kernels
    rtol = args.rtol if args.rtol is not None else (2e-4 if dtype == torch.float32 else 3e-2)
    worst = max(abs(a - b) / max(abs(b), 1e-6) for a, b in zip(ours, theirs))
    record = {"model": args.model, "world": state.num_processes, "device": device.type, "dtype": str(dtype).split(".")[-1],
              "steps": args.steps, "max_rel_loss_diff": worst, "rtol": rtol, "accelerate": ours_stats, "torch_fsdp2": theirs_stats,
              "first_loss": theirs[0], "last_loss": theirs[-1]}
    if state.is_main_process:
        print(json.dumps(record), flush=True)
        if args.out:
            with open(args.out, "w") as f:
                json.dump({**record, "losses_accelerate": ours, "losses_torch_fsdp2": theirs}, f, indent=1)
    assert worst <= rtol, f"loss curves differ: {worst:.3e} > {rtol:.1e}\naccelerate {ours}\ntorch      {theirs}"
    assert theirs[-1] < theirs[0], "the model did not learn"
    state.wait_for_everyone()
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
