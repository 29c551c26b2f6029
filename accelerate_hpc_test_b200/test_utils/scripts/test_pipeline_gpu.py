"""Pipeline-parallel inference on GPUs: ``prepare_pippy`` over N ranks (one stage per GPU, NCCL p2p between stages) must produce the
logits of the single-GPU forward.  Also times a Llama-3-8B-shaped 8-layer model against its single-GPU forward.

Launch: ``torchrun --nproc-per-node N --master-addr 127.0.0.1 -m accelerate_hpc_test_b200.test_utils.scripts.test_pipeline_gpu``
"""

import json
import os

import torch

from accelerate_hpc_test_b200 import PartialState
from accelerate_hpc_test_b200.inference import prepare_pippy
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM


def main():
    state = PartialState()
    dev = state.device
    world = state.num_processes
    # ---- correctness: tiny model, bf16 kernels on every stage
    cfg = LlamaConfig.tiny(num_hidden_layers=max(4, world))
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=dev)
    ids = torch.randint(0, cfg.vocab_size, (4, 128), generator=torch.Generator().manual_seed(3)).to(dev)
    with torch.no_grad():
        want = model(input_ids=ids).logits.float()
    piped = prepare_pippy(model, num_chunks=2)
    with torch.no_grad():
        out = piped(input_ids=ids) if state.is_main_process else piped()
    if state.is_last_process:
        err = ((out.float() - want).norm() / want.norm()).item()
        assert out.shape == want.shape and err < 2e-2, err
        print(f"pipeline logits match the single-GPU forward on {world} stages (rel err {err:.1e}); stage ranges "
              f"{model.pippy_stage.lo}..{model.pippy_stage.hi}", flush=True)
    state.wait_for_everyone()
    # ---- timing: 8 layers of Llama-3-8B, 8 sequences of 2048 tokens, 4 micro-batches in flight
    big = LlamaConfig.llama3_8b(num_hidden_layers=8)
    torch.manual_seed(1)
    model = LlamaForCausalLM(big, dtype=torch.bfloat16, device=dev)
    ids = torch.randint(0, big.vocab_size, (8, 2048), generator=torch.Generator().manual_seed(4)).to(dev)

    def time_it(fn, iters=5):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        state.wait_for_everyone()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / iters], device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return t.item()

    with torch.no_grad():
        single = time_it(lambda: model(input_ids=ids).logits)
        piped = prepare_pippy(model, num_chunks=4)
        pipe = time_it(lambda: piped(input_ids=ids) if state.is_main_process else piped())
    if state.is_main_process:
        row = {"op": "pipeline inference, 8 x Llama-3-8B layers + head, 8 x 2048 tokens", "stages": world, "micro_batches": 4,
               "single_gpu_forward_ms": round(single, 2), "pipelined_ms": round(pipe, 2), "speedup": round(single / pipe, 3)}
        print(json.dumps(row), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(row, open(f"gpurun_out/pipeline_n{world}.json", "w"))
    state.wait_for_everyone()
    print("PIPELINE CHECKS PASSED") if state.is_main_process else None


if __name__ == "__main__":
    main()
