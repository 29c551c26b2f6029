"""Distributed inference by splitting the *inputs*: every process takes its share of the prompts with
``PartialState.split_between_processes`` and the results are gathered (reference examples/inference/distributed/*.py).

    accelerate launch --num_processes 2 --cpu examples/inference/distributed_split.py --cpu
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch

from accelerate_hpc_test_b200 import PartialState
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.utils import gather_object


@torch.no_grad()
def greedy_generate(model, ids, new_tokens):
    for _ in range(new_tokens):
        logits = model(ids, return_logits=True).logits[:, -1]
        ids = torch.cat([ids, logits.argmax(-1, keepdim=True)], dim=1)
    return ids


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--num-prompts", type=int, default=5)
    args = p.parse_args()
    state = PartialState(cpu=args.cpu)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512, num_attention_heads=4, num_key_value_heads=2)
    torch.manual_seed(0)  # same weights everywhere
    model = LlamaForCausalLM(cfg, dtype=torch.float32 if args.cpu else torch.bfloat16, device=state.device).eval()
    g = torch.Generator().manual_seed(1)
    prompts = [torch.randint(0, cfg.vocab_size, (1, 8), generator=g) for _ in range(args.num_prompts)]
    results = []
    # apply_padding=True would repeat the last prompt so that every rank gets the same count (needed before `gather`)
    with state.split_between_processes(prompts) as mine:
        for ids in mine:
            out = greedy_generate(model, ids.to(state.device), new_tokens=4)
            results.append(out[0, -4:].tolist())
    everything = gather_object(results)
    if state.is_main_process:
        assert len(everything) == args.num_prompts
        for i, toks in enumerate(everything):
            print(f"prompt {i}: generated {toks}")


if __name__ == "__main__":
    main()
