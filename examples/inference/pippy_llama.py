"""Pipeline-parallel inference with ``prepare_pippy``: the model is cut into one stage per GPU, rank 0 feeds the inputs,
the last rank returns the output (reference examples/inference/pippy/llama.py).

    accelerate launch --num_processes 2 examples/inference/pippy_llama.py
"""

import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch

from accelerate_hpc_test_b200 import PartialState, prepare_pippy
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM


def main():
    state = PartialState()
    cfg = LlamaConfig.llama3_1b()
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device="cpu").eval()
    ids = torch.randint(0, cfg.vocab_size, (2, 128))
    model = prepare_pippy(model, split_points="auto", example_args=(ids,), gather_output=True)
    with torch.no_grad():
        out = model(ids.to(state.device))
    if state.is_main_process:
        print("logits on rank 0 after gather:", None if out is None else tuple(out[0].shape if isinstance(out, (tuple, list)) else out.shape))
    state.wait_for_everyone()


if __name__ == "__main__":
    main()
