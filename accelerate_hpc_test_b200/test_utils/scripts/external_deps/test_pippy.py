"""Pipeline-parallel inference: ``prepare_pippy`` cuts a model into one stage per process; the last stage (or every stage, with
``gather_output=True``) must hold the logits of the un-split model.  Checked on the native Llama (block-level stages), on a plain
``nn.Sequential`` and with keyword inputs.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.external_deps.test_pippy
"""

import torch

from accelerate_hpc_test_b200 import PartialState
from accelerate_hpc_test_b200.inference import prepare_pippy
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.utils import set_seed


def llama(state, gather_output):
    set_seed(0)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=max(4, state.num_processes))
    on_gpu = state.device.type == "cuda"
    model = LlamaForCausalLM(cfg, dtype=torch.bfloat16 if on_gpu else torch.float32, device=state.device)
    ids = torch.randint(0, cfg.vocab_size, (2 * state.num_processes, 16), generator=torch.Generator().manual_seed(1)).to(state.device)
    with torch.no_grad():
        want = model(input_ids=ids).logits.float()
    piped = prepare_pippy(model, num_chunks=state.num_processes, gather_output=gather_output)
    with torch.no_grad():
        out = piped(input_ids=ids) if state.is_main_process else piped()
    if gather_output or state.is_last_process:
        assert out is not None and out.shape == want.shape
        assert torch.allclose(out.float(), want, atol=3e-2 if on_gpu else 1e-5), (out.float() - want).abs().max()
    else:
        assert out is None
    assert len(model.hf_split_points) == state.num_processes - 1


def sequential(state):
    set_seed(1)
    layers = [torch.nn.Linear(32, 32) if i % 2 == 0 else torch.nn.Tanh() for i in range(4 * state.num_processes)]
    model = torch.nn.Sequential(*layers).to(state.device)
    x = torch.randn(4 * state.num_processes, 32, generator=torch.Generator().manual_seed(2)).to(state.device)
    with torch.no_grad():
        want = model(x)
    piped = prepare_pippy(model, gather_output=True)
    with torch.no_grad():
        out = piped(x) if state.is_main_process else piped()
        again = piped(2 * x) if state.is_main_process else piped()  # stages are reusable
    assert torch.allclose(out, want, atol=1e-5) and again.shape == want.shape and not torch.allclose(again, out)


def main():
    state = PartialState()
    if state.num_processes < 2:
        raise SystemExit("a pipeline needs at least two stages: launch with at least two processes")
    for gather_output in (False, True):
        llama(state, gather_output)
        state.wait_for_everyone()
        state.print(f"Llama split into {state.num_processes} stages reproduces the un-split logits (gather_output={gather_output})")
    sequential(state)
    state.wait_for_everyone()
    state.print("test_pippy: all passed")


if __name__ == "__main__":
    main()
