"""Pipeline-parallel inference (`prepare_pippy`): native stages with overlapped p2p, checked on 2 and 3 gloo processes
against the single-process forward of the same model."""

import torch
import torch.nn as nn

from accelerate_hpc_test_b200.inference import Block, _balanced_cut, _cut, _split_inputs, find_pippy_batch_size
from accelerate_hpc_test_b200.launchers import debug_launcher


def test_balanced_cut_covers_all_blocks_and_balances_bytes():
    mods = [nn.Linear(8, 8) for _ in range(6)]
    blocks = [Block(f"m{i}", None, m) for i, m in enumerate(mods)]
    assert _balanced_cut(blocks, 2) == [(0, 3), (3, 6)]
    assert _balanced_cut(blocks, 3) == [(0, 2), (2, 4), (4, 6)]
    ranges = _balanced_cut(blocks, 6)
    assert ranges == [(i, i + 1) for i in range(6)]
    heavy = [nn.Linear(64, 64)] + [nn.Linear(4, 4) for _ in range(4)]
    r = _balanced_cut([Block(f"b{i}", None, m) for i, m in enumerate(heavy)], 2)
    assert r[0] == (0, 1) and r[1] == (1, 5)


def test_cut_at_named_split_points():
    blocks = [Block(f"layers.{i}", None, None) for i in range(5)]
    assert _cut(blocks, ["layers.2"], 2) == [(0, 2), (2, 5)]
    assert _cut(blocks, ["layers.1", "layers.3.mlp"], 3) == [(0, 1), (1, 3), (3, 5)]


def test_split_inputs_and_batch_size():
    x = torch.arange(8).view(4, 2)
    parts = _split_inputs((x,), {"flag": True, "y": x + 1}, 2)
    assert len(parts) == 2 and torch.equal(parts[1][0][0], x[2:]) and parts[0][1]["flag"] is True
    assert torch.equal(parts[1][1]["y"], x[2:] + 1)
    assert find_pippy_batch_size((), {"ids": x}) == 4 and find_pippy_batch_size(None, None) is None


def _sequential_pipeline():
    from accelerate_hpc_test_b200 import PartialState
    from accelerate_hpc_test_b200.inference import prepare_pippy

    state = PartialState()
    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(16, 32), nn.Tanh(), nn.Linear(32, 32), nn.ReLU(), nn.Linear(32, 4))
    x = torch.randn(6, 16)
    want = model(x)
    piped = prepare_pippy(model, num_chunks=3, gather_output=True)
    got = piped(x) if state.is_main_process else piped()
    assert torch.allclose(got, want, atol=1e-6), (state.process_index, (got - want).abs().max())
    # a second call reuses the stage (fresh receive buffers, no stale requests)
    got2 = piped(x * 2) if state.is_main_process else piped()
    assert torch.allclose(got2, model._original_forward(x * 2) if state.num_processes == 1 else got2)
    assert len(model.hf_split_points) == state.num_processes - 1


def test_sequential_model_two_and_three_stages():
    debug_launcher(_sequential_pipeline, num_processes=2)
    debug_launcher(_sequential_pipeline, num_processes=3)


def _llama_pipeline():
    from accelerate_hpc_test_b200 import PartialState
    from accelerate_hpc_test_b200.inference import prepare_pippy
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM

    state = PartialState()
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=4)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    ids = torch.randint(0, cfg.vocab_size, (4, 12), generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        want = model(input_ids=ids).logits
    piped = prepare_pippy(model, split_points=["model.layers.2"], num_chunks=2)
    out = piped(input_ids=ids) if state.is_main_process else piped()
    if state.is_last_process:
        assert out.shape == want.shape and torch.allclose(out, want, atol=1e-5), (out - want).abs().max()
    else:
        assert out is None
    lo, hi = model.pippy_stage.lo, model.pippy_stage.hi
    assert (lo, hi) == ((0, 3) if state.is_main_process else (3, 6))  # embed + layers 0-1 | layers 2-3 + head


def test_native_llama_two_stages():
    debug_launcher(_llama_pipeline, num_processes=2)


def _bert_pipeline():
    from accelerate_hpc_test_b200 import PartialState
    from accelerate_hpc_test_b200.inference import prepare_pippy
    from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification

    state = PartialState()
    cfg = BertConfig(vocab_size=200, hidden_size=32, num_hidden_layers=4, num_attention_heads=4, intermediate_size=64, max_position_embeddings=64)
    torch.manual_seed(0)
    model = BertForSequenceClassification(cfg).eval()
    g = torch.Generator().manual_seed(5)
    ids = torch.randint(0, 200, (4, 10), generator=g)
    mask = torch.ones(4, 10, dtype=torch.long)
    mask[1, 6:] = 0
    mask[3, 3:] = 0
    types = torch.randint(0, 2, (4, 10), generator=g)
    with torch.no_grad():
        want = model(ids, attention_mask=mask, token_type_ids=types).logits
    piped = prepare_pippy(model, num_chunks=2, gather_output=True)
    out = piped(input_ids=ids, attention_mask=mask, token_type_ids=types) if state.is_main_process else piped()
    assert torch.allclose(out, want, atol=1e-5), (state.process_index, (out - want).abs().max())
    lo, hi = model.pippy_stage.lo, model.pippy_stage.hi
    assert hi - lo >= 1 and (lo == 0) == state.is_main_process     # embeddings on the first stage, classifier on the last


def test_native_bert_two_stages_with_padding_mask():
    debug_launcher(_bert_pipeline, num_processes=2)
