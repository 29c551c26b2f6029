"""Ulysses sequence parallelism and context parallelism on 2 CPU (gloo) ranks against the unsharded model
(reference: accelerator.py:2358-2410 Ulysses wiring, :1618-1680 context parallel; tests/test_dataclasses + examples
torch_native_parallelism/nd_parallel.py for the user-facing flow)."""

import torch

from accelerate_hpc_test_b200.launchers import debug_launcher


def _models():
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2,
                           num_attention_heads=4, num_key_value_heads=2)
    torch.manual_seed(0)
    ref = LlamaForCausalLM(cfg, dtype=torch.float32)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    return cfg, ref, model


def _batch(cfg, S=16):
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, cfg.vocab_size, (2, S), generator=g)
    lab = ids.clone()
    lab[0, :5] = -100
    return ids, lab


def _check_grads(model, ref, world):
    import torch.distributed as dist

    for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
        g = p.grad.clone()
        dist.all_reduce(g)
        g /= world  # what the data-parallel engine does over the group holding the sequence shards
        assert torch.allclose(g, q.grad, atol=2e-5, rtol=1e-4), (n, (g - q.grad).abs().max().item())


def _ulysses_parity():
    import torch.distributed as dist
    from accelerate_hpc_test_b200 import PartialState
    from accelerate_hpc_test_b200.parallel.ulysses import UlyssesDataLoaderAdapter, apply_ulysses

    st = PartialState()
    cfg, ref, model = _models()
    model = apply_ulysses(model, dist.group.WORLD)
    ids, lab = _batch(cfg)
    adapter = UlyssesDataLoaderAdapter([{"input_ids": ids, "labels": lab}], dist.group.WORLD)
    (batch,) = list(adapter)
    assert batch["input_ids"].shape == (2, 8) and "shift_labels" in batch and "labels" not in batch
    assert batch["position_ids"][0, 0].item() == 8 * st.process_index
    out = model(**batch)
    out.loss.backward()
    r = ref(ids, labels=lab)
    r.loss.backward()
    assert abs(out.loss.item() - r.loss.item()) < 1e-5, (out.loss.item(), r.loss.item())
    _check_grads(model, ref, 2)
    # plain `labels` on a sequence slice: the model fetches the boundary label from the next rank
    model.zero_grad()
    c = 8
    sl = slice(st.process_index * c, (st.process_index + 1) * c)
    out2 = model(ids[:, sl], labels=lab[:, sl])
    assert abs(out2.loss.item() - r.loss.item()) < 1e-5


def _cp_parity():
    import torch.distributed as dist
    from accelerate_hpc_test_b200 import PartialState
    from accelerate_hpc_test_b200.parallel.context_parallel import (apply_context_parallel, context_parallel_buffers,
                                                                     shard_sequence, unshard_sequence)

    st = PartialState()
    cfg, ref, model = _models()
    model = apply_context_parallel(model, dist.group.WORLD)
    ids, lab = _batch(cfg)
    full = torch.arange(16).unsqueeze(0)
    parts = [shard_sequence(full, 1, r, 2) for r in range(2)]
    assert parts[0].tolist() == [[0, 1, 2, 3, 12, 13, 14, 15]] and torch.equal(unshard_sequence(parts, 1), full)
    shift = torch.full_like(lab, -100)
    shift[:, :-1] = lab[:, 1:]
    ids_l, shift_l = ids.clone(), shift.clone()
    with context_parallel_buffers(dist.group.WORLD, [ids_l, shift_l], [1, 1], no_restore_buffers={shift_l}):
        assert ids_l.shape == (2, 8)
        out = model(ids_l, shift_labels=shift_l)
        out.loss.backward()
    assert ids_l.shape == (2, 16) and torch.equal(ids_l, ids) and shift_l.shape == (2, 8)
    r = ref(ids, labels=lab)
    r.loss.backward()
    assert abs(out.loss.item() - r.loss.item()) < 1e-5, (out.loss.item(), r.loss.item())
    _check_grads(model, ref, 2)
    model.zero_grad()
    out2 = model(shard_sequence(ids, 1, st.process_index, 2), labels=shard_sequence(lab, 1, st.process_index, 2))
    assert abs(out2.loss.item() - r.loss.item()) < 1e-5


def _accelerator_sp():
    """End to end through Accelerator + ParallelismConfig(sp_size=2): sharded engine over the sp axis."""
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    cfg, ref, model = _models()
    pc = ParallelismConfig(sp_size=2, sp_backend="b200")
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(parallelism_config=pc, fsdp_plugin=plugin)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    ids, lab = _batch(cfg)
    dl = torch.utils.data.DataLoader([{"input_ids": ids[i], "labels": lab[i]} for i in range(2)], batch_size=2)
    model, opt, dl = acc.prepare(model, opt, dl)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-3)
    for batch in dl:
        assert batch["input_ids"].shape == (2, 8)
        out = model(**batch)
        acc.backward(out.loss)
        opt.step()
        opt.zero_grad()
    r = ref(ids, labels=lab)
    r.loss.backward()
    ropt.step()
    assert abs(out.loss.item() - r.loss.item()) < 1e-5
    sd = model.full_state_dict()
    worst = max(((sd[n] - p).norm() / p.norm()).item() for n, p in ref.named_parameters())
    assert worst < 1e-4, worst


def test_ulysses_two_processes():
    debug_launcher(_ulysses_parity, num_processes=2)


def test_context_parallel_two_processes():
    debug_launcher(_cp_parity, num_processes=2)


def test_accelerator_sequence_parallel_two_processes():
    debug_launcher(_accelerator_sp, num_processes=2)
