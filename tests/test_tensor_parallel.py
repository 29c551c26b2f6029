"""Tensor parallelism on 2 CPU (gloo) ranks: sharded Llama vs the unsharded model — loss, replicated-parameter
gradients after the TP all-reduce hook, and the local slices of column/row-sharded weight gradients
(reference: accelerator.py:1592-1607 `_prepare_tp`, tests/tp/)."""

import torch

from accelerate_hpc_test_b200.launchers import debug_launcher


def _tp_parity():
    import torch.distributed as dist
    from accelerate_hpc_test_b200 import PartialState
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.parallel.tensor_parallel import apply_tensor_parallel
    st = PartialState()
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2)
    torch.manual_seed(0); ref = LlamaForCausalLM(cfg, dtype=torch.float32)
    torch.manual_seed(0); model = LlamaForCausalLM(cfg, dtype=torch.float32)
    model = apply_tensor_parallel(model, dist.group.WORLD)
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, cfg.vocab_size, (2, 16), generator=g)
    lab = ids.clone(); lab[0, :5] = -100
    out = model(ids, labels=lab); out.loss.backward()
    r = ref(ids, labels=lab); r.loss.backward()
    assert abs(out.loss.item() - r.loss.item()) < 1e-5, (out.loss.item(), r.loss.item())
    # replicated params: grads equal after the hook all-reduce
    for n in ["model.norm.weight", "model.embed_tokens.weight", "lm_head.weight", "model.layers.0.input_layernorm.weight"]:
        a = dict(model.named_parameters())[n].grad; b = dict(ref.named_parameters())[n].grad
        assert torch.allclose(a, b, atol=1e-5), (n, (a-b).abs().max())
    # sharded params: compare the local slice
    w = model.model.layers[1].self_attn.o_proj.weight.grad
    wr = ref.model.layers[1].self_attn.o_proj.weight.grad
    n = wr.shape[1] // 2
    assert torch.allclose(w, wr[:, st.process_index*n:(st.process_index+1)*n], atol=1e-5)
    gu = model.model.layers[0].mlp.gate_up_proj.weight.grad; gur = ref.model.layers[0].mlp.gate_up_proj.weight.grad
    I = cfg.intermediate_size; h = I // 2; k = st.process_index
    assert torch.allclose(gu, torch.cat([gur[:I][k*h:(k+1)*h], gur[I:][k*h:(k+1)*h]]), atol=1e-5)


def test_tensor_parallel_llama_two_processes():
    debug_launcher(_tp_parity, num_processes=2)


def _nd_parallel_parity():
    """ParallelismConfig(tp_size=2, dp_shard_size=2): tensor-parallel model inside the sharded engine (BASELINE config 4,
    reference examples/torch_native_parallelism/nd_parallel.py) against a single-process replay with the global batch."""
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2)
    torch.manual_seed(0); ref = LlamaForCausalLM(cfg, dtype=torch.float32)
    torch.manual_seed(0); model = LlamaForCausalLM(cfg, dtype=torch.float32)
    pc = ParallelismConfig(tp_size=2, dp_shard_size=2)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap", transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(parallelism_config=pc, fsdp_plugin=plugin)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.0)
    model, opt = acc.prepare(model, opt)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-3, weight_decay=0.0)
    g = torch.Generator().manual_seed(3)
    ids_all = torch.randint(0, cfg.vocab_size, (4, 16), generator=g)
    dp_rank = acc.torch_device_mesh.get_local_rank("dp_shard")
    ids = ids_all[dp_rank*2:(dp_rank+1)*2]
    for step in range(2):
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step(); opt.zero_grad()
        ropt.zero_grad()
        for r in range(2):
            sub = ids_all[r*2:(r+1)*2]
            (ref(input_ids=sub, labels=sub).loss / 2).backward()
        ropt.step()
    rl = ref(input_ids=ids, labels=ids).loss.item()
    ml = model(input_ids=ids, labels=ids).loss.item()
    assert abs(ml - rl) < 2e-4 * abs(rl), (ml, rl)


def test_nd_parallel_tp2_x_fsdp2_four_processes():
    debug_launcher(_nd_parallel_parity, num_processes=4)


def _transformers_llama_under_tp():
    """A transformers Llama handed to ``prepare`` under tp=2: it is rebuilt as the native model (same weights), the optimizer that was
    created on the transformers parameters — two groups, decay / no decay — is re-pointed, and training matches transformers on one
    process (the reference accepts a ``tp_plan="auto"`` transformers model here)."""
    from transformers import LlamaConfig as HFConfig
    from transformers import LlamaForCausalLM as HFLlama

    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig

    hf_cfg = HFConfig(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=64, rope_theta=10000.0)

    def groups(m):
        decay = [p for n, p in m.named_parameters() if p.dim() > 1]
        rest = [p for n, p in m.named_parameters() if p.dim() <= 1]
        return [{"params": decay, "weight_decay": 0.1}, {"params": rest, "weight_decay": 0.0}]

    torch.manual_seed(0)
    reference = HFLlama(hf_cfg)
    torch.manual_seed(0)
    model = HFLlama(hf_cfg)
    acc = Accelerator(parallelism_config=ParallelismConfig(tp_size=2))
    opt = torch.optim.AdamW(groups(model), lr=1e-3)
    ropt = torch.optim.AdamW(groups(reference), lr=1e-3)
    model, opt = acc.prepare(model, opt)
    native = acc.unwrap_model(model)
    assert type(native).__name__ == "LlamaForCausalLM" and type(native).__module__.startswith("accelerate_hpc_test_b200")
    inner = opt.optimizer
    assert [g["weight_decay"] for g in inner.param_groups] == [0.1, 0.0]
    assert sum(len(g["params"]) for g in inner.param_groups) > 0
    ids = torch.randint(0, 128, (2, 16), generator=torch.Generator().manual_seed(3))
    for step in range(3):
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()
        opt.zero_grad()
        ref_loss = reference(input_ids=ids, labels=ids).loss
        ref_loss.backward()
        ropt.step()
        ropt.zero_grad()
        assert abs(out.loss.item() - ref_loss.item()) < 2e-4 * abs(ref_loss.item()), (step, out.loss.item(), ref_loss.item())


def test_transformers_llama_is_adopted_for_tensor_parallelism():
    debug_launcher(_transformers_llama_under_tp, num_processes=2)


def test_from_transformers_refuses_what_the_native_kernels_do_not_implement():
    import pytest
    from transformers import LlamaConfig as HFConfig
    from transformers import LlamaForCausalLM as HFLlama

    from accelerate_hpc_test_b200.models import LlamaForCausalLM

    small = dict(vocab_size=64, hidden_size=32, intermediate_size=64, num_hidden_layers=1, num_attention_heads=2, num_key_value_heads=1,
                 max_position_embeddings=32)
    tied = HFLlama(HFConfig(**small, tie_word_embeddings=True))
    native = LlamaForCausalLM.from_transformers(tied)
    assert native.lm_head.weight is native.model.embed_tokens.weight
    ids = torch.randint(0, 64, (1, 8))
    with torch.no_grad():
        assert torch.allclose(native(input_ids=ids).logits, tied(input_ids=ids).logits, atol=1e-5)
    with pytest.raises(NotImplementedError, match="attention_bias"):
        LlamaForCausalLM.from_transformers(HFLlama(HFConfig(**small, attention_bias=True)))
    with pytest.raises(NotImplementedError, match="only Llama-architecture"):
        LlamaForCausalLM.from_transformers(torch.nn.Linear(2, 2).__class__.__mro__[0].__new__(torch.nn.Linear) if False else type("M", (), {"config": type("C", (), {"model_type": "gpt2"})()})())
