"""2-process CPU (gloo) runs of the in-package distributed scripts through ``debug_launcher`` — the
reference's "cluster on a laptop" pattern (reference tests/test_cpu.py:21-27, tests/test_grad_sync.py)."""

import torch

from accelerate_hpc_test_b200.launchers import debug_launcher
from accelerate_hpc_test_b200.test_utils.scripts import test_ops, test_sync


def test_ops_two_processes():
    debug_launcher(test_ops.main, num_processes=2)


def test_grad_sync_two_processes():
    debug_launcher(test_sync.main, num_processes=2)


def test_ops_single_process():
    test_ops.main()


def _fsdp_parity():
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.optim import FusedAdamW, grad_global_norm
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=3)
    torch.manual_seed(0)
    ref = LlamaForCausalLM(cfg, dtype=torch.float32)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(fsdp_plugin=plugin)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    assert type(model).__name__ == "B200ShardedModule" and acc.is_fsdp2
    ropt = FusedAdamW(ref.parameters(), lr=1e-3, weight_decay=0.01)
    g = torch.Generator().manual_seed(7)
    for step in range(3):
        ids_all = torch.randint(0, cfg.vocab_size, (4, 16), generator=g)
        ids = ids_all[acc.process_index * 2 : acc.process_index * 2 + 2]
        opt.zero_grad()
        with acc.accumulate(model):
            out = model(input_ids=ids, labels=ids)
            acc.backward(out.loss)
            norm = acc.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
        ropt.zero_grad()
        for r in range(2):
            sub = ids_all[r * 2 : r * 2 + 2]
            (ref(input_ids=sub, labels=sub).loss / 2).backward()
        rnorm, coef, _ = grad_global_norm(ref.parameters(), 1.0)
        ropt.set_clip_coef(coef)
        ropt.step()
        assert abs(float(norm) - float(rnorm)) < 1e-4 * float(rnorm)
    sd = model.full_state_dict()
    worst = max(((sd[n] - p).norm() / p.norm()).item() for n, p in ref.named_parameters())
    assert worst < 1e-5, worst


def test_sharded_engine_matches_single_process_two_processes():
    debug_launcher(_fsdp_parity, num_processes=2)


def _fsdp_checkpoint(tmpdir):
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(fsdp_plugin=plugin)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    model, opt = acc.prepare(model, opt)
    ids = torch.randint(0, cfg.vocab_size, (2, 16), generator=torch.Generator().manual_seed(acc.process_index))

    def step():
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()
        return out.loss.item()

    step()
    acc.save_state(tmpdir)
    a = [step() for _ in range(2)]
    after = model.full_state_dict()
    acc.load_state(tmpdir)
    b = [step() for _ in range(2)]
    again = model.full_state_dict()
    assert a == b, (a, b)
    for k in after:
        assert torch.equal(after[k], again[k]), k
    import os

    if acc.is_main_process:
        assert os.path.isdir(os.path.join(tmpdir, "pytorch_model_fsdp_0")) and os.path.isdir(os.path.join(tmpdir, "optimizer_0"))
        from accelerate_hpc_test_b200.parallel.fsdp import merge_sharded_checkpoint

        merged = merge_sharded_checkpoint(os.path.join(tmpdir, "pytorch_model_fsdp_0"))
        assert set(merged) == set(n for n, _ in model.module.named_parameters())


def test_sharded_checkpoint_resume_two_processes(tmp_path):
    debug_launcher(_fsdp_checkpoint, args=(str(tmp_path),), num_processes=2)


def _gathered_parameters():
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin
    from accelerate_hpc_test_b200.utils.operations import GatheredParameters

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    ref = {n: p.detach().clone() for n, p in model.named_parameters()}
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(fsdp_plugin=plugin)
    model = acc.prepare(model)
    inner = acc.unwrap_model(model)
    w = inner.model.layers[1].mlp.down_proj.weight
    assert w.numel() == 0  # sharded away between uses
    with GatheredParameters([w], modifier_rank=0):
        assert torch.equal(w, ref["model.layers.1.mlp.down_proj.weight"])
        w.data.mul_(2.0)  # every rank applies the same modification; it is written back to the shards
    assert w.numel() == 0
    with GatheredParameters([w]):
        assert torch.allclose(w, 2.0 * ref["model.layers.1.mlp.down_proj.weight"])


def test_gathered_parameters_two_processes():
    debug_launcher(_gathered_parameters, num_processes=2)
