"""Engine-style (Megatron-LM) API on the native stack: ``engine(**batch)`` is a full optimisation step over
micro-batches (reference utils/megatron_lm.py:924-1158, tests rely on examples/by_feature/megatron_lm_gpt_pretraining.py)."""

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import (GPTTrainStep, MegatronEngine, MegatronLMOptimizerWrapper, MegatronLMSchedulerWrapper,
                                            avg_losses_across_data_parallel_group)


def test_engine_step_equals_manual_accumulation():
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    torch.manual_seed(0)
    a = LlamaForCausalLM(cfg, dtype=torch.float32)
    torch.manual_seed(0)
    b = LlamaForCausalLM(cfg, dtype=torch.float32)
    acc = Accelerator(cpu=True)
    oa = FusedAdamW(a.parameters(), lr=1e-2)
    ob = FusedAdamW(b.parameters(), lr=1e-2)
    sched = torch.optim.lr_scheduler.LambdaLR(oa, lambda s: 1.0 / (1 + s))
    engine = MegatronEngine(acc, a, MegatronLMOptimizerWrapper(oa), MegatronLMSchedulerWrapper(sched, oa))
    engine.num_micro_batches = 2
    assert isinstance(engine.train_step_handler, GPTTrainStep)
    ids = torch.randint(0, cfg.vocab_size, (4, 16))
    engine.train()
    loss = engine(input_ids=ids)
    # manual: two micro-batches, averaged
    ob.zero_grad()
    ref = 0.0
    for i in range(2):
        out = b(ids[2 * i : 2 * i + 2], labels=ids[2 * i : 2 * i + 2])
        (out.loss / 2).backward()
        ref += out.loss.item() / 2
    ob.step()
    assert abs(float(loss) - ref) < 1e-5
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)
    assert abs(oa.param_groups[0]["lr"] - 0.5e-2) < 1e-9  # the engine stepped the schedule
    engine.eval()
    ev = engine(input_ids=ids)
    assert torch.isfinite(ev) and engine.log_eval_results()
    assert avg_losses_across_data_parallel_group([torch.tensor(2.0)]).item() == 2.0
