"""Two GPUs: a ``transformers`` Llama handed to ``Accelerator.prepare`` under ``ParallelismConfig(tp_size=2)`` and *no* sharding plugin.

Checks, on the GPU kernels: the model is rebuilt as the native Llama, the tensor-parallel layers run the fused all-gather -> GEMM /
GEMM -> reduce-scatter kernels inside the sharded engine with a shard group of one (no data-parallel collectives), the two-group
AdamW created on the transformers parameters is re-pointed, and three training steps follow the un-parallel transformers model
(bf16, so a few percent).  Launched by ``tests/test_multigpu.py`` with torchrun.
"""

import os

import torch
import torch.distributed as dist


def main():
    from transformers import LlamaConfig as HFConfig
    from transformers import LlamaForCausalLM as HFLlama

    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.ops import _lib
    from accelerate_hpc_test_b200.parallelism_config import ParallelismConfig

    acc = Accelerator(mixed_precision="bf16", parallelism_config=ParallelismConfig(tp_size=2))
    assert acc.num_processes == 2 and acc.device.type == "cuda"
    hf_cfg = HFConfig(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=512, rope_theta=10000.0)

    def groups(m):
        decay = [p for _, p in m.named_parameters() if p.dim() > 1]
        rest = [p for _, p in m.named_parameters() if p.dim() <= 1]
        return [{"params": decay, "weight_decay": 0.1}, {"params": rest, "weight_decay": 0.0}]

    torch.manual_seed(0)
    reference = HFLlama(hf_cfg).to(device=acc.device, dtype=torch.bfloat16)
    torch.manual_seed(0)
    model = HFLlama(hf_cfg).to(dtype=torch.bfloat16)
    opt = torch.optim.AdamW(groups(model), lr=1e-3)
    ropt = torch.optim.AdamW(groups(reference), lr=1e-3)
    model, opt = acc.prepare(model, opt)
    native = acc.unwrap_model(model)
    assert type(native).__module__.startswith("accelerate_hpc_test_b200") and type(model).__name__ == "B200ShardedModule", type(model)
    assert [g["weight_decay"] for g in opt.optimizer.param_groups] == [0.1, 0.0]
    ids = torch.randint(0, 512, (2, 256), generator=torch.Generator().manual_seed(3)).to(acc.device)
    _lib.reset_launch_count()
    ours, theirs = [], []
    for _ in range(3):
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()
        opt.zero_grad()
        ref_loss = reference(input_ids=ids, labels=ids).loss
        ref_loss.backward()
        ropt.step()
        ropt.zero_grad()
        ours.append(out.loss.item())
        theirs.append(ref_loss.item())
    torch.cuda.synchronize()
    assert _lib.launch_count() > 0, "the native kernels did not run"
    for step, (a, b) in enumerate(zip(ours, theirs)):
        assert abs(a - b) < 3e-2 * abs(b), (step, ours, theirs)
    assert ours[-1] < ours[0]
    acc.print(f"tp adoption on GPUs ok: losses {ours} vs transformers {theirs}; native launches {_lib.launch_count()}")
    acc.print("TP ADOPTION GPU CHECK PASSED")
    acc.end_training()
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")
    main()
