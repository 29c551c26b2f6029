"""Step time of a Llama-3-8B-shaped model (LAYERS env, default 8) with the FusedAdamW update inside backward vs after it."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW

acc = Accelerator(mixed_precision="bf16")
cfg = LlamaConfig.llama3_8b(num_hidden_layers=int(os.environ.get("LAYERS", 8)))
model = LlamaForCausalLM(cfg, dtype=torch.bfloat16, device=acc.device)
opt = FusedAdamW(model.parameters(), lr=3e-5, weight_decay=0.01)
model, opt = acc.prepare(model, opt)
ids = torch.randint(0, cfg.vocab_size, (2, 2048), device=acc.device)


def step():
    opt.zero_grad()
    out = model(input_ids=ids, labels=ids)
    acc.backward(out.loss)
    opt.step()


for i in range(3):
    step()
    print("step", i, "mode", opt.optimizer._bwd_mode, flush=True)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(5):
    step()
e.record()
torch.cuda.synchronize()
print(f"mode={opt.optimizer._bwd_mode} env={os.environ.get('ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD', 'auto')}: {s.elapsed_time(e) / 5:.2f} ms/step")
