"""Full vs regional ``torch.compile`` of a *user* model made of repeated blocks: cold-start compile time and steady-state forward time.
Counterpart of the reference's ``benchmarks/torch.compile/regional_compilation.py``.  (The framework's own models are hand-written
kernels and are never compiled; this measures the option offered for stock ``nn.Module`` models.)

    python benchmarks/regional_compilation.py --layers 12 --hidden 512                 # inductor, GPU if there is one
    python benchmarks/regional_compilation.py --backend eager --layers 4 --cpu         # structure check without code generation
"""

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch import nn


class Block(nn.Module):
    def __init__(self, hidden):
        super().__init__()
        self.norm1, self.norm2 = nn.LayerNorm(hidden), nn.LayerNorm(hidden)
        self.attn = nn.MultiheadAttention(hidden, num_heads=max(1, hidden // 64), batch_first=True)
        self.mlp = nn.Sequential(nn.Linear(hidden, 4 * hidden), nn.GELU(), nn.Linear(4 * hidden, hidden))

    def forward(self, x):
        h = self.norm1(x)
        x = x + self.attn(h, h, h, need_weights=False)[0]
        return x + self.mlp(self.norm2(x))


class Stack(nn.Module):
    def __init__(self, layers, hidden, vocab=1024):
        super().__init__()
        self.embed = nn.Embedding(vocab, hidden)
        self.layers = nn.ModuleList(Block(hidden) for _ in range(layers))
        self.head = nn.Linear(hidden, vocab)

    def forward(self, ids):
        x = self.embed(ids)
        for layer in self.layers:
            x = layer(x)
        return self.head(x)


def measure(make, ids, iters, sync):
    torch._dynamo.reset()
    model = make()
    t0 = time.perf_counter()
    with torch.no_grad():
        model(ids)
    sync()
    cold = time.perf_counter() - t0
    with torch.no_grad():
        for _ in range(3):
            model(ids)
        sync()
        t0 = time.perf_counter()
        for _ in range(iters):
            model(ids)
        sync()
    return cold, (time.perf_counter() - t0) / iters


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    p.add_argument("--layers", type=int, default=8)
    p.add_argument("--hidden", type=int, default=256)
    p.add_argument("--batch", type=int, default=4)
    p.add_argument("--seq-len", type=int, default=128)
    p.add_argument("--iters", type=int, default=10)
    p.add_argument("--backend", default="inductor")
    p.add_argument("--cpu", action="store_true")
    args = p.parse_args(argv)

    from accelerate_hpc_test_b200.utils import compile_regions

    device = torch.device("cuda" if torch.cuda.is_available() and not args.cpu else "cpu")
    sync = torch.cuda.synchronize if device.type == "cuda" else (lambda: None)
    torch.manual_seed(0)
    base = Stack(args.layers, args.hidden).to(device).eval()
    ids = torch.randint(0, 1024, (args.batch, args.seq_len), device=device)
    with torch.no_grad():
        want = base(ids)

    rows = {}
    for name, make in [("eager", lambda: base), ("full", lambda: torch.compile(base, backend=args.backend)),
                       ("regional", lambda: compile_regions(base, backend=args.backend))]:
        cold, steady = measure(make, ids, args.iters, sync)
        rows[name] = {"first_call_s": round(cold, 3), "steady_ms": round(steady * 1e3, 3)}
    regional = compile_regions(base, backend=args.backend)
    with torch.no_grad():
        err = (regional(ids) - want).abs().max().item()
    record = {"backend": args.backend, "device": device.type, "layers": args.layers, "hidden": args.hidden, **rows,
              "compile_time_ratio_full_over_regional": round(rows["full"]["first_call_s"] / max(rows["regional"]["first_call_s"], 1e-9), 2),
              "regional_max_abs_err": err, "regional_keeps_original": regional._orig_mod is base}
    print(json.dumps(record), flush=True)
    assert err < 1e-3 and record["regional_keeps_original"]
    return record


if __name__ == "__main__":
    main()
