"""Single-GPU program that launches every hot kernel of the training step exactly once inside a profiler range — the target of
``ncu --set full --profile-from-start off``.  Shapes are the Llama-3-8B per-layer shapes of the bench (4096 tokens).

The multi-GPU kernels are launched on a one-rank group (their peer tables point at this GPU's own memory): that shows their
instruction mix, occupancy and memory pipeline, but NVLink counters stay at zero — a multi-rank program may not run under ncu.

    python benchmarks/ncu_targets.py                      # plain run first (must exit 0)
    ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/prof python benchmarks/ncu_targets.py
"""
import ctypes
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from accelerate_hpc_test_b200.ops import functional as F
from accelerate_hpc_test_b200.parallel.precision import gemm_mxfp8, quantize_mx

T, H, I, V = 4096, 4096, 14336, 128256
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
bf = torch.bfloat16
x = torch.randn(T, H, device=dev, dtype=bf)
w = torch.randn(2 * I, H, device=dev, dtype=bf) * 0.02
dy = torch.randn(T, 2 * I, device=dev, dtype=bf)
B, S, Hq, Hkv, D = 2, 2048, 32, 8, 128
qkv = (torch.randn(B * S, (Hq + 2 * Hkv) * D, device=dev) * 0.5).to(bf)
scale = 1 / math.sqrt(D)
o, lse = F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale)
do = torch.randn_like(o)
wn = torch.ones(H, device=dev, dtype=bf)
res = torch.randn(T, H, device=dev, dtype=bf)
y, s_, rstd = F.rmsnorm_fwd(x, wn, 1e-5, residual=res)
gu = torch.randn(T, 2 * I, device=dev, dtype=bf)
act = F.swiglu_fwd(gu)
cos, sin = F.rope_tables(S, D, 500000.0, dev)
pos = torch.arange(S, device=dev, dtype=torch.int32).repeat(B)
n = 4096 * 14336
master = torch.randn(n, device=dev)
m1, v1 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
g = torch.randn(n, device=dev).to(bf)
sh = torch.empty(n, device=dev, dtype=bf)
hh = (torch.randn(T, H, device=dev) * 0.5).to(bf).requires_grad_(True)
wv = (torch.randn(V, H, device=dev) * 0.02).to(bf).requires_grad_(True)
labels = torch.randint(0, V, (T,), device=dev)
xq, wq = quantize_mx(x), quantize_mx(w)

# one-rank "group" for the peer-memory kernels
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29577")
dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
from accelerate_hpc_test_b200.parallel.symm import SymmetricWorkspace  # noqa: E402

layer = 218_112_000 // 8 * 8
ws = SymmetricWorkspace(dist.group.WORLD, layer * 2 * 2 + w.numel() * 2 + 64 * 2**20, dev)
soff, shard = ws.alloc_tensor((layer,), bf)
goff, grads = ws.alloc_tensor((layer,), bf)
shard.normal_()
grads.normal_()
full = torch.empty(layer, dtype=bf, device=dev)
lm, lm1, lm2 = torch.randn(layer, device=dev), torch.zeros(layer, device=dev), torch.zeros(layer, device=dev)
ch = [ws.new_channel() for _ in range(4)]
lib = ws.lib
segs = (ctypes.c_longlong * 3)(0, 0, layer)


def rs_adamw():
    e = ws.next_epoch(ch[0])
    ws._epochs[ch[1]] = e
    rc = lib.b200_reduce_scatter_adamw_segs(ws.peer_array, 0, 1, 0, goff, segs, 1, ctypes.c_void_p(lm.data_ptr()), ctypes.c_void_p(lm1.data_ptr()),
                                            ctypes.c_void_p(lm2.data_ptr()), soff, None, 1e-3, 0.9, 0.999, 1e-8, 0.01, 3, 1.0, 1, None,
                                            ch[0], ch[1], e, ws.counter_ptr(ch[0]), 148, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, rc


woff, wshard = ws.alloc_tensor(tuple(w.shape), bf)
wshard.copy_(w)
wfull = torch.empty_like(w)
ready = torch.zeros(8, dtype=torch.int64, device=dev)
calls = [0]
yg = torch.empty(T, 2 * I, dtype=bf, device=dev)


def gemm_gather():
    """Forward GEMM whose gather warp brings the weight rows in (mode 1: the MMA of an n-tile waits for that tile's rows)."""
    calls[0] += 1
    rc = lib.b200_gemm2_bf16_gather(ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(wfull.data_ptr()), ctypes.c_void_p(yg.data_ptr()), None,
                                    T, 2 * I, H, H, H, 2 * I, 0, 0, 0, 0, ws.peer_array, 0, 1, woff, ctypes.c_void_p(wfull.data_ptr()),
                                    w.numel() * 2, 2 * I, 1, ctypes.c_void_p(ready.data_ptr()), calls[0] * w.numel() * 2,
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, rc


def ops():
    yield "gemm fwd", lambda: F.gemm(x, w)
    yield "gemm dgrad", lambda: F.gemm(dy, w, b_mn_major=True)
    yield "gemm wgrad", lambda: F.gemm(dy, x, a_mn_major=True, b_mn_major=True)
    yield "flash fwd", lambda: F.flash_attn_fwd(qkv, B, S, Hq, Hkv, D, True, scale)
    yield "flash bwd", lambda: F.flash_attn_bwd(do, qkv, o, lse, B, S, Hq, Hkv, D, True, scale)
    yield "rmsnorm fwd", lambda: F.rmsnorm_fwd(x, wn, 1e-5, residual=res)
    yield "rmsnorm bwd", lambda: F.rmsnorm_bwd(y, s_, wn, rstd, dres=res)
    yield "swiglu fwd", lambda: F.swiglu_fwd(gu)
    yield "swiglu bwd", lambda: F.swiglu_bwd(gu, act)
    yield "rope", lambda: F.rope_inplace(qkv, cos, sin, pos, Hq + Hkv, D)
    yield "adamw", lambda: F.adamw_flat_(master, m1, v1, g, sh, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=3)
    yield "lm head + CE", lambda: F.fused_linear_cross_entropy(hh, wv, labels, -100, 4096).backward()
    yield "mx quantize", lambda: quantize_mx(x)
    yield "mxfp8 gemm", lambda: gemm_mxfp8(xq, wq)
    yield "gemm fwd + weight gather (1 rank)", gemm_gather
    yield "all-gather pull (1 rank)", lambda: ws.allgather_pull(soff, layer * 2, full, ctas=148)
    yield "reduce-scatter + adamw (1 rank)", rs_adamw


# TARGETS=tc: tensor-core kernels only; TARGETS=mem: the bandwidth-bound and comm kernels; default: all (a full-set report
# of all of them with sources imported is ~85 MB — more than one gpurun call brings back)
which = os.environ.get("TARGETS", "all")
TC = ("gemm", "flash", "mxfp8", "lm head")
selected = [(n_, f_) for n_, f_ in ops() if which == "all" or (which == "tc") == n_.startswith(TC)]
for name, fn in selected:
    for _ in range(2):
        fn()
torch.cuda.synchronize()
torch.cuda.profiler.start()
for name, fn in selected:
    fn()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
dist.destroy_process_group()
