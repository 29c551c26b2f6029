"""Multi-GPU checks of the dim-0-sharded FSDP data path: segment-table all-gather, segment-table reduce-scatter(+AdamW) and the
weight gather that rides on a GEMM (forward: n-gated, dgrad: k-gated, "carry": no gating).

Launch: ``torchrun --nproc-per-node N --master-addr 127.0.0.1 -m accelerate_hpc_test_b200.test_utils.scripts.test_fsdp_gather_gpu``
Oracles are torch.distributed (NCCL) collectives and fp32 matmuls; timings are CUDA events, max over ranks, with the clocks seen.
"""

import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

from accelerate_hpc_test_b200.ops import functional as F
from accelerate_hpc_test_b200.ops._lib import check, cur_stream, ptr
from accelerate_hpc_test_b200.parallel.symm import SymmetricWorkspace
from accelerate_hpc_test_b200.utils.diagnostics import ClockProbe


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp(min=1e-12)).item()


_PROBE = None


class sample_clocks:
    """NVML clock / power / throttle samples taken while the timed launches of the block are in flight (the kernels are far
    shorter than the 200 ms period nvidia-smi polls at)."""

    def __enter__(self):
        global _PROBE
        if _PROBE is None:
            _PROBE = ClockProbe()
        _PROBE.reset()
        return self

    def __exit__(self, *exc):
        return False

    def summary(self):
        return _PROBE.summary()


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
        if _PROBE is not None:
            _PROBE.sample()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def seg_table(rows):
    arr = (ctypes.c_longlong * (3 * len(rows)))()
    for i, r in enumerate(rows):
        arr[3 * i], arr[3 * i + 1], arr[3 * i + 2] = r
    return arr


def check_segs(ws, dev, rank, world):
    """A unit made of three pieces per rank: [rows0, K0] weight, [rows1, K1] weight, flat tail."""
    lib = ws.lib
    shapes = [(world * 96, 256), (world * 40, 512)]
    tail = world * 1000
    pieces = [s[0] // world * s[1] for s in shapes] + [tail // world]
    pieces = [(p + 7) // 8 * 8 for p in pieces]
    shard_n = sum(pieces)
    full_n = shard_n * world
    torch.manual_seed(5)
    full_ref = torch.randn(full_n, device=dev).to(torch.bfloat16)  # same on every rank
    # full layout: segment i at full_off[i], rank q's piece at + q * piece
    full_off, soff, o, so = [], [], 0, 0
    for p in pieces:
        full_off.append(o)
        soff.append(so)
        o += p * world
        so += p
    shard_off, shard = ws.alloc_tensor((shard_n,), torch.bfloat16)
    for i, p in enumerate(pieces):
        shard[soff[i] : soff[i] + p].copy_(full_ref[full_off[i] + rank * p : full_off[i] + (rank + 1) * p])
    torch.cuda.synchronize()
    dist.barrier()
    out = torch.zeros(full_n, device=dev, dtype=torch.bfloat16)
    tab = seg_table([(soff[i] * 2, full_off[i] * 2, pieces[i] * 2) for i in range(len(pieces))])
    check(lib.b200_allgather_pull_segs(ws.peer_array, rank, world, shard_off, tab, len(pieces), ptr(out), -1, 0, 32, cur_stream()), "ag_segs")
    torch.cuda.synchronize()
    assert torch.equal(out, full_ref), "segment all-gather moved wrong data"
    log("allgather_pull_segs ok")

    # reduce-scatter (+AdamW) with the same table: every rank's gradient is full-layout
    goff, grad = ws.alloc_tensor((full_n,), torch.bfloat16)
    g_local = (torch.randn(full_n, device=dev) * 0.1).to(torch.bfloat16)
    grad.copy_(g_local)
    gsum = g_local.float().clone()
    dist.all_reduce(gsum)
    want = torch.cat([gsum[full_off[i] + rank * p : full_off[i] + (rank + 1) * p] for i, p in enumerate(pieces)]) / world
    gout = torch.zeros(shard_n, device=dev, dtype=torch.float32)
    ch_a, ch_d = ws.new_channel(), ws.new_channel()
    rtab = seg_table([(soff[i], full_off[i], pieces[i]) for i in range(len(pieces))])
    torch.cuda.synchronize()
    dist.barrier()
    e = ws.next_epoch(ch_a)
    ws._epochs[ch_d] = e
    check(lib.b200_reduce_scatter_adamw_segs(ws.peer_array, rank, world, rank, goff, rtab, len(pieces), None, None, None, shard_off, ptr(gout),
                                             0.0, 0.9, 0.999, 1e-8, 0.0, 1, 1.0 / world, 0, None, ch_a, ch_d, e, ws.counter_ptr(ch_a), 64,
                                             cur_stream()), "rs_segs")
    torch.cuda.synchronize()
    err = rel(gout, want)
    assert err < 1e-2, ("reduce_scatter_segs", err)
    # fused AdamW on the pieces against the flat fp32 formula
    master = torch.cat([full_ref[full_off[i] + rank * p : full_off[i] + (rank + 1) * p] for i, p in enumerate(pieces)]).float()
    m = torch.zeros_like(master)
    v = torch.zeros_like(master)
    mr, vr, pr = m.clone(), v.clone(), master.clone()
    F.adamw_flat_(pr, mr, vr, want.clone(), None, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.1, step=1)
    grad.copy_(g_local)
    torch.cuda.synchronize()
    dist.barrier()
    e = ws.next_epoch(ch_a)
    ws._epochs[ch_d] = e
    check(lib.b200_reduce_scatter_adamw_segs(ws.peer_array, rank, world, rank, goff, rtab, len(pieces), ptr(master), ptr(m), ptr(v), shard_off, None,
                                             1e-2, 0.9, 0.999, 1e-8, 0.1, 1, 1.0 / world, 1, None, ch_a, ch_d, e, ws.counter_ptr(ch_a), 64,
                                             cur_stream()), "rs_adamw_segs")
    torch.cuda.synchronize()
    assert rel(master, pr) < 1e-3, ("rs_adamw_segs master", rel(master, pr))
    assert rel(shard, pr) < 1e-2, ("rs_adamw_segs shard", rel(shard, pr))
    log(f"reduce_scatter_adamw_segs ok (reduce err {err:.1e})")
    if not ws.has_multicast:
        log("NVLS multicast not available on this workspace: skipping the multimem reduce-scatter / all-reduce checks")
        return
    # the same reduction done by the switch (multimem.ld_reduce): reduce only, then with AdamW
    mc = ctypes.c_void_p(ws.mc_base)
    grad.copy_(g_local)
    gout.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    e = ws.next_epoch(ch_a)
    ws._epochs[ch_d] = e
    check(lib.b200_reduce_scatter_adamw_nvls(ws.peer_array, mc, rank, world, rank, goff, rtab, len(pieces), None, None, None, shard_off, ptr(gout),
                                             0.0, 0.9, 0.999, 1e-8, 0.0, 1, 1.0 / world, 0, None, ch_a, ch_d, e, ws.counter_ptr(ch_a), 64,
                                             cur_stream()), "rs_nvls")
    torch.cuda.synchronize()
    err = rel(gout, want)
    assert err < 1e-2 and (gout - want).abs().max().item() < 2e-2 * want.abs().max().item() + 1e-3, ("reduce_scatter_nvls", err)
    master2 = torch.cat([full_ref[full_off[i] + rank * p : full_off[i] + (rank + 1) * p] for i, p in enumerate(pieces)]).float()
    m2, v2 = torch.zeros_like(master2), torch.zeros_like(master2)
    torch.cuda.synchronize()
    dist.barrier()
    e = ws.next_epoch(ch_a)
    ws._epochs[ch_d] = e
    check(lib.b200_reduce_scatter_adamw_nvls(ws.peer_array, mc, rank, world, rank, goff, rtab, len(pieces), ptr(master2), ptr(m2), ptr(v2), shard_off, None,
                                             1e-2, 0.9, 0.999, 1e-8, 0.1, 1, 1.0 / world, 1, None, ch_a, ch_d, e, ws.counter_ptr(ch_a), 64,
                                             cur_stream()), "rs_adamw_nvls")
    torch.cuda.synchronize()
    assert rel(master2, pr) < 2e-3, ("rs_adamw_nvls master", rel(master2, pr))
    log(f"reduce_scatter_adamw_nvls ok (reduce err {err:.1e})")
    # one-pass NVLS all-reduce, in place
    n = world * 8 * 4099
    aoff, abuf = ws.alloc_tensor((n,), torch.bfloat16)
    src = torch.randn(n, device=dev).to(torch.bfloat16)
    ref = src.float().clone()
    dist.all_reduce(ref)
    ref /= world
    ca, cd = ws.new_channel(), ws.new_channel()
    for it in range(3):
        abuf.copy_(src)
        torch.cuda.synchronize()
        dist.barrier()
        ws.allreduce_nvls_(aoff, n, 1.0 / world, ca, cd)
        torch.cuda.synchronize()
        assert rel(abuf, ref) < 2e-2, ("allreduce_nvls", it, rel(abuf, ref))
    log("allreduce_nvls ok")


class WeightGather:
    """One dim-0-sharded weight [N, K] in symmetric memory + the gathered buffer + counters."""

    def __init__(self, ws, dev, rank, world, N, K, seed):
        self.ws, self.rank, self.world, self.N, self.K = ws, rank, world, N, K
        torch.manual_seed(seed)
        self.w_ref = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)  # identical on every rank
        self.rows = N // world
        self.off, self.piece = ws.alloc_tensor((self.rows, K), torch.bfloat16)
        self.piece.copy_(self.w_ref[rank * self.rows : (rank + 1) * self.rows])
        self.full = torch.zeros(N, K, device=dev, dtype=torch.bfloat16)
        self.ready = torch.zeros(8, dtype=torch.int64, device=dev)
        self.calls = 0

    def reset(self):
        self.full.zero_()  # the gather warp also places this rank's own piece

    def gemm(self, a, b, mode, out=None, b_mn=False, a_mn=False, gathered=None):
        """GEMM(a, b) carrying the gather of this weight (mode 1/2: b is self.full)."""
        self.calls += 1
        if a_mn:
            K, M = a.shape
        else:
            M, K = a.shape
        N = b.shape[1] if b_mn else b.shape[0]
        if out is None:
            out = torch.empty(M, N, device=a.device, dtype=torch.bfloat16)
        piece_bytes = self.rows * self.K * 2
        check(self.ws.lib.b200_gemm2_bf16_gather(
            ptr(a), ptr(b), ptr(out), None, M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn), 0, 0,
            self.ws.peer_array, self.rank, self.world, self.off, ptr(self.full if gathered is None else gathered), piece_bytes, self.rows, mode,
            ptr(self.ready), self.calls * piece_bytes, cur_stream()), "gemm2_gather")
        return out


def check_gemm_gather(ws, dev, rank, world):
    T = 512
    for (N, K) in ((world * 256, 512), (world * 96, 320), (world * 1024, 1024)):
        wg = WeightGather(ws, dev, rank, world, N, K, seed=N + K)
        torch.manual_seed(rank)
        x = torch.randn(T, K, device=dev).to(torch.bfloat16)
        dy = torch.randn(T, N, device=dev).to(torch.bfloat16)
        for it in range(3):
            # forward, n-gated: y = x W^T with W arriving piece by piece
            wg.reset()
            torch.cuda.synchronize()
            dist.barrier()
            y = wg.gemm(x, wg.full, mode=1)
            torch.cuda.synchronize()
            assert torch.equal(wg.full, wg.w_ref), ("mode 1 gathered wrong bytes", N, K, it)
            e = rel(y, x.float() @ wg.w_ref.float().t())
            assert e < 1e-2, ("mode 1 gemm", N, K, it, e)
            assert (y.float() - x.float() @ wg.w_ref.float().t()).abs().max().item() < 0.5, "mode 1 max-abs"
            # dgrad, k-gated: dx = dy W
            wg.reset()
            torch.cuda.synchronize()
            dist.barrier()
            dx = wg.gemm(dy, wg.full, mode=2, b_mn=True)
            torch.cuda.synchronize()
            assert torch.equal(wg.full, wg.w_ref), ("mode 2 gathered wrong bytes", N, K, it)
            e = rel(dx, dy.float() @ wg.w_ref.float())
            assert e < 1e-2, ("mode 2 gemm", N, K, it, e)
            # carry: an unrelated wgrad GEMM moves the weight for a later kernel
            wg.reset()
            torch.cuda.synchronize()
            dist.barrier()
            dw = wg.gemm(dy, x, mode=3, a_mn=True, b_mn=True, gathered=wg.full)
            torch.cuda.synchronize()
            assert torch.equal(wg.full, wg.w_ref), ("mode 3 gathered wrong bytes", N, K, it)
            e = rel(dw, dy.float().t() @ x.float())
            assert e < 1e-2, ("mode 3 gemm", N, K, it, e)
        dist.barrier()
    log("gemm + weight gather (n-gated / k-gated / carry) ok")


def bench_gemm_gather(ws, dev, rank, world, results):
    """Llama-3-8B layer shapes, 4096 tokens per rank: fused gather+GEMM vs NCCL all-gather + cuBLAS, and the plain GEMM."""
    T = 4096
    shapes = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336)}
    for name, (N, K) in shapes.items():
        wg = WeightGather(ws, dev, rank, world, N, K, seed=1)
        wg.reset()
        x = torch.randn(T, K, device=dev).to(torch.bfloat16)
        dy = torch.randn(T, N, device=dev).to(torch.bfloat16)
        full_nccl = torch.empty(N, K, device=dev, dtype=torch.bfloat16)
        y = torch.empty(T, N, device=dev, dtype=torch.bfloat16)
        dx = torch.empty(T, K, device=dev, dtype=torch.bfloat16)
        dw = torch.empty(N, K, device=dev, dtype=torch.bfloat16)

        def nccl_fwd():
            dist.all_gather_into_tensor(full_nccl, wg.piece)
            torch.matmul(x, full_nccl.t(), out=y)

        def nccl_bwd():
            dist.all_gather_into_tensor(full_nccl, wg.piece)
            torch.matmul(dy, full_nccl, out=dx)

        with sample_clocks() as clk:
            t_fwd = timed(lambda: wg.gemm(x, wg.full, mode=1, out=y))
            t_bwd = timed(lambda: wg.gemm(dy, wg.full, mode=2, b_mn=True, out=dx))
            t_carry = timed(lambda: wg.gemm(dy, x, mode=3, a_mn=True, b_mn=True, out=dw, gathered=wg.full))
            t_plain = timed(lambda: F.gemm(x, wg.full, out=y))
            t_plain_w = timed(lambda: F.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=dw))
            t_nccl_f = timed(nccl_fwd)
            t_nccl_b = timed(nccl_bwd)
            t_ag = timed(lambda: dist.all_gather_into_tensor(full_nccl, wg.piece))
        flops = 2.0 * T * N * K
        link = N * K * 2 * (world - 1) / world
        row = dict(op=name, N=N, K=K, tokens=T, world=world, fused_fwd_ms=round(t_fwd, 4), fused_dgrad_ms=round(t_bwd, 4),
                   wgrad_carry_ms=round(t_carry, 4), plain_gemm_ms=round(t_plain, 4), plain_wgrad_ms=round(t_plain_w, 4),
                   nccl_ag_plus_cublas_fwd_ms=round(t_nccl_f, 4), nccl_ag_plus_cublas_dgrad_ms=round(t_nccl_b, 4), nccl_ag_ms=round(t_ag, 4),
                   speedup_fwd=round(t_nccl_f / t_fwd, 3), speedup_dgrad=round(t_nccl_b / t_bwd, 3),
                   gemm_tflops_fwd=round(flops / t_fwd / 1e9, 1), link_gbs_fwd=round(link / t_fwd / 1e6, 1), clocks=clk.summary())
        results.append(row)
        log(json.dumps(row))
        del wg


def bench_reductions(ws, dev, rank, world, results):
    """All-reduce (25 MB DDP bucket, 512 MiB) and reduce-scatter+AdamW of one Llama-3-8B layer: TMA pull, NVLS, NCCL."""
    lib = ws.lib
    for label, nbytes in (("allreduce_25MB", 25 * 2**20), ("allreduce_512MiB", 512 * 2**20)):
        n = nbytes // 2 // (world * 8) * (world * 8)
        off, buf = ws.alloc_tensor((n,), torch.bfloat16)
        buf.normal_()
        chans = [ws.new_channel() for _ in range(3)]
        ref = torch.randn(n, device=dev).to(torch.bfloat16)
        with sample_clocks() as clk:
            t_tma = timed(lambda: ws.allreduce_tma_(off, n, 1.0 / world, chans))
            nv_ctas = int(os.environ.get("NVLS_CTAS", "148"))
            t_nvls = timed(lambda: ws.allreduce_nvls_(off, n, 1.0 / world, chans[0], chans[2], ctas=nv_ctas)) if ws.has_multicast else float("nan")
            t_nccl = timed(lambda: dist.all_reduce(ref))
        bus = lambda t: 2.0 * (world - 1) / world * n * 2 / t / 1e6  # noqa: E731
        row = dict(op=label, world=world, bytes=n * 2, ours_tma_ms=round(t_tma, 4), ours_nvls_ms=round(t_nvls, 4), nccl_ms=round(t_nccl, 4),
                   busbw_tma_gbs=round(bus(t_tma), 1), busbw_nvls_gbs=round(bus(t_nvls), 1), busbw_nccl_gbs=round(bus(t_nccl), 1),
                   nvls_vs_nccl=round(t_nccl / t_nvls, 3), clocks=clk.summary())
        results.append(row)
        log(json.dumps(row))
    # reduce-scatter + AdamW of one decoder layer (218.1 M parameters)
    shard_n = 218_112_000 // world // 8 * 8
    full_n = shard_n * world
    goff, grad = ws.alloc_tensor((full_n,), torch.bfloat16)
    soff, shard = ws.alloc_tensor((shard_n,), torch.bfloat16)
    grad.normal_(std=0.01)
    master = torch.zeros(shard_n, device=dev)
    m, v = torch.zeros_like(master), torch.zeros_like(master)
    ch_a, ch_d = ws.new_channel(), ws.new_channel()
    tab = seg_table([(0, 0, shard_n)])

    def ours(fn_name, *extra):
        def run():
            e = ws.next_epoch(ch_a)
            ws._epochs[ch_d] = e
            fn = getattr(lib, fn_name)
            check(fn(ws.peer_array, *extra, rank, world, rank, goff, tab, 1, ptr(master), ptr(m), ptr(v), soff, None, 1e-4, 0.9, 0.999, 1e-8,
                     0.01, 1, 1.0 / world, 1, None, ch_a, ch_d, e, ws.counter_ptr(ch_a), 148, cur_stream()), fn_name)
        return run

    gfull = torch.randn(full_n, device=dev).to(torch.bfloat16)
    gsh = torch.empty(shard_n, device=dev, dtype=torch.bfloat16)

    def nccl():
        dist.reduce_scatter_tensor(gsh, gfull)
        F.adamw_flat_(master, m, v, gsh, shard, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=1)

    with sample_clocks() as clk:
        t_tma = timed(ours("b200_reduce_scatter_adamw_segs"))
        t_nvls = timed(ours("b200_reduce_scatter_adamw_nvls", ctypes.c_void_p(ws.mc_base))) if ws.has_multicast else float("nan")
        t_nccl = timed(nccl)
    row = dict(op="reduce_scatter+adamw one layer", world=world, params=full_n, ours_tma_ms=round(t_tma, 4), ours_nvls_ms=round(t_nvls, 4),
               nccl_rs_plus_adamw_kernel_ms=round(t_nccl, 4), nvls_vs_nccl=round(t_nccl / t_nvls, 3), tma_vs_nccl=round(t_nccl / t_tma, 3),
               clocks=clk.summary())
    results.append(row)
    log(json.dumps(row))


def main():
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    ws = SymmetricWorkspace(dist.group.WORLD, int(6 * 2**30), dev)
    log(f"symmetric workspace backend: {ws.backend}, multicast: {ws.has_multicast}")
    check_segs(ws, dev, rank, world)
    check_gemm_gather(ws, dev, rank, world)
    results = []
    if "--bench" in sys.argv:
        bench_reductions(ws, dev, rank, world, results)
        bench_gemm_gather(ws, dev, rank, world, results)
        if rank == 0:
            os.makedirs("gpurun_out", exist_ok=True)
            with open(f"gpurun_out/fsdp_gather_n{world}.json", "w") as f:
                json.dump(results, f, indent=1)
    torch.cuda.synchronize()
    dist.barrier()
    ws.close()
    log("ALL OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
