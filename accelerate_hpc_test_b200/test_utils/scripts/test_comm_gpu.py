"""Multi-GPU checks of the peer-memory collectives and the native DDP / sharded engines.

Launch: ``torchrun --nproc-per-node N --master-addr 127.0.0.1 -m accelerate_hpc_test_b200.test_utils.scripts.test_comm_gpu``
Every check compares against torch.distributed (NCCL) or a single-process replay of the same math.
"""

import json
import os
import sys

import torch
import torch.distributed as dist


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


RESULTS = []  # rows of profiles/fused_paths_n{world}.json
NVLINK_GBS = 900.0      # NVLink 5 per direction and GPU (nominal)
try:
    _mp = json.load(open(os.path.join(os.path.dirname(__file__), "..", "..", "..", "MEASURED_PEAKS.json")))
    PEAK_TFLOPS, HBM_GBS = float(_mp["bf16_tflops"]), float(_mp["hbm_gbs"])
except Exception:  # noqa: BLE001
    PEAK_TFLOPS, HBM_GBS = 2250.0, 7700.0


def record(row, name, ours_ms, baseline, baseline_ms, flops=0.0, link_bytes=0.0, hbm_bytes=0.0):
    """Roofline of a fused path = the slower of its math at the measured bf16 peak and its bytes over the link / HBM."""
    roof_ms = max(flops / (PEAK_TFLOPS * 1e9), link_bytes / (NVLINK_GBS * 1e6), hbm_bytes / (HBM_GBS * 1e6))
    RESULTS.append(dict(row=row, op=name, ours_ms=round(ours_ms, 4), baseline=baseline, baseline_ms=round(baseline_ms, 4),
                        speedup_vs_baseline=round(baseline_ms / ours_ms, 3), roofline_ms=round(roof_ms, 4),
                        roofline_fraction=round(roof_ms / ours_ms, 3), clocks=getattr(ours_ms, "clocks", None),
                        baseline_clocks=getattr(baseline_ms, "clocks", None)))


class Ms(float):
    """Milliseconds that remember the clocks / throttle reasons sampled while they were measured (this rank's GPU)."""

    def __new__(cls, value, clocks=None):
        self = super().__new__(cls, value)
        self.clocks = clocks
        return self


_PROBE = None


def timed(fn, iters=10, warm=3):
    """ms per call: CUDA events on the launching stream, max over ranks; NVML clock samples while the calls are in flight."""
    global _PROBE
    if _PROBE is None:
        from accelerate_hpc_test_b200.utils.diagnostics import ClockProbe

        _PROBE = ClockProbe()
    _PROBE.reset()
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
        _PROBE.sample()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return Ms(t.item(), _PROBE.summary())


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp(min=1e-12)).item()


def check_collectives(ws, dev, rank, world):
    from accelerate_hpc_test_b200.ops._lib import check, ptr
    import ctypes

    torch.manual_seed(100 + rank)
    # ---- all-reduce bf16 / fp32, in place in symmetric memory
    for dtype, n in ((torch.bfloat16, world * 8 * 1000), (torch.float32, world * 4 * 777), (torch.bfloat16, world * 8 * 2_000_000)):
        off, buf = ws.alloc_tensor((n,), dtype)
        src = torch.randn(n, device=dev).to(dtype)
        ref = src.float().clone()
        dist.all_reduce(ref)
        ref /= world
        ch_a, ch_d = ws.new_channel(), ws.new_channel()
        for it in range(3):
            buf.copy_(src)
            torch.cuda.synchronize()
            dist.barrier()
            ws.allreduce_(off, n, dtype, 1.0 / world, ch_a, ch_d)
            torch.cuda.synchronize()
            err = rel(buf, ref)
            assert err < (2e-2 if dtype == torch.bfloat16 else 1e-5), ("allreduce", dtype, n, it, err)
    log("allreduce ok")
    # ---- all-gather pull
    n = 8 * 12345
    off, shard = ws.alloc_tensor((n,), torch.bfloat16)
    shard.copy_(torch.randn(n, device=dev).to(torch.bfloat16))
    torch.cuda.synchronize()
    dist.barrier()
    out = torch.empty(world * n, dtype=torch.bfloat16, device=dev)
    ws.allgather_pull(off, n * 2, out)
    ref = torch.empty_like(out)
    dist.all_gather_into_tensor(ref, shard)
    torch.cuda.synchronize()
    assert torch.equal(out, ref), "allgather"
    log("allgather ok")
    # ---- all-to-all pull
    blk = 8 * 999
    off, send = ws.alloc_tensor((world * blk,), torch.bfloat16)
    send.copy_(torch.randn(world * blk, device=dev).to(torch.bfloat16))
    torch.cuda.synchronize()
    dist.barrier()
    out = torch.empty(world * blk, dtype=torch.bfloat16, device=dev)
    ch_a, ch_d = ws.new_channel(), ws.new_channel()
    ws.alltoall_pull(off, blk * 2, out, ch_a, ch_d)
    ref = torch.empty_like(out)
    dist.all_to_all_single(ref, send.clone())
    torch.cuda.synchronize()
    assert torch.equal(out, ref), "alltoall"
    log("alltoall ok")
    # ---- reduce-scatter + AdamW vs NCCL reduce_scatter + reference AdamW
    from accelerate_hpc_test_b200.ops import functional as F

    n = 8 * 4001
    goff, grad = ws.alloc_tensor((world * n,), torch.bfloat16)
    poff, pshard = ws.alloc_tensor((n,), torch.bfloat16)
    grad.copy_(torch.randn(world * n, device=dev).to(torch.bfloat16))
    master = torch.randn(n, device=dev)
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    rs_ref = torch.empty(n, device=dev)
    dist.reduce_scatter_tensor(rs_ref, grad.float())
    rs_ref /= world
    pm, mm, vm = master.clone(), m.clone(), v.clone()
    kw = dict(lr=1e-2, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1)
    F.adamw_flat_(pm, mm, vm, rs_ref, None, step=1, **kw)
    ch_a, ch_d = ws.new_channel(), ws.new_channel()
    e = ws.next_epoch(ch_a)
    torch.cuda.synchronize()
    dist.barrier()
    check(ws.lib.b200_reduce_scatter_adamw(ws.peer_array, rank, world, goff, n, ptr(master), ptr(m), ptr(v), poff, None,
                                           kw["lr"], kw["beta1"], kw["beta2"], kw["eps"], kw["weight_decay"], 1, 1.0 / world, 1, None,
                                           ch_a, ch_d, e, ws.counter_ptr(ch_a), 16, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert rel(master, pm) < 1e-4, ("rs+adamw master", rel(master, pm))
    assert rel(pshard, pm) < 1e-2, ("rs+adamw shard", rel(pshard, pm))
    log("reduce_scatter+adamw ok")
    # ---- bandwidth numbers (CUDA events, max over ranks)
    res = {}
    n = 2**28  # 512 MiB of bf16
    off, buf = ws.alloc_tensor((n,), torch.bfloat16)
    buf.normal_()
    nccl_buf = buf.clone()
    ch_a, ch_d = ws.new_channel(), ws.new_channel()
    # correctness of the TMA all-reduce first (small integers: the bf16 sum is exact, so the comparison is bitwise)
    chs = (ws.new_channel(), ws.new_channel(), ws.new_channel())
    nt = world * 8 * 1031
    toff, tbuf = ws.alloc_tensor((nt,), torch.bfloat16)
    vals = torch.randint(-8, 8, (nt,), device=dev).to(torch.bfloat16)
    tbuf.copy_(vals)
    tref = vals.clone()
    torch.cuda.synchronize()
    dist.barrier()
    ws.allreduce_tma_(toff, nt, 1.0, chs)
    dist.all_reduce(tref)
    torch.cuda.synchronize()
    assert torch.equal(tbuf, tref), "allreduce_tma"
    log("allreduce_tma ok")
    chs2 = (ws.new_channel(), ws.new_channel(), ws.new_channel())
    candidates = [("ours_allreduce", lambda: ws.allreduce_(off, n, torch.bfloat16, 1.0 / world, ch_a, ch_d, ctas=64)),
                  ("ours_allreduce_tma", lambda: ws.allreduce_tma_(off, n, 1.0 / world, chs2)),
                  ("nccl_allreduce", lambda: dist.all_reduce(nccl_buf))]
    if ws.has_multicast:
        ch_n = (ws.new_channel(), ws.new_channel())
        candidates.insert(2, ("ours_allreduce_nvls", lambda: ws.allreduce_nvls_(off, n, 1.0 / world, ch_n[0], ch_n[1])))
    for name, fn in candidates:
        t = timed(fn, iters=5)
        busbw = 2 * (world - 1) / world * n * 2 / (t * 1e-3) / 1e9
        res[name] = (t, busbw)
        log(f"{name}: {t:.3f} ms, bus bw {busbw:.1f} GB/s ({n * 2 / 2**20:.0f} MiB), clocks {t.clocks}")
    record("W7", "all-reduce 512 MiB bf16, scale fused (TMA: reduce own slice + pull the rest)", res["ours_allreduce_tma"][0], "ncclAllReduce",
           res["nccl_allreduce"][0],
           link_bytes=2 * (world - 1) / world * n * 2)
    if "ours_allreduce_nvls" in res:
        record("W7", "all-reduce 512 MiB bf16, scale fused (NVLS: multimem.ld_reduce own slice, multimem.st to all)", res["ours_allreduce_nvls"][0],
               "ncclAllReduce", res["nccl_allreduce"][0], link_bytes=2 * (world - 1) / world * n * 2)
    # all-gather of one Llama-3-8B decoder layer (436 MB gathered) vs NCCL
    layer_elems = 218_112_000 // (world * 8) * (world * 8)
    sh = layer_elems // world
    soff, shard = ws.alloc_tensor((sh,), torch.bfloat16)
    shard.normal_()
    full = torch.empty(layer_elems, dtype=torch.bfloat16, device=dev)
    full2 = torch.empty_like(full)
    torch.cuda.synchronize()
    dist.barrier()
    for ctas in (16, 32, 64, 148):
        for name, fn in ((f"ours_allgather_ctas{ctas}", lambda: ws.allgather_pull(soff, sh * 2, full, ctas=ctas)),):
            t = timed(fn, iters=5)
            log(f"{name}: {t:.3f} ms, ingress {(world - 1) / world * layer_elems * 2 / (t * 1e-3) / 1e9:.1f} GB/s")
    t_nccl_ag = timed(lambda: dist.all_gather_into_tensor(full2, shard), iters=5)
    log(f"nccl_allgather: {t_nccl_ag:.3f} ms, ingress {(world - 1) / world * layer_elems * 2 / (t_nccl_ag * 1e-3) / 1e9:.1f} GB/s")
    t_ag = timed(lambda: ws.allgather_pull(soff, sh * 2, full, ctas=148), iters=5)
    record("W9", "all-gather of one Llama-3-8B layer (436 MB, TMA bulk pull, 148 CTAs)", t_ag, "ncclAllGather", t_nccl_ag,
           link_bytes=(world - 1) / world * layer_elems * 2)
    # ---- W8: reduce-scatter + AdamW on one layer's gradients vs NCCL reduce-scatter + our flat AdamW kernel
    nsh = sh
    g_off, g_all = ws.alloc_tensor((layer_elems,), torch.bfloat16)
    p_off, p_sh = ws.alloc_tensor((nsh,), torch.bfloat16)
    g_all.normal_()
    mst, m1, v1 = torch.randn(nsh, device=dev), torch.zeros(nsh, device=dev), torch.zeros(nsh, device=dev)
    ch_a2, ch_d2 = ws.new_channel(), ws.new_channel()
    hp = dict(lr=1e-4, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1)

    def ours_rs_adamw():
        e_ = ws.next_epoch(ch_a2)
        ws._epochs[ch_d2] = e_
        check(ws.lib.b200_reduce_scatter_adamw(ws.peer_array, rank, world, g_off, nsh, ptr(mst), ptr(m1), ptr(v1), p_off, None, hp["lr"],
                                               hp["beta1"], hp["beta2"], hp["eps"], hp["weight_decay"], 2, 1.0 / world, 1, None, ch_a2, ch_d2, e_,
                                               ws.counter_ptr(ch_a2), 148, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))

    g_red = torch.empty(nsh, dtype=torch.bfloat16, device=dev)
    sh_out = torch.empty(nsh, dtype=torch.bfloat16, device=dev)

    def nccl_rs_adamw():
        dist.reduce_scatter_tensor(g_red, g_all)
        F.adamw_flat_(mst, m1, v1, g_red, sh_out, step=2, grad_scale=1.0 / world, **hp)

    t_o, t_b = timed(ours_rs_adamw, iters=5), timed(nccl_rs_adamw, iters=5)
    log(f"rs+adamw one layer: ours {t_o:.3f} ms, nccl reduce_scatter + adamw kernel {t_b:.3f} ms")
    record("W8", "reduce-scatter fused with sharded AdamW (one layer, 436 MB of gradients)", t_o, "ncclReduceScatter + AdamW kernel", t_b,
           link_bytes=(world - 1) / world * layer_elems * 2, hbm_bytes=layer_elems * 2 + nsh * 26.0)
    assert torch.equal(full, full2)
    return res


def check_ddp(dev, rank, world, tied: bool = False):
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.optim import FusedAdamW
    from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState

    # tied=True: lm_head and embed_tokens share one Parameter, which then receives two native gradient deliveries per step
    cfg = LlamaConfig.tiny(tie_word_embeddings=tied)
    torch.manual_seed(0)
    ref = LlamaForCausalLM(cfg, device=dev)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, device=dev)
    acc = Accelerator(mixed_precision="bf16")
    opt = FusedAdamW(model.parameters(), lr=1e-3, weight_decay=0.0)
    model, opt = acc.prepare(model, opt)
    assert type(model).__name__ == "B200DataParallel", type(model)
    ref_opt = FusedAdamW(ref.parameters(), lr=1e-3, weight_decay=0.0)
    g = torch.Generator().manual_seed(5)
    for step in range(4):
        ids_all = torch.randint(0, cfg.vocab_size, (world * 2, 128), generator=g).to(dev)
        ids = ids_all[rank * 2 : rank * 2 + 2]
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        opt.step()
        # single-process replay: average of the per-rank losses == gradient averaging
        ref_opt.zero_grad()
        tot = 0
        for r in range(world):
            sub = ids_all[r * 2 : r * 2 + 2]
            l = ref(input_ids=sub, labels=sub).loss / world
            l.backward()
        ref_opt.step()
    torch.cuda.synchronize()
    worst = max(rel(p, q) for p, q in zip(model.module.parameters(), ref.parameters()))
    assert worst < 2e-2, ("ddp param mismatch", tied, worst)
    if tied:
        assert model.module.lm_head.weight is model.module.model.embed_tokens.weight
    mode = opt.optimizer._bwd_mode
    if os.environ.get("ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD", "auto") == "auto":
        assert mode == "active", mode  # steps 3 and 4 updated every bucket right behind its all-reduce
    log(f"ddp ok (tied_embeddings={tied}, worst rel err {worst:.2e}, update behind the all-reduce: {mode})")
    model.release()
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


def check_fsdp(dev, rank, world, fused_env):
    os.environ["ACCELERATE_B200_FUSED_OPTIM_IN_BACKWARD"] = fused_env
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.optim import FusedAdamW
    from accelerate_hpc_test_b200.state import AcceleratorState, GradientState
    from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin

    cfg = LlamaConfig.tiny(num_hidden_layers=3)
    torch.manual_seed(0)
    ref = LlamaForCausalLM(cfg, device=dev)
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, device=dev)
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                            transformer_cls_names_to_wrap=["LlamaDecoderLayer"])
    acc = Accelerator(mixed_precision="bf16", fsdp_plugin=plugin)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.01)
    model, opt = acc.prepare(model, opt)
    assert type(model).__name__ == "B200ShardedModule", type(model)
    ref_opt = FusedAdamW(ref.parameters(), lr=1e-3, weight_decay=0.01)
    g = torch.Generator().manual_seed(7)
    losses, ref_losses = [], []
    for step in range(4):
        ids_all = torch.randint(0, cfg.vocab_size, (world * 2, 128), generator=g).to(dev)
        ids = ids_all[rank * 2 : rank * 2 + 2]
        opt.zero_grad()
        out = model(input_ids=ids, labels=ids)
        acc.backward(out.loss)
        if fused_env == "0":
            acc.clip_grad_norm_(model.parameters(), 1e9)
        opt.step()
        losses.append(out.loss.item())
        ref_opt.zero_grad()
        rl = 0.0
        for r in range(world):
            sub = ids_all[r * 2 : r * 2 + 2]
            l = ref(input_ids=sub, labels=sub).loss
            (l / world).backward()
            if r == rank:
                rl = l.item()
        ref_opt.step()
        ref_losses.append(rl)
    torch.cuda.synchronize()
    sd = model.full_state_dict(cpu=False)
    worst = max(rel(sd[n], p) for n, p in ref.named_parameters())
    log(f"fsdp(fused={fused_env}) losses {losses} ref {ref_losses}")
    assert all(abs(a - b) < 3e-2 * abs(b) for a, b in zip(losses, ref_losses)), (losses, ref_losses)
    assert worst < 3e-2, ("fsdp param mismatch", worst)
    log(f"fsdp ok (fused={fused_env}, worst rel err {worst:.2e})")
    model.release()
    AcceleratorState._reset_state(True)
    GradientState._reset_state()


def check_tp(dev, rank, world):
    """Fused all-gather->GEMM / GEMM->reduce-scatter against NCCL + a plain matmul, then a TP Llama against one GPU."""
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.parallel.tensor_parallel import TPComm, apply_tensor_parallel

    torch.manual_seed(11 + rank)
    H, N, t = 512, 1024, 256
    comm = TPComm(dist.group.WORLD, dev, H, 4096, fused=True)
    for it in range(3):
        x = torch.randn(t, H, device=dev).to(torch.bfloat16)
        w = (torch.randn(N, H, device=dev) * 0.05).to(torch.bfloat16)
        dist.broadcast(w, 0)
        y, full = comm.ag_gemm(x, w)
        ref_full = torch.empty(t * world, H, device=dev, dtype=torch.bfloat16)
        dist.all_gather_into_tensor(ref_full, x)
        torch.cuda.synchronize()
        assert torch.equal(full, ref_full), "fused all-gather moved wrong data"
        assert rel(y, ref_full.float() @ w.float().t()) < 1e-2, ("ag_gemm", rel(y, ref_full.float() @ w.float().t()))
        # GEMM -> reduce-scatter: x2 [T, H] differs per rank (partial products)
        x2 = torch.randn(t * world, H, device=dev).to(torch.bfloat16)
        out = comm.gemm_rs(x2, w)
        part = (x2.float() @ w.float().t())
        ref = torch.empty(t, N, device=dev)
        dist.reduce_scatter_tensor(ref, part)
        torch.cuda.synchronize()
        assert rel(out, ref) < 2e-2, ("gemm_rs", rel(out, ref))
    log("fused ag->gemm and gemm->rs ok")
    # timing at Llama-3-8B shapes: x [T/tp, 4096] -> qkv/gate_up column parallel
    H, T = 4096, 8192
    big = TPComm(dist.group.WORLD, dev, H, T, fused=True)
    x = torch.randn(T // world, H, device=dev).to(torch.bfloat16)
    w = (torch.randn(28672 // world, H, device=dev) * 0.02).to(torch.bfloat16)
    xf = torch.randn(T, H, device=dev).to(torch.bfloat16)
    wr = (torch.randn(H, 14336 // world, device=dev) * 0.02).to(torch.bfloat16)
    xr = torch.randn(T, 14336 // world, device=dev).to(torch.bfloat16)
    full = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
    rs_out = torch.empty(T // world, H, device=dev, dtype=torch.bfloat16)

    def nccl_ag_gemm():
        dist.all_gather_into_tensor(full, x)
        return full @ w.t()

    def nccl_gemm_rs():
        yy = xr @ wr.t()
        dist.reduce_scatter_tensor(rs_out, yy)
        return rs_out

    from accelerate_hpc_test_b200.ops import functional as F
    tms = {}
    for name, fn in (("fused_ag_gemm", lambda: big.ag_gemm(x, w)), ("nccl_ag_then_cublas", nccl_ag_gemm), ("gemm_only", lambda: F.gemm(xf, w)),
                     ("fused_gemm_rs", lambda: big.gemm_rs(xr, wr)), ("cublas_then_nccl_rs", nccl_gemm_rs), ("gemm_only_rs_shape", lambda: F.gemm(xr, wr))):
        tms[name] = timed(fn, iters=10)
        log(f"tp {name}: {tms[name]:.3f} ms")
    record("W10", f"all-gather -> GEMM (x [{T // world}x{H}] per rank, W [{28672 // world}x{H}])", tms["fused_ag_gemm"],
           "ncclAllGather + cuBLASLt", tms["nccl_ag_then_cublas"], flops=2.0 * T * (28672 // world) * H,
           link_bytes=(world - 1) / world * T * H * 2)
    record("W11", f"GEMM -> reduce-scatter (x [{T}x{14336 // world}], W [{H}x{14336 // world}])", tms["fused_gemm_rs"],
           "cuBLASLt + ncclReduceScatter", tms["cublas_then_nccl_rs"], flops=2.0 * T * H * (14336 // world),
           link_bytes=(world - 1) / world * T * H * 2)
    # whole model: TP Llama vs single-GPU replica
    cfg = LlamaConfig.tiny(num_attention_heads=max(4, 2 * world), num_key_value_heads=max(2, world), hidden_size=512, intermediate_size=1024)
    torch.manual_seed(0)
    ref = LlamaForCausalLM(cfg, device=dev)
    torch.manual_seed(0)
    model = apply_tensor_parallel(LlamaForCausalLM(cfg, device=dev), dist.group.WORLD)
    g = torch.Generator().manual_seed(5)
    ids = torch.randint(0, cfg.vocab_size, (2, 256), generator=g).to(dev)
    out = model(ids, labels=ids)
    out.loss.backward()
    r = ref(ids, labels=ids)
    r.loss.backward()
    torch.cuda.synchronize()
    assert abs(out.loss.item() - r.loss.item()) < 2e-2 * abs(r.loss.item()), (out.loss.item(), r.loss.item())
    gw = model.model.layers[1].self_attn.o_proj.weight.grad
    gr = ref.model.layers[1].self_attn.o_proj.weight.grad
    n = gr.shape[1] // world
    assert rel(gw, gr[:, rank * n : (rank + 1) * n]) < 8e-2, rel(gw, gr[:, rank * n : (rank + 1) * n])
    assert rel(model.lm_head.weight.grad, ref.lm_head.weight.grad) < 8e-2
    log(f"tp llama ok (loss {out.loss.item():.4f} vs {r.loss.item():.4f})")


def check_seq(dev, rank, world):
    """Ulysses exchange kernel against all_to_all_single + permutes; SP / CP Llama against one GPU."""
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.parallel.context_parallel import apply_context_parallel, shard_sequence
    from accelerate_hpc_test_b200.parallel.ulysses import SPComm, apply_ulysses

    comm = SPComm(dist.group.WORLD, dev)
    torch.manual_seed(21 + rank)
    B, S_loc, Hq, Hkv, D = 2, 192, 2 * world, world, 128
    widths = (Hq * D, Hkv * D, Hkv * D)
    x = torch.randn(B * S_loc, sum(widths), device=dev).to(torch.bfloat16)
    for it in range(3):
        y = comm.exchange(x, 0, B, S_loc, widths)
        y_ref = comm._exchange_reference(x, 0, B, S_loc, widths)
        assert torch.equal(y, y_ref), "seq->head exchange mismatch"
        z = comm.exchange(y, 1, B, S_loc, widths)
        assert torch.equal(z, x), "head->seq exchange is not the inverse"
        assert torch.equal(comm._exchange_reference(y, 1, B, S_loc, widths), x)
    # timing at the Llama-3-8B shape (per layer: qkv forward exchange)
    Bq, Sl, hq, hkv = 1, 8192 // world, 32, 8
    wd = (hq * D, hkv * D, hkv * D)
    xb = torch.randn(Bq * Sl, sum(wd), device=dev).to(torch.bfloat16)
    for name, fn in (("exchange_kernel", lambda: comm.exchange(xb, 0, Bq, Sl, wd)),
                     ("nccl_all_to_all+permutes", lambda: comm._exchange_reference(xb, 0, Bq, Sl, wd))):
        tt = timed(fn, iters=10)
        log(f"ulysses {name}: {tt:.3f} ms ({xb.numel() * 2 * (world - 1) / world / tt / 1e6:.0f} GB/s ingress), clocks {tt.clocks}")
    # the same bytes in the other direction (attention output -> sequence shards), where rows are cut into `world` pieces
    yb = comm.exchange(xb, 0, Bq, Sl, wd)
    for name, fn in (("exchange_kernel head->seq", lambda: comm.exchange(yb, 1, Bq, Sl, wd)),
                     ("nccl_all_to_all+permutes head->seq", lambda: comm._exchange_reference(yb, 1, Bq, Sl, wd))):
        tt = timed(fn, iters=10)
        log(f"ulysses {name}: {tt:.3f} ms ({xb.numel() * 2 * (world - 1) / world / tt / 1e6:.0f} GB/s ingress)")
    log("ulysses exchange ok")
    # W12 / W13: QKV projection fused with the sequence->head all-to-all (also the dgrad of the output projection)
    from accelerate_hpc_test_b200.ops import functional as F

    Hm = 4096
    xa = torch.randn(Bq * Sl, Hm, device=dev).to(torch.bfloat16)
    wqkv = (torch.randn(sum(wd), Hm, device=dev) * 0.02).to(torch.bfloat16)
    fused = comm.gemm_scatter(xa, wqkv, Bq, Sl, wd)
    unfused = comm._exchange_reference(F.gemm(xa, wqkv), 0, Bq, Sl, wd)
    assert torch.equal(fused, unfused), "fused QKV projection + all-to-all differs from GEMM + all_to_all_single"
    t_f = timed(lambda: comm.gemm_scatter(xa, wqkv, Bq, Sl, wd))
    t_u = timed(lambda: comm._exchange_reference(xa @ wqkv.t(), 0, Bq, Sl, wd))
    log(f"ulysses fused qkv-projection + all-to-all: {t_f:.3f} ms vs cuBLAS + NCCL all_to_all + permutes {t_u:.3f} ms")
    record("W12", f"Ulysses QKV projection + all-to-all ({Bq * Sl} tokens per rank, W [{sum(wd)}x{Hm}])", t_f,
           "cuBLASLt + all_to_all_single + permutes", t_u, flops=2.0 * Bq * Sl * sum(wd) * Hm,
           link_bytes=(world - 1) / world * Bq * Sl * sum(wd) * 2)
    wo = (torch.randn(Hm, hq * D, device=dev) * 0.02).to(torch.bfloat16)
    dyo = torch.randn(Bq * Sl, Hm, device=dev).to(torch.bfloat16)
    t_f = timed(lambda: comm.gemm_scatter(dyo, wo, Bq, Sl, (hq * D,), b_mn_major=True))
    t_u = timed(lambda: comm._exchange_reference(dyo @ wo, 0, Bq, Sl, (hq * D,)))
    log(f"ulysses fused out-projection dgrad + all-to-all: {t_f:.3f} ms vs cuBLAS + NCCL {t_u:.3f} ms")
    record("W13", "Ulysses output-projection dgrad + all-to-all", t_f, "cuBLASLt + all_to_all_single + permutes", t_u,
           flops=2.0 * Bq * Sl * hq * D * Hm, link_bytes=(world - 1) / world * Bq * Sl * hq * D * 2)

    cfg = LlamaConfig.tiny(num_attention_heads=2 * world, num_key_value_heads=world, hidden_size=512, intermediate_size=1024)
    torch.manual_seed(0)
    ref = LlamaForCausalLM(cfg, device=dev)
    g = torch.Generator().manual_seed(5)
    S = 512
    ids = torch.randint(0, cfg.vocab_size, (2, S), generator=g).to(dev)
    r = ref(ids, labels=ids)
    r.loss.backward()

    def compare(model, tag, out):
        torch.cuda.synchronize()
        assert abs(out.loss.item() - r.loss.item()) < 2e-2 * abs(r.loss.item()), (tag, out.loss.item(), r.loss.item())
        for name in ("model.layers.1.self_attn.qkv_proj.weight", "model.layers.0.mlp.down_proj.weight", "lm_head.weight"):
            gw = dict(model.named_parameters())[name].grad.float().clone()
            dist.all_reduce(gw)
            gw /= world
            gr = dict(ref.named_parameters())[name].grad
            assert rel(gw, gr) < 8e-2, (tag, name, rel(gw, gr))
        log(f"{tag} llama ok (loss {out.loss.item():.4f} vs {r.loss.item():.4f})")

    torch.manual_seed(0)
    sp_model = apply_ulysses(LlamaForCausalLM(cfg, device=dev), dist.group.WORLD)
    c = S // world
    out = sp_model(ids[:, rank * c : (rank + 1) * c].contiguous(), labels=ids[:, rank * c : (rank + 1) * c].contiguous())
    out.loss.backward()
    compare(sp_model, "ulysses", out)

    torch.manual_seed(0)
    cp_model = apply_context_parallel(LlamaForCausalLM(cfg, device=dev), dist.group.WORLD)
    loc = shard_sequence(ids, 1, rank, world).contiguous()
    out = cp_model(loc, labels=loc)
    out.loss.backward()
    compare(cp_model, "context-parallel", out)


def main():
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    from accelerate_hpc_test_b200.parallel.symm import SymmetricWorkspace

    what = sys.argv[1:] or ["coll", "ddp", "fsdp", "tp", "seq"]
    if "coll" in what:
        ws = SymmetricWorkspace(dist.group.WORLD, 3 * 2**30, dev)
        check_collectives(ws, dev, rank, world)
        ws.close()
    if "ddp" in what:
        check_ddp(dev, rank, world)
        check_ddp(dev, rank, world, tied=True)
    if "fsdp" in what:
        check_fsdp(dev, rank, world, "0")
        check_fsdp(dev, rank, world, "1")
    if "tp" in what:
        check_tp(dev, rank, world)
    if "seq" in what:
        check_seq(dev, rank, world)
    dist.barrier()
    if rank == 0 and RESULTS:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/fused_paths_n{world}.json", "w") as f:
            json.dump({"world": world, "peaks": {"bf16_tflops": PEAK_TFLOPS, "nvlink_gbs_per_direction": NVLINK_GBS, "hbm_gbs": HBM_GBS},
                       "rows": RESULTS}, f, indent=1)
        for r_ in RESULTS:
            log(json.dumps(r_))
    log("ALL MULTI-GPU CHECKS PASSED")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
