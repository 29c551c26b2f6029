"""Multi-GPU checks of the peer-memory collectives and the native DDP / sharded engines.

Launch: ``torchrun --nproc-per-node N --master-addr 127.0.0.1 -m accelerate_hpc_test_b200.test_utils.scripts.test_comm_gpu``
Every check compares against torch.distributed (NCCL) or a single-process replay of the same math.
"""

import os
import sys
import time

import torch
import torch.distributed as dist


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


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
    for name, fn in (("ours_allreduce", lambda: ws.allreduce_(off, n, torch.bfloat16, 1.0 / world, ch_a, ch_d, ctas=64)),
                     ("nccl_allreduce", lambda: dist.all_reduce(nccl_buf))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        s, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            fn()
        e2.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e2) / 5], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        busbw = 2 * (world - 1) / world * n * 2 / (t.item() * 1e-3) / 1e9
        res[name] = (t.item(), busbw)
        log(f"{name}: {t.item():.3f} ms, bus bw {busbw:.1f} GB/s ({n * 2 / 2**20:.0f} MiB)")
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
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            s, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(5):
                fn()
            e2.record()
            torch.cuda.synchronize()
            t = torch.tensor([s.elapsed_time(e2) / 5], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            log(f"{name}: {t.item():.3f} ms, ingress {(world - 1) / world * layer_elems * 2 / (t.item() * 1e-3) / 1e9:.1f} GB/s")
    for _ in range(3):
        dist.all_gather_into_tensor(full2, shard)
    torch.cuda.synchronize()
    dist.barrier()
    s, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        dist.all_gather_into_tensor(full2, shard)
    e2.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e2) / 5], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    log(f"nccl_allgather: {t.item():.3f} ms, ingress {(world - 1) / world * layer_elems * 2 / (t.item() * 1e-3) / 1e9:.1f} GB/s")
    assert torch.equal(full, full2)
    return res


def check_ddp(dev, rank, world):
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.optim import FusedAdamW
    from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState

    cfg = LlamaConfig.tiny()
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
    for step in range(3):
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
    assert worst < 2e-2, ("ddp param mismatch", worst)
    log(f"ddp ok (worst rel err {worst:.2e})")
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


def main():
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    from accelerate_hpc_test_b200.parallel.symm import SymmetricWorkspace

    what = sys.argv[1:] or ["coll", "ddp", "fsdp"]
    if "coll" in what:
        ws = SymmetricWorkspace(dist.group.WORLD, 3 * 2**30, dev)
        check_collectives(ws, dev, rank, world)
        ws.close()
    if "ddp" in what:
        check_ddp(dev, rank, world)
    if "fsdp" in what:
        check_fsdp(dev, rank, world, "0")
        check_fsdp(dev, rank, world, "1")
    dist.barrier()
    log("ALL MULTI-GPU CHECKS PASSED")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
