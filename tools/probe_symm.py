"""Probe: does torch.distributed._symmetric_memory give peer + multicast pointers on this box?"""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
try:
    t = symm_mem.empty(64 * 2**20, dtype=torch.uint8, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    print(rank, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast_ptr", hex(hdl.multicast_ptr), "signal_pad", [hex(p) for p in hdl.signal_pad_ptrs][:2], "size", hdl.buffer_size, flush=True)
    # sanity: all_reduce via multimem op if available
    x = t.view(torch.bfloat16)[:1024 * 1024]
    x.fill_(rank + 1)
    dist.barrier(); torch.cuda.synchronize()
    try:
        torch.ops.symm_mem.multimem_all_reduce_(x, "sum", dist.group.WORLD.group_name)
        torch.cuda.synchronize()
        print(rank, "multimem_all_reduce ->", float(x[0]), "expect", world * (world + 1) / 2, flush=True)
    except Exception as e:
        print(rank, "multimem op failed:", repr(e)[:300], flush=True)
except Exception as e:
    print(rank, "symm_mem failed:", repr(e)[:500], flush=True)
dist.barrier()
dist.destroy_process_group()
