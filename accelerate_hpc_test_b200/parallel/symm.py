"""Symmetric (peer-mapped) device memory: the data plane of the fused collectives.

Every rank owns one buffer of the same size and maps every peer's buffer, so a kernel on GPU *r* can ``ld`` / ``st`` any
byte of any peer over NVLink.  Two ways to get there, tried in this order:

* **VMM + NVLS multicast** (``torch.distributed._symmetric_memory``: ``cuMemCreate`` allocations exchanged as file
  descriptors and bound to a ``cuMulticast`` object).  Besides the per-peer addresses this yields ``mc_base``, the
  *multicast alias* of the buffer: ``multimem.ld_reduce`` through it returns the sum over all ranks computed inside the
  NVSwitch, ``multimem.st`` stores to all ranks at once (csrc/comm.cu, the ``*_nvls`` kernels).
* **CUDA IPC** (``cudaMalloc`` + ``cudaIpcGetMemHandle``): peer addresses only, ``mc_base == 0``; used when the first
  path is not available (``ACCELERATE_B200_SYMM_BACKEND=ipc`` forces it).

The first bytes of the buffer are the signal pad the kernels order themselves with; the rest is handed out by a bump
allocator that every rank drives identically, so equal allocations land at equal offsets ("symmetric").

There is no reference counterpart: the reference's data plane is NCCL's internal buffers
(SURVEY.md §5.8).  The kernels that use this memory are in ``csrc/comm.cu``.
"""

from __future__ import annotations

import ctypes
import itertools
import os

import torch

from ..ops import _lib
from ..ops._lib import check


def peer_domain(group=None) -> tuple:
    """(ok, reason): can the ranks of ``group`` map each other's memory?  CUDA-IPC handles do not cross hosts and the signal
    pads are sized for 8 ranks (one NVLink / NVSwitch domain), so larger or multi-node groups must use the NCCL engines."""
    import socket

    group = group if group is not None else torch.distributed.group.WORLD
    world = torch.distributed.get_world_size(group)
    if world > 8:
        return False, f"{world} ranks exceed one 8-GPU NVLink domain"
    hosts = [None] * world
    torch.distributed.all_gather_object(hosts, socket.gethostname(), group=group)
    if len(set(hosts)) > 1:
        return False, f"ranks span {len(set(hosts))} hosts ({sorted(set(hosts))[:3]}...)"
    return True, ""


class _RawCudaBuffer:
    """Object exposing ``__cuda_array_interface__`` so torch can wrap raw device memory without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


class SymmetricWorkspace:
    """One peer-mapped buffer per rank plus channel / epoch bookkeeping for the comm kernels."""

    def __init__(self, group, nbytes: int, device: torch.device):
        self.group = group
        self.world = torch.distributed.get_world_size(group)
        self.rank = torch.distributed.get_rank(group)
        self.device = device
        lib = _lib.require_gpu_lib()
        self.lib = lib
        max_ranks, n_channels, pad = ctypes.c_int(), ctypes.c_int(), ctypes.c_longlong()
        check(lib.b200_comm_limits(ctypes.byref(max_ranks), ctypes.byref(n_channels), ctypes.byref(pad)))
        if self.world > max_ranks.value:
            raise RuntimeError(f"symmetric workspace supports up to {max_ranks.value} ranks per NVLink domain, got {self.world}")
        self.num_channels = n_channels.value
        self.pad_bytes = (pad.value + 4095) // 4096 * 4096
        self.nbytes = int(nbytes) + self.pad_bytes
        self.mc_base = 0
        self._opened = []
        self._symm_tensor = None
        self.backend = "ipc"
        # a group of one has no peers to map and no switch to reduce in: plain device memory through the IPC path
        if os.environ.get("ACCELERATE_B200_SYMM_BACKEND", "auto") != "ipc" and self.world > 1 and self._try_vmm_multicast(group, device):
            self.backend = "vmm"
        else:
            self._alloc_ipc(group, device, lib)
        self.peer_array = (ctypes.c_void_p * self.world)(*self.peer_bases)
        self._cursor = self.pad_bytes
        self._channel_iter = itertools.count(1)
        self._epochs: dict[int, int] = {}
        self._counters = torch.zeros(self.num_channels, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        torch.distributed.barrier(group=group)
        self.closed = False

    def _try_vmm_multicast(self, group, device) -> bool:
        """Allocate through torch's symmetric-memory allocator (VMM handles + NVLS multicast binding).  Every rank must reach
        the same verdict, so success is agreed on with an all-reduce before anything is kept."""
        ok, t, hdl = 1, None, None
        try:
            import torch.distributed._symmetric_memory as symm_mem

            if torch.distributed.get_backend(group) != "nccl":
                raise RuntimeError("symmetric memory needs an NCCL group")
            nbytes = (self.nbytes + (1 << 21) - 1) >> 21 << 21
            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
            hdl = symm_mem.rendezvous(t, group)
            if len(hdl.buffer_ptrs) != self.world:
                raise RuntimeError("unexpected peer count")
        except Exception:  # noqa: BLE001 - any failure means: use CUDA IPC instead
            ok = 0
        flag = torch.tensor([ok], device=device, dtype=torch.int32)
        torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            del hdl, t
            return False
        t.zero_()
        self._symm_tensor, self._symm_handle = t, hdl
        self._whole = t[: self.nbytes]
        self.peer_bases = [int(x) for x in hdl.buffer_ptrs]
        self.base = self.peer_bases[self.rank]
        try:
            self.mc_base = int(hdl.multicast_ptr or 0)
        except Exception:  # noqa: BLE001
            self.mc_base = 0
        return True

    def _alloc_ipc(self, group, device, lib):
        base = ctypes.c_void_p()
        check(lib.b200_symm_alloc(ctypes.byref(base), self.nbytes), "symm_alloc")
        self.base = base.value
        handle = (ctypes.c_ubyte * 64)()
        check(lib.b200_ipc_get_handle(ctypes.c_void_p(self.base), handle), "ipc_get_handle")
        handles = [None] * self.world
        torch.distributed.all_gather_object(handles, bytes(handle), group=group)
        self.peer_bases = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.peer_bases.append(self.base)
                continue
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            p = ctypes.c_void_p()
            check(lib.b200_ipc_open_handle(buf, ctypes.byref(p)), "ipc_open_handle")
            self.peer_bases.append(p.value)
            self._opened.append(p.value)
        self._whole = torch.as_tensor(_RawCudaBuffer(self.base, self.nbytes), device=device)

    @property
    def has_multicast(self) -> bool:
        return self.mc_base != 0 and os.environ.get("ACCELERATE_B200_NVLS", "1") != "0"

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def create_default(cls, state) -> "SymmetricWorkspace":
        gb = float(os.environ.get("ACCELERATE_B200_SYMM_GB", "4"))
        return cls(torch.distributed.group.WORLD, int(gb * 2**30), state.device)

    # ------------------------------------------------------------------ allocation
    def alloc(self, nbytes: int, align: int = 256) -> int:
        """Reserve ``nbytes`` and return the byte offset (identical on every rank when called in lock-step)."""
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError(
                f"symmetric workspace exhausted: need {nbytes} bytes at offset {off}, capacity {self.nbytes}. "
                "Raise ACCELERATE_B200_SYMM_GB."
            )
        self._cursor = off + nbytes
        return off

    def tensor(self, offset: int, shape, dtype: torch.dtype) -> torch.Tensor:
        """A tensor view of my own buffer at ``offset``."""
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return self._whole[offset : offset + nbytes].view(dtype).view(*shape)

    def alloc_tensor(self, shape, dtype: torch.dtype):
        n = 1
        for s in shape:
            n *= int(s)
        off = self.alloc(n * torch.empty((), dtype=dtype).element_size())
        return off, self.tensor(off, shape, dtype)

    def new_channel(self) -> int:
        c = next(self._channel_iter)
        if c >= self.num_channels:
            raise RuntimeError("out of signal channels")
        self._epochs[c] = 0
        return c

    def next_epoch(self, channel: int) -> int:
        self._epochs[channel] += 1
        return self._epochs[channel]

    def epoch(self, channel: int) -> int:
        return self._epochs[channel]

    def counter_ptr(self, channel: int) -> ctypes.c_void_p:
        return ctypes.c_void_p(self._counters.data_ptr() + 4 * channel)

    # ------------------------------------------------------------------ collectives (thin launch wrappers)
    def barrier(self, channel: int, stream=None):
        e = self.next_epoch(channel)
        check(self.lib.b200_barrier_all(self.peer_array, self.rank, self.world, channel, e, _stream(stream)), "barrier_all")

    def allgather_pull(self, shard_off: int, shard_bytes: int, out: torch.Tensor, wait_channel: int = -1, wait_epoch: int = 0,
                       ctas: int = 32, stream=None):
        check(self.lib.b200_allgather_pull(self.peer_array, self.rank, self.world, shard_off, shard_bytes, _lib.ptr(out),
                                           wait_channel, wait_epoch, ctas, _stream(stream)), "allgather_pull")

    def allreduce_(self, off: int, numel: int, dtype: torch.dtype, scale: float, ch_arrive: int, ch_done: int, ctas: int = 32,
                   stream=None):
        e = self.next_epoch(ch_arrive)
        self._epochs[ch_done] = e
        check(self.lib.b200_allreduce(self.peer_array, self.rank, self.world, off, numel, int(dtype == torch.bfloat16), float(scale),
                                      ch_arrive, ch_done, e, self.counter_ptr(ch_arrive), ctas, _stream(stream)), "allreduce")

    def allreduce_nvls_(self, off: int, numel: int, scale: float, ch_arrive: int, ch_done: int, ctas: int = 148, stream=None):
        """In-place bf16 all-reduce with the sum done inside the NVSwitch (needs ``has_multicast``)."""
        e = self.next_epoch(ch_arrive)
        self._epochs[ch_done] = e
        check(self.lib.b200_allreduce_nvls(self.peer_array, ctypes.c_void_p(self.mc_base), self.rank, self.world, off, numel, float(scale),
                                           ch_arrive, ch_done, e, self.counter_ptr(ch_arrive), ctas, _stream(stream)), "allreduce_nvls")

    def allreduce_tma_(self, off: int, numel: int, scale: float, channels, ctas: int = 148, stream=None):
        """In-place bf16 all-reduce on the TMA engine (reduce my slice, then pull the others).  ``channels`` = three channel
        ids reserved for this buffer (arrive / between the phases / done)."""
        ch_a, ch_m, ch_d = channels
        e = self.next_epoch(ch_a)
        self._epochs[ch_m] = e
        self._epochs[ch_d] = e
        check(self.lib.b200_allreduce_tma(self.peer_array, self.rank, self.world, off, numel, float(scale), ch_a, ch_m, ch_d, e,
                                          self.counter_ptr(ch_a), ctas, _stream(stream)), "allreduce_tma")

    def alltoall_pull(self, send_off: int, block_bytes: int, out: torch.Tensor, ch_arrive: int, ch_done: int, ctas: int = 32,
                      stream=None):
        e = self.next_epoch(ch_arrive)
        self._epochs[ch_done] = e
        check(self.lib.b200_alltoall_pull(self.peer_array, self.rank, self.world, send_off, block_bytes, _lib.ptr(out), ch_arrive,
                                          ch_done, e, self.counter_ptr(ch_arrive), ctas, _stream(stream)), "alltoall_pull")

    def close(self):
        if self.closed:
            return
        self.closed = True
        try:
            torch.cuda.synchronize(self.device)
            self._whole = None
            if self.backend == "vmm":
                self._symm_handle = None
                self._symm_tensor = None
            else:
                for p in self._opened:
                    self.lib.b200_ipc_close_handle(ctypes.c_void_p(p))
                self.lib.b200_symm_free(ctypes.c_void_p(self.base))
        except Exception:  # pragma: no cover
            pass


def _stream(stream) -> ctypes.c_void_p:
    if stream is None:
        return _lib.cur_stream()
    return ctypes.c_void_p(stream.cuda_stream)
