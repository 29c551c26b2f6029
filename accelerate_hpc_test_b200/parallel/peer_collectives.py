"""All-gather and reduce-scatter of activation-sized tensors on the symmetric workspace, for the model-parallel engines
(context parallel K/V exchange, tensor-parallel token gathers) — the kernels of csrc/comm.cu instead of NCCL calls.

* ``all_gather(x)``: my rows are copied into a staging slot of the workspace, one device barrier publishes them, the TMA pull
  kernel assembles ``[world, *x.shape]`` from the peers' slots.  Two staging slots alternate, so the single barrier of call
  *k+1* is also the proof that every rank finished reading the slot written in call *k-1*.
* ``reduce_scatter(x)``: ``x`` is ``[world, rows, w]`` (partial sums for every rank's rows); it is staged the same way and
  rank *r* receives the sum over ranks of slice *r* — added inside the NVSwitch (``multimem.ld_reduce``) when the workspace
  has a multicast mapping, else pulled and summed through the TMA engine.

The workspace grows on demand (collectively: every rank sees the same sizes).  On CPU / gloo the same calls use c10d.
"""

from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops._lib import check, cur_stream, ptr


class PeerCollectives:
    def __init__(self, group, device):
        self.group = group
        self.device = torch.device(device)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.ws = None
        self._cap = 0
        self._calls = 0

    @property
    def native(self) -> bool:
        return self.device.type == "cuda" and self.world > 1

    def _ensure(self, nbytes: int):
        need = (nbytes + 255) // 256 * 256
        if self.ws is not None and need <= self._cap:
            return
        from .symm import SymmetricWorkspace

        if self.ws is not None:
            torch.cuda.synchronize(self.device)
            self.ws.close()
        self._cap = max(need, 32 * 2**20)
        self.ws = SymmetricWorkspace(self.group, 2 * self._cap + 2**20, self.device)
        self._slots = [self.ws.alloc(self._cap), self.ws.alloc(self._cap)]
        self._ch_bar = self.ws.new_channel()
        self._ch_a, self._ch_d = self.ws.new_channel(), self.ws.new_channel()

    def _stage(self, x: torch.Tensor) -> int:
        self._ensure(x.numel() * x.element_size())
        off = self._slots[self._calls & 1]
        self._calls += 1
        self.ws.tensor(off, x.shape, x.dtype).copy_(x)
        return off

    # ------------------------------------------------------------------ all-gather
    def all_gather(self, x: torch.Tensor) -> torch.Tensor:
        """``x [..] -> [world, ..]`` (same dtype)."""
        x = x.contiguous()
        out = torch.empty((self.world,) + tuple(x.shape), dtype=x.dtype, device=x.device)
        if not self.native or (x.numel() * x.element_size()) % 16:
            dist.all_gather_into_tensor(out.view(-1, *x.shape[1:]) if x.dim() > 0 else out, x, group=self.group)
            return out
        off = self._stage(x)
        self.ws.barrier(self._ch_bar)  # every rank's slot is written (and everybody is done with the slot of two calls ago)
        self.ws.allgather_pull(off, x.numel() * x.element_size(), out, ctas=64)
        return out

    # ------------------------------------------------------------------ reduce-scatter
    def reduce_scatter(self, x: torch.Tensor) -> torch.Tensor:
        """``x [world, rows, w]`` bf16 -> ``[rows, w]``: sum over ranks of slice ``rank``."""
        x = x.contiguous()
        piece = x[0].numel()
        if not self.native or x.dtype != torch.bfloat16 or piece % 8:
            out = torch.empty(tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
            if x.is_cuda:
                dist.reduce_scatter_tensor(out, x.view(-1, *x.shape[2:]), group=self.group)
            else:  # gloo has no reduce_scatter
                dist.all_reduce(x, group=self.group)
                out.copy_(x[self.rank])
            return out
        off = self._stage(x)
        ws = self.ws
        acc = torch.empty(piece, dtype=torch.float32, device=x.device)
        tab = (ctypes.c_longlong * 3)(0, 0, piece)
        e = ws.next_epoch(self._ch_a)
        ws._epochs[self._ch_d] = e
        common = (self.rank, self.world, self.rank, off, tab, 1, None, None, None, 0, ptr(acc), 0.0, 0.9, 0.999, 1e-8, 0.0, 1, 1.0, 0, None,
                  self._ch_a, self._ch_d, e, ws.counter_ptr(self._ch_a), 64, cur_stream())
        if ws.has_multicast:
            check(ws.lib.b200_reduce_scatter_adamw_nvls(ws.peer_array, ctypes.c_void_p(ws.mc_base), *common), "reduce_scatter_nvls")
        else:
            check(ws.lib.b200_reduce_scatter_adamw_segs(ws.peer_array, *common), "reduce_scatter_segs")
        return acc.view(x.shape[1:]).to(x.dtype)

    def close(self):
        if self.ws is not None:
            self.ws.close()
            self.ws = None
