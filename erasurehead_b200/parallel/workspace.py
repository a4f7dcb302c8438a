"""Peer-mapped workspace: the NVLink replacement for the reference's MPI message buffers.

Every GPU allocates one block (layout in ``include/erasurehead.h``) holding the double-buffered
parameters, the parameter flag, and -- used on the root only -- the per-(worker, message)
gradient slots with their arrival flags.  CUDA IPC handles are exchanged once through
``torch.distributed``; afterwards

* every worker GPU stores its encoded gradient straight into the *root's* slot and releases the
  arrival flag (replaces ``comm.Isend([g, MPI.DOUBLE], dest=0, tag=i)``, naive.py:150);
* the root stores new parameters into *every* GPU's buffer and releases the parameter flag
  (replaces the ``Isend(beta)`` loop, naive.py:97-98).

The mapping covers the GPUs of rank 0's node.  A rank on another node owns an identical block that
only its own kernels touch; ``parallel/bridge.py`` carries its messages and parameters over NCCL.
"""
from __future__ import annotations

import ctypes as C
from typing import List

from ..ops._lib import WsLayout, check, lib, stream_ptr
from .context import DistContext


class Workspace:
    def __init__(self, ctx: DistContext, n_workers: int, n_msgs: int, d: int):
        self.ctx = ctx
        L = lib()
        self.lay = WsLayout()
        check(L.eh_ws_layout_make(n_workers, n_msgs, d, C.byref(self.lay)), "eh_ws_layout_make")
        self._local = C.c_void_p()
        handle = C.create_string_buffer(64)
        check(L.eh_ws_create(C.byref(self.lay), C.byref(self._local), handle), "eh_ws_create")
        self.local = self._local.value
        self._opened: List[int] = []
        handles = ctx.all_gather_object(bytes(handle.raw))
        # root maps every peer of its node (parameter push); every rank of that node maps the root (gradient
        # ship).  Ranks on other nodes keep to their own block: their kernels "ship" into the local slots and the
        # NCCL bridge (parallel/bridge.py) moves messages and parameters between the blocks.
        self.peers: List[int] = [0] * ctx.world
        self.peers[ctx.rank] = self.local
        if ctx.world > 1 and not ctx.is_remote:
            want = ctx.domain_ranks if ctx.is_root else [0]
            for r in want:
                if r == ctx.rank:
                    continue
                p = C.c_void_p()
                check(L.eh_ws_open(handles[r], C.byref(p)), f"eh_ws_open(rank {r})")
                self.peers[r] = p.value
                self._opened.append(p.value)
        self.root = self.local if ctx.is_remote else self.peers[0]
        self.clock_offset_ns: List[int] = [0] * ctx.world      # peer clock - root clock (root only)
        self.clock_rtt_ns: List[int] = [0] * ctx.world

    def sync_clocks(self, iters: int = 64) -> None:
        """Measure every peer GPU's ``%globaltimer`` offset against the root's (one ping-pong kernel
        pair per peer through the control words).  Worker kernels stamp message completion with
        their own GPU's clock; the decoder subtracts these offsets to put the stamps on its clock."""
        import torch
        ctx, L = self.ctx, lib()
        if ctx.world == 1:
            return
        self.reset()
        torch.cuda.synchronize()
        ctx.barrier()
        if ctx.is_root:
            out = torch.zeros(2 * ctx.world, dtype=torch.int64, device=ctx.device)
            for r in ctx.domain_ranks[1:]:
                check(L.eh_clock_sync_root(C.byref(self.lay), self.local, self.peers[r], r << 16, iters,
                                           out.data_ptr() + 8 * r, out.data_ptr() + 8 * (ctx.world + r), stream_ptr()),
                      "eh_clock_sync_root")
            torch.cuda.synchronize()
            v = out.cpu().tolist()
            self.clock_offset_ns = [int(x) for x in v[:ctx.world]]
            self.clock_rtt_ns = [int(x) for x in v[ctx.world:]]
            if any(x < 0 for x in self.clock_rtt_ns[1:]):
                raise RuntimeError(f"clock alignment with a peer GPU timed out (round trips {self.clock_rtt_ns})")
        elif not ctx.is_remote:
            check(L.eh_clock_sync_peer(C.byref(self.lay), self.local, self.root, ctx.rank << 16, iters, stream_ptr()),
                  "eh_clock_sync_peer")
            torch.cuda.synchronize()
        ctx.barrier()
        self.reset()
        torch.cuda.synchronize()

    def reset(self) -> None:
        check(lib().eh_ws_reset(C.byref(self.lay), self.local, stream_ptr()), "eh_ws_reset")

    def error(self) -> int:
        e = C.c_int(0)
        check(lib().eh_ws_error(C.byref(self.lay), self.local, C.byref(e), stream_ptr()), "eh_ws_error")
        return e.value

    def close(self) -> None:
        L = lib()
        for p in self._opened:
            L.eh_ws_close(p)
        self._opened = []
        if self.local:
            L.eh_ws_destroy(self.local)
            self.local = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
