"""Bridge between NVLink domains: ranks on other nodes than rank 0.

The reference runs one MPI rank per EC2 host (tools/pytorch_ec2.py, run_approx_coding.sh:47-49):
parameters and gradients cross the network every round (naive.py:97-98,150).  Inside one node this
framework replaces those messages with peer stores into a mapped workspace; a GPU on another node
cannot be mapped, so for such a *remote* rank the same two transfers are NCCL point-to-point
operations between the two workspace blocks, enqueued on CUDA streams like every other piece of a
round -- the host still only enqueues, nothing on the device protocol changes:

* a remote rank's gradient kernels ship into its OWN slots; the slot is then sent to rank 0, which
  receives it into the same slot of its block and, behind the receive on the same stream, launches
  ``eh_bridge_arrival``: the arrival record the decoder polls (round stamp + completion time on the
  root's clock, i.e. the network transfer is part of that worker's completion time);
* after the decoder of a round the root sends the new parameters to every remote rank, which
  receives them into its parameter buffer and releases its own parameter flag
  (``eh_bridge_beta_flag``) for the kernels queued behind it.

Every (root, remote) pair has its own process group, hence its own NCCL communicator and stream:
a slow or dead node delays only its own transfers, and the decoder's stop rule skips its messages
exactly as it skips a straggler inside the node.  Receives go to per-remote streams on the root so
that the root's own gradient kernels and the decoder never wait for them in stream order.

On the root the receives and arrival kernels of a round are enqueued before that round's decoder kernel:
the decoder spins on the arrival records, and nothing a spinning kernel waits for should be issued after it
(the first version, issued the other way round, hung in round 0 -- most likely a kernel queued behind the spinning
decoder in a shared hardware queue).

Per pair the operations are issued in one fixed order on both sides (message(s) of round r, then
the parameters of round r+1), which is what NCCL point-to-point requires.  A remote rank waits for
the sends of round r-2 before a kernel overwrites those slots again.  The root does not wait for
its parameter sends: a remote rank that is rounds behind may read a parameter buffer the decoder
is already rewriting, but every message computed from it carries a stamp the decoder has passed
and is never read (the same rule that discards a straggler's late message inside the node).
"""
from __future__ import annotations

import ctypes as C
from collections import deque
from typing import Dict, List, Tuple

from ..schemes import MSG_CODED, MSG_PRIVATE


def messages_of(spec, ctx, rank: int) -> List[Tuple[int, int]]:
    """(worker, message kind) pairs a rank sends per round, in the order its jobs are built
    (engine.py: per local worker, private part first, partial_coded.py:229,234)."""
    out = []
    for w in range(spec.n):
        if ctx.owner(w) != rank:
            continue
        for kind in (MSG_PRIVATE, MSG_CODED):
            if any(sg.msg == kind for sg in spec.segments[w]):
                out.append((w, kind))
    return out


_PAIR_GROUPS: Dict[Tuple[int, int], object] = {}      # (world group, remote rank) -> process group {0, rank}


def pair_group(x: int):
    """The two-rank group of rank 0 and remote rank ``x``.  Creation is collective over the whole job, so every
    rank calls this for every remote rank in the same order (the engines of a job are built in lock-step)."""
    import torch.distributed as dist
    key = (id(dist.group.WORLD), x)                     # a re-initialised process group gets new pair groups
    if key not in _PAIR_GROUPS:
        _PAIR_GROUPS[key] = dist.new_group(ranks=[0, x])
    return _PAIR_GROUPS[key]


class NcclBridge:
    """Owned by an :class:`~erasurehead_b200.engine.Engine` whose process group spans several nodes."""

    def __init__(self, engine):
        import torch
        import torch.distributed as dist
        from ..ops import _lib
        self.torch, self.dist, self._lib = torch, dist, _lib
        self.eng = engine
        ctx, ws = engine.ctx, engine.ws
        self.ctx, self.ws = ctx, ws
        self.remote = ctx.remote_ranks
        # one group per (root, remote) pair; every rank takes part in every creation, in the same order
        self.groups: Dict[int, object] = {x: pair_group(x) for x in self.remote}
        L = _lib.lib()
        lay = ws.lay
        dpad = int(lay.d_pad)
        self._slots: Dict[Tuple[int, int, int], object] = {}
        self._beta = [engine._tensor_from_ptr(int(L.eh_ws_beta_ptr(C.byref(lay), ws.local, par)), dpad) for par in (0, 1)]
        self.msgs: Dict[int, List[Tuple[int, int]]] = {}
        if ctx.is_root:
            _, hi = torch.cuda.Stream.priority_range()
            self.streams = {x: torch.cuda.Stream(device=ctx.device, priority=hi) for x in self.remote}
            for x in self.remote:
                self.msgs[x] = messages_of(engine.spec, ctx, x)
        elif ctx.is_remote:
            self.msgs[ctx.rank] = messages_of(engine.spec, ctx, ctx.rank)
            self._sends = deque()
        for lst in self.msgs.values():
            for (w, m) in lst:
                for par in (0, 1):
                    p = L.eh_ws_slot_ptr(C.byref(lay), ws.local, w, m, par)
                    if not p:
                        raise RuntimeError(f"bridge: slot ({w},{m}) outside the workspace")
                    self._slots[(w, m, par)] = engine._tensor_from_ptr(int(p), dpad)

    @property
    def active(self) -> bool:
        """Whether this rank takes part in transfers (ranks of the root's node other than rank 0 do not)."""
        return self.ctx.is_root or self.ctx.is_remote

    # ------------------------------------------------------------------ parameters of the first round
    def publish(self, round_: int) -> None:
        """After ``eh_decoder_publish`` on the root: the first parameters of a run reach the remote ranks."""
        if not self.active:
            return
        par = round_ & 1
        if self.ctx.is_root:
            for x in self.remote:
                self.dist.isend(self._beta[par], dst=x, group=self.groups[x])
        else:
            self._recv_beta(par, round_ + 1)

    def _recv_beta(self, par: int, stamp: int) -> None:
        L, lay = self._lib.lib(), self.ws.lay
        self.dist.irecv(self._beta[par], src=0, group=self.groups[self.ctx.rank]).wait()     # stream-level wait
        self._lib.check(L.eh_bridge_beta_flag(C.byref(lay), self.ws.local, stamp, self._lib.stream_ptr()),
                        "eh_bridge_beta_flag")

    # ------------------------------------------------------------------ one round
    def before_round(self, round_: int) -> None:
        """Enqueued BEFORE the kernels of ``round_``.

        Root: the receives of this round's remote messages and their arrival records.  They are issued ahead of
        the decoder kernel of the round on purpose: the decoder spins until the arrival records appear, and
        nothing a spinning kernel waits for may be issued behind it (two streams can share a hardware queue, where
        the later kernel would never start).
        Remote rank: the sends of round - 2 used the slots this round's kernels overwrite."""
        if self.ctx.is_root:
            torch, dist = self.torch, self.dist
            L, lay = self._lib.lib(), self.ws.lay
            par, stamp = round_ & 1, round_ + 1
            for x in self.remote:
                st = self.streams[x]
                with torch.cuda.stream(st):
                    for (w, m) in self.msgs[x]:
                        dist.irecv(self._slots[(w, m, par)], src=x, group=self.groups[x]).wait()   # stream-level wait
                        self._lib.check(L.eh_bridge_arrival(C.byref(lay), self.ws.local, w, m, stamp, st.cuda_stream),
                                        "eh_bridge_arrival")
        elif self.ctx.is_remote:
            keep = len(self.msgs[self.ctx.rank])
            while len(self._sends) > keep:
                self._sends.popleft().wait()

    def after_round(self, round_: int) -> None:
        """Enqueued AFTER the kernels of ``round_``: the root sends the next parameters (the calling stream has
        joined the decoder of this round), a remote rank sends its messages and receives those parameters."""
        if not self.active:
            return
        dist = self.dist
        par, nxt, stamp = round_ & 1, (round_ + 1) & 1, round_ + 1
        if self.ctx.is_root:
            for x in self.remote:
                dist.isend(self._beta[nxt], dst=x, group=self.groups[x])
        else:
            for (w, m) in self.msgs[self.ctx.rank]:
                self._sends.append(dist.isend(self._slots[(w, m, par)], dst=0, group=self.groups[self.ctx.rank]))
            self._recv_beta(nxt, stamp + 1)

    def debug_state(self) -> dict:
        """Flag words of this rank's block, copied on a side stream (readable while kernels are spinning)."""
        torch = self.torch
        lay = self.ws.lay
        if not hasattr(self, "_dbg_stream"):
            self._dbg_stream = torch.cuda.Stream(device=self.ctx.device)
        n_ctl, n_flag = 8 * 32, int(lay.n_workers) * 2 * 32          # 128-byte words
        with torch.cuda.stream(self._dbg_stream):
            ctl = self.eng._tensor_from_ptr(self.ws.local + int(lay.off_ctl), n_ctl).view(torch.int32).cpu()
            flg = self.eng._tensor_from_ptr(self.ws.local + int(lay.off_flags), n_flag).view(torch.int32).cpu()
            self._dbg_stream.synchronize()
        return {"rank": self.ctx.rank, "beta_flag": int(ctl[0]), "error": int(ctl[32]),
                "arrival_stamps": flg.view(int(lay.n_workers), 2, 32)[:, :, 0].tolist()}

    def drain(self) -> None:
        """Before a host synchronisation: make the calling stream wait for everything in flight."""
        if self.ctx.is_root:
            cur = self.torch.cuda.current_stream()
            for st in self.streams.values():
                cur.wait_stream(st)
        elif self.ctx.is_remote:
            while self._sends:
                self._sends.popleft().wait()
