"""One process per GPU: ``torch.distributed`` (NCCL on GPUs, gloo on CPU) is the control plane.

The reference launches ``mpirun -np n_procs`` with rank 0 as a dedicated master
(main.py:16-18, 55-57).  Here ``world_size`` is the number of GPUs; the ``n_procs - 1`` logical
workers are mapped round-robin onto the ranks (``owner(w) = w % world``) and rank 0 additionally
hosts the decoder.  Inside one node (one NVLink / NVSwitch domain) NCCL is used only for bootstrap
(handle exchange, barriers) and by the measured baseline (``parallel/nccl_baseline.py``) -- the hot
path is peer stores.  Ranks on other nodes than rank 0 (the reference's deployment is one MPI rank
per EC2 host, tools/pytorch_ec2.py) are reached through stream-ordered NCCL point-to-point
transfers (``parallel/bridge.py``); ``node_of`` records which rank lives where.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any, List, Optional

import torch
import torch.distributed as dist


@dataclass
class DistContext:
    rank: int = 0
    world: int = 1
    device: Any = None
    backend: str = "none"
    node_of: Optional[List[int]] = None        # node id of every rank (None: one node)

    @property
    def is_root(self) -> bool:
        return self.rank == 0

    def owner(self, worker: int) -> int:
        return worker % self.world

    def local_workers(self, n_workers: int) -> List[int]:
        return [w for w in range(n_workers) if self.owner(w) == self.rank]

    # ---- topology: the root's node is the peer-memory domain, every other rank is "remote"
    def node(self, rank: int) -> int:
        return 0 if self.node_of is None else self.node_of[rank]

    @property
    def domain_ranks(self) -> List[int]:
        """Ranks whose GPUs share rank 0's node (peer-mapped workspace), rank 0 first."""
        return [r for r in range(self.world) if self.node(r) == self.node(0)]

    @property
    def remote_ranks(self) -> List[int]:
        """Ranks on other nodes: served by the NCCL bridge."""
        return [r for r in range(self.world) if self.node(r) != self.node(0)]

    @property
    def is_remote(self) -> bool:
        return self.node(self.rank) != self.node(0)

    def barrier(self) -> None:
        if self.world > 1:
            if self.backend == "nccl":
                dist.barrier(device_ids=[self.device.index])
            else:
                dist.barrier()

    def all_gather_object(self, obj) -> list:
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj)
        return out

    def broadcast_object(self, obj, src: int = 0):
        if self.world == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src)
        return box[0]


def init_distributed(device: Optional[str] = None) -> DistContext:
    """Join the process group described by the torchrun environment (RANK / WORLD_SIZE /
    LOCAL_RANK / MASTER_ADDR / MASTER_PORT); single-process when those are absent."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = torch.cuda.is_available() and device != "cpu"
    dev = torch.device("cuda", local % max(1, torch.cuda.device_count())) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    backend = "none"
    if world > 1:
        backend = "nccl" if use_cuda else "gloo"
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            kw = {"device_id": dev} if use_cuda else {}
            dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return DistContext(rank=rank, world=world, device=dev, backend=backend, node_of=node_ids(rank, world))


def node_ids(rank: int, world: int) -> Optional[List[int]]:
    """Node id of every rank, in order of first appearance.  ``EH_NODE_SIZE=k`` declares k consecutive
    ranks per node (what torchrun produces; also lets one machine stand in for several nodes in tests);
    otherwise ranks are grouped by host name (one collective)."""
    if world == 1:
        return None
    k = int(os.environ.get("EH_NODE_SIZE", "0") or 0)
    if k > 0:
        ids = [r // k for r in range(world)]
    else:
        import socket
        hosts = [None] * world
        dist.all_gather_object(hosts, socket.gethostname())
        first = {}
        ids = [first.setdefault(h, len(first)) for h in hosts]
    return ids if len(set(ids)) > 1 else None
