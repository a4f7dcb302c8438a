"""Baseline B2: the same schemes with library calls only.

Per round, host-driven: NCCL broadcast of the parameters, each rank computes its workers'
encoded gradients with two cuBLAS GEMVs (dense data) or two cuSPARSE CSR matvecs (one-hot data)
+ elementwise kernels (X is read twice), NCCL gather of
the messages to rank 0, which waits out the injected delay of the last message it needs (a
device spin, same delay table as the product), decodes with torch ops and applies the update.
This is what "a solution that only calls NCCL" looks like; it is measured by
``bench.py --impl nccl`` and never used by the product path.

Because a gather is a collective, every worker's message is waited for each round: injected
delays can be emulated, genuine stragglers cannot be skipped.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import oracle
from ..config import RunConfig
from ..delays import delay_table
from ..models import get_model
from ..schemes import MSG_CODED, MSG_PRIVATE, build_scheme
from .context import DistContext, init_distributed


class NcclBaselineEngine:
    def __init__(self, cfg: RunConfig, data=None, ctx: Optional[DistContext] = None, delays=None):
        self.cfg, self.ctx = cfg, ctx or init_distributed()
        self.spec = build_scheme(cfg)
        self.model = get_model(cfg.model_name, cfg.classes)
        self.data = data
        dev = self.ctx.device
        self.dev = dev
        spec, n, d = self.spec, self.spec.n, cfg.n_cols
        self.delays = delays if delays is not None else delay_table(
            cfg.rounds, n, mode=cfg.delay_mode, scale=cfg.delay_scale, n_stragglers=cfg.n_stragglers,
            enabled=bool(cfg.add_delay))
        if cfg.delay_carryover and cfg.add_delay:
            self.delays = oracle.carryover_delays(spec, self.delays)
        self.mine = self.ctx.local_workers(n)
        self.parts = {}
        for p in sorted({sg.partition for w in self.mine for sg in spec.segments[w]}):
            self.parts[p] = self._partition(p)
        self.beta = torch.as_tensor(oracle.initial_beta(spec, d, cfg.seed), dtype=torch.float64, device=dev)
        self.u = torch.zeros_like(self.beta)
        self.beta32 = self.beta.float()
        self.eta = cfg.learning_rates()
        # every rank sends the same number of message rows (pad with zeros) so gather is regular
        counts = [len([w for w in range(n) if w % self.ctx.world == r]) for r in range(self.ctx.world)]
        self.max_local = max(counts)
        self.sendbuf = torch.zeros((self.max_local, spec.n_msgs, d), dtype=torch.float32, device=dev)
        self.recv = ([torch.zeros_like(self.sendbuf) for _ in range(self.ctx.world)]
                     if self.ctx.is_root and self.ctx.world > 1 else None)
        self.launches = 0
        self.h2d_bytes = self.d2h_bytes = 0
        self._cursor = 0
        self.timeset = []
        self._clock_ghz = 1.9
        self._pin_out = None

    def _csr(self, idx):
        """One-hot ELL indices -> torch sparse CSR (and its transpose), the library format (cuSPARSE)."""
        rows, nnz = idx.shape
        crow = torch.arange(0, rows * nnz + 1, nnz, device=self.dev, dtype=torch.int64)
        X = torch.sparse_csr_tensor(crow, idx.reshape(-1).long(), torch.ones(rows * nnz, device=self.dev),
                                    size=(rows, self.cfg.n_cols))
        return X, X.t().to_sparse_csr()

    def _partition(self, p):
        cfg = self.cfg
        task = 1 if self.model.name == "linear" else 0
        if self.data is None:
            from ..ops import synth
            if cfg.is_real:                     # --synthetic one-hot shape: same generator and seed as the product
                from ..data.shapes import DATASET_SHAPES
                nnz = cfg.synthetic_nnz or DATASET_SHAPES[cfg.dataset]["nnz"]
                idx, y = synth.synth_onehot(self.spec.rows_per_part, cfg.n_cols, nnz, cfg.seed or 0, p, task, 0, self.dev)
                return self._csr(idx), y
            X, y = synth.synth_gmm(self.spec.rows_per_part, cfg.n_cols, cfg.seed or 0, p, task, self.dev)
            return X[:, :cfg.n_cols], y
        y = torch.as_tensor(self.data.labels(p), dtype=torch.float32, device=self.dev)
        if self.data.is_sparse:
            return self._csr(torch.from_numpy(np.ascontiguousarray(self.data.ell(p))).to(self.dev)), y
        X = torch.as_tensor(np.asarray(self.data.partition(p)), dtype=torch.float32, device=self.dev)
        return X, y

    def launches_per_round(self) -> int:
        return 0          # library kernels only: none of them are ours

    def prepare(self) -> None:
        torch.cuda.synchronize()
        self.ctx.barrier()

    def _grad(self, X, y, coeff):
        if isinstance(X, tuple):                              # one-hot data: two cuSPARSE CSR matvecs
            Xs, Xt = X
            p = torch.mv(Xs, self.beta32)
            r = -2.0 * (y - p) if self.model.name == "linear" else -y / (torch.exp(p * y) + 1.0)
            return torch.mv(Xt, coeff * r)
        p = X @ self.beta32                                   # cuBLAS GEMV
        if self.model.name == "linear":
            r = -2.0 * (y - p)
        else:
            r = -y / (torch.exp(p * y) + 1.0)
        return X.t() @ (coeff * r)                            # cuBLAS GEMV (transposed): second pass over X

    def _round(self, i: int) -> None:
        ctx, spec = self.ctx, self.spec
        if ctx.world > 1:
            dist.broadcast(self.beta32, src=0)
        self.sendbuf.zero_()
        for li, w in enumerate(self.mine):
            for sg in spec.segments[w]:
                X, y = self.parts[sg.partition]
                self.sendbuf[li, sg.msg if spec.n_msgs > 1 else 0] += self._grad(X, y, sg.coeff)
        if ctx.world > 1:
            dist.gather(self.sendbuf, self.recv, dst=0)
        if ctx.is_root:
            a, priv, t_stop, _, _ = oracle.plan_round(spec, self.delays[i], round_index=i)
            if t_stop > 0:
                torch.cuda._sleep(int(t_stop * self._clock_ghz * 1e9))
            bufs = self.recv if ctx.world > 1 else [self.sendbuf]
            g = torch.zeros_like(self.beta)
            for w in range(spec.n):
                r, li = w % ctx.world, w // ctx.world
                if a[w] != 0.0:
                    g += float(a[w]) * bufs[r][li, MSG_CODED].double()
                if spec.n_msgs > 1 and priv[w] != 0.0:
                    g += bufs[r][li, MSG_PRIVATE].double()
            eta, alpha = float(self.eta[i]), float(self.cfg.alpha)
            gm = eta * spec.lr_scale / self.cfg.n_rows
            if self.cfg.update_rule == "GD":
                self.beta = (1.0 - 2.0 * alpha * eta) * self.beta - gm * g
            else:
                theta = 2.0 / (i + 2.0)
                yt = (1.0 - theta) * self.beta + theta * self.u
                bn = yt - gm * g - (2.0 * alpha * eta) * self.beta
                self.u = self.beta + (bn - self.beta) * (1.0 / theta)
                self.beta = bn
            self.beta32.copy_(self.beta)

    def enqueue(self, n_rounds: int) -> None:
        for _ in range(n_rounds):
            self._round(self._cursor)
            self._cursor += 1

    def run(self, rounds: int):
        """Run ``rounds`` rounds keeping every iterate and a CUDA-event time per round (max over ranks).
        Returns ``(betaset [R, d] float64, timeset [R] seconds)`` on the root, else ``(None, timeset)``."""
        self.prepare()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(rounds + 1)]
        keep = torch.zeros((rounds, self.cfg.n_cols), dtype=torch.float64, device=self.dev) if self.ctx.is_root else None
        ev[0].record()
        for r in range(rounds):
            self._round(self._cursor)
            self._cursor += 1
            if keep is not None:
                keep[r].copy_(self.beta)
            ev[r + 1].record()
        torch.cuda.synchronize()
        ts = torch.tensor([ev[r].elapsed_time(ev[r + 1]) * 1e-3 for r in range(rounds)], dtype=torch.float64,
                          device=self.dev)
        if self.ctx.world > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        return (keep.cpu().numpy() if keep is not None else None), ts.cpu().numpy()

    def step(self):
        i = self._cursor
        self.enqueue(1)
        if self.ctx.is_root:
            if self._pin_out is None:
                self._pin_out = torch.empty(self.cfg.n_cols, dtype=torch.float64).pin_memory()
                self.d2h_bytes = self.cfg.n_cols * 8
                self.h2d_bytes = 0
            self._pin_out.copy_(self.beta, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return self._pin_out.numpy(), 0.0
        return None

    def finish(self) -> None:
        torch.cuda.synchronize()

    def close(self) -> None:
        torch.cuda.synchronize()
