"""The training engine: one code path for all seven schemes, two models, GD/AGD.

Reference counterpart: the ten ``*_logistic_regression`` / ``*_linear_regression`` drivers
 -- data assignment, master loop, worker loop and evaluation epilogue each.  Here

* the scheme is a :class:`~erasurehead_b200.schemes.SchemeSpec` row;
* worker compute + encode + ship is one fused kernel per message (``csrc/grad_dense.cu`` /
  ``grad_ell.cu``), the master side is one decoder kernel per round (``csrc/decoder.cu``);
* the round loop itself runs natively (``eh_run_rounds``): the host enqueues all rounds and
  synchronises once; ordering between GPUs is carried by stamped flags in peer-mapped memory.

``Engine(device='cpu')`` (or a machine without CUDA) executes the float64 oracle instead, so
the CLI, result writers, checkpointing and multi-process host logic are testable anywhere.
"""
from __future__ import annotations

import ctypes as C
import json
import time
from math import comb
from typing import Dict, Optional

import numpy as np

from . import codes, oracle
from .config import RunConfig
from .delays import delay_table
from .models import get_model
from .parallel.context import DistContext, init_distributed
from .schemes import (DECODE_LSTSQ, MSG_CODED, MSG_PRIVATE, SchemeSpec, build_scheme)

NEVER_NS = (1 << 64) - 1


class _nvtx:
    """NVTX range around an engine phase (shows up in Nsight Systems / ncu --nvtx); no-op on CPU."""

    def __init__(self, engine, name: str):
        self.on = bool(getattr(engine, "on_gpu", False))
        self.nvtx = engine.torch.cuda.nvtx if self.on else None
        self.name = name

    def __enter__(self):
        if self.on:
            self.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        if self.on:
            self.nvtx.range_pop()
        return False
DECODE_TABLE_LIMIT = 50000        # rows of the precomputed decode table (C(n, s))


def _pad32(d: int) -> int:
    return (d + 31) // 32 * 32


def _mc_cn(C_: int) -> int:
    if not 1 <= C_ <= 32:
        raise ValueError("multi-column models support 1..32 columns on the GPU")
    return 16 if C_ <= 16 else 32


def _mc_flat(d: int, C_: int) -> int:
    return _mc_cn(C_) * _pad32(d)


def pack_mc(W: np.ndarray) -> np.ndarray:
    """``[..., d, C]`` -> flat class-major padded ``[..., CN * ldw]`` (layout of csrc/grad_mc_tc.cu)."""
    d, C_ = W.shape[-2:]
    out = np.zeros(W.shape[:-2] + (_mc_cn(C_), _pad32(d)), dtype=W.dtype)
    out[..., :C_, :d] = np.swapaxes(W, -1, -2)
    return out.reshape(W.shape[:-2] + (-1,))


def unpack_mc(flat: np.ndarray, d: int, C_: int) -> np.ndarray:
    W = flat.reshape(flat.shape[:-1] + (_mc_cn(C_), _pad32(d)))[..., :C_, :d]
    return np.ascontiguousarray(np.swapaxes(W, -1, -2))


ELL_MC_CP = 8      # padded classes per feature of the one-hot multi-column kernel (csrc/grad_ell_mc.cu)


def pack_mc_ell(W: np.ndarray) -> np.ndarray:
    """``[..., d, C]`` -> flat column-major ``[..., d * 8]`` (layout of csrc/grad_ell_mc.cu)."""
    d, C_ = W.shape[-2:]
    if C_ > ELL_MC_CP:
        raise ValueError(f"multi-column models on one-hot features support up to {ELL_MC_CP} columns")
    out = np.zeros(W.shape[:-2] + (d, ELL_MC_CP), dtype=W.dtype)
    out[..., :C_] = W
    return out.reshape(W.shape[:-2] + (-1,))


def unpack_mc_ell(flat: np.ndarray, d: int, C_: int) -> np.ndarray:
    return np.ascontiguousarray(flat.reshape(flat.shape[:-1] + (d, ELL_MC_CP))[..., :C_])


class Engine:
    """Owns the device state of one run on one rank."""

    def __init__(self, cfg: RunConfig, data=None, ctx: Optional[DistContext] = None,
                 device: Optional[str] = None, delays: Optional[np.ndarray] = None):
        self.cfg = cfg
        self.ctx = ctx or init_distributed(device)
        self.model = get_model(cfg.model_name, cfg.classes)
        # the code matrix must be identical on every rank: seed it, or let the root draw it
        B = None
        if cfg.seed is None and cfg.scheme in ("coded", "partial_coded"):
            B = self.ctx.broadcast_object(
                codes.cyclic_code(cfg.n_workers, cfg.n_stragglers) if self.ctx.is_root else None)
        self.spec: SchemeSpec = build_scheme(cfg, B)
        self.data = data
        self.first_round = 0
        self.on_gpu = self.ctx.device.type == "cuda"
        self.synth_shape = self._synthetic_onehot_shape() if self.data is None else None
        if data is None and cfg.synthetic and not self.on_gpu:
            # no GPU: --synthetic falls back to the host generators (generate_data.py distribution / data/shapes.py)
            task = {"linear": "linear", "softmax": "softmax"}.get(self.model.name, "logistic")
            kw = dict(classes=self.model.n_out) if task == "softmax" else {}
            if self.synth_shape is not None:
                from .data import generate_onehot
                self.data = generate_onehot(cfg.n_rows, cfg.n_cols, self.synth_shape["nnz"], cfg.n_partitions,
                                            seed=cfg.seed or 0, task=task, name=cfg.dataset, **kw)
                self.synth_shape = None
            else:
                from .data import generate_gmm
                self.data = generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=cfg.seed or 0, task=task, **kw)
        beta0 = oracle.initial_beta(self.spec, cfg.n_cols, cfg.seed, self.model.n_out)
        self.beta0 = self.ctx.broadcast_object(beta0 if self.ctx.is_root else None)
        self.u0 = np.zeros_like(self.beta0)
        if cfg.resume:
            self._load_checkpoint(cfg.resume)
        # delay rows are indexed by (round - first_round); the stream is a function of the round
        self.delays = delays if delays is not None else delay_table(
            cfg.rounds - self.first_round, self.spec.n, mode=cfg.delay_mode, scale=cfg.delay_scale,
            n_stragglers=cfg.n_stragglers, fixed_sleep=cfg.delay_fixed, enabled=bool(cfg.add_delay),
            first_round=self.first_round)
        if cfg.delay_carryover and cfg.add_delay:
            # backlog of workers that were still inside their delay when a round ended (starts empty on a resume)
            self.delays = oracle.carryover_delays(self.spec, self.delays)
        self.result: Optional[oracle.RunResult] = None
        self.launches = 0
        # multi-column models (softmax): parameters travel as one flat class-major vector
        self.mc = self.model.n_out > 1
        self.sparse = bool(self.data.is_sparse) if self.data is not None else self.synth_shape is not None
        if self.mc and self.sparse:
            if self.model.n_out > ELL_MC_CP:
                raise ValueError(f"multi-column models on one-hot features support up to {ELL_MC_CP} columns")
            self.d_eff = cfg.n_cols * ELL_MC_CP
        else:
            self.d_eff = _mc_flat(cfg.n_cols, self.model.n_out) if self.mc else cfg.n_cols
        if self.on_gpu:
            self._setup_gpu()

    def _synthetic_onehot_shape(self):
        """``--synthetic`` with ``is_real=1``: one-hot data of the named dataset's shape, generated on the
        device (ones per row from data/shapes.py; any other dataset name needs ``--synthetic-nnz``)."""
        cfg = self.cfg
        if not (cfg.synthetic and cfg.is_real):
            return None
        from .data.shapes import DATASET_SHAPES
        nnz = cfg.synthetic_nnz or DATASET_SHAPES.get(cfg.dataset, {}).get("nnz", 0)
        if not nnz:
            raise ValueError(f"--synthetic with is_real=1 needs a known dataset shape {sorted(DATASET_SHAPES)} "
                             "or --synthetic-nnz=K (ones per row)")
        return {"nnz": int(nnz)}

    def _pack(self, W: np.ndarray) -> np.ndarray:
        """Host ``[..., d, C]`` parameters -> the flat device layout of this run's multi-column kernel."""
        return pack_mc_ell(W) if self.sparse else pack_mc(W)

    def _unpack(self, flat: np.ndarray) -> np.ndarray:
        f = unpack_mc_ell if self.sparse else unpack_mc
        return f(flat, self.cfg.n_cols, self.model.n_out)

    # ------------------------------------------------------------------ device set-up
    def _setup_gpu(self) -> None:
        import torch
        from .ops import _lib, grad as gops
        from .parallel.workspace import Workspace

        cfg, spec, ctx = self.cfg, self.spec, self.ctx
        dev = ctx.device
        self.torch = torch
        L = _lib.lib()
        d = cfg.n_cols
        self.ws = Workspace(ctx, spec.n, spec.n_msgs, self.d_eff)
        lay = self.ws.lay
        if cfg.decoder not in ("concurrent", "inline"):
            raise ValueError("decoder must be concurrent or inline")
        # worker kernels stamp completion with their own GPU's clock: align the clocks once
        self.ws.sync_clocks()
        # the decoder's SMs: the root's gradient kernels leave them free (no effect on other ranks)
        self.dec_ctas = int(L.eh_decoder_ctas(int(lay.d_pad))) if cfg.decoder == "concurrent" else 0
        # with ranks on other nodes the root also keeps a few SMs free for the NCCL receives of their messages,
        # so that those do not queue behind the root's own gradient kernels
        n_remote = len(ctx.remote_ranks)
        br_sms = min(8, 2 * n_remote) if cfg.decoder == "concurrent" else 0
        _lib.check(L.eh_set_reserved_sms(self.dec_ctas + br_sms if ctx.is_root else 0), "eh_set_reserved_sms")
        if n_remote and (cfg.cuda_graph or cfg.beta_push == "multicast"):
            raise ValueError("ranks on several nodes: --cuda-graph and --beta-push=multicast need one NVLink domain")
        # ---- partitions resident on this GPU (deduplicated across local workers)
        self.parts: Dict[int, tuple] = {}
        mine = ctx.local_workers(spec.n)
        needed = sorted({sg.partition for w in mine for sg in spec.segments[w]})
        self.bf16 = cfg.dtype == "bf16" and not self.sparse
        if self.bf16 and self.mc and _pad32(cfg.n_cols) > 640:
            raise ValueError("--dtype=bf16 with a multi-column model needs n_cols <= 640 "
                             "(one-pass tensor-core kernel); use fp32 features for wider models")
        if cfg.dtype not in ("fp32", "bf16"):
            raise ValueError("dtype must be fp32 or bf16")
        for p in needed:
            self.parts[p] = self._device_partition(p)
        # ---- worker jobs (private message first: it is sent first, partial_coded.py:229,234)
        jobs = []
        for w in mine:
            for kind in (MSG_PRIVATE, MSG_CODED):
                segs = [sg for sg in spec.segments[w] if sg.msg == kind]
                if not segs:
                    continue
                if len(segs) > _lib.EH_MAX_SEGS:
                    raise NotImplementedError(
                        f"more than {_lib.EH_MAX_SEGS} partitions per message (worker {w}); "
                        "use fewer partitions per worker")
                j = _lib.WorkerJob()
                if self.mc:
                    jkind = 5 if self.sparse else (4 if self.bf16 else 2)
                else:
                    jkind = 1 if self.sparse else (3 if self.bf16 else 0)
                j.worker, j.msg, j.sparse, j.model = w, kind, jkind, self.model.device_id
                j.classes = self.model.n_out
                j.n_segs, j.rows_per_seg, j.d = len(segs), spec.rows_per_part, d
                X0 = self.parts[segs[0].partition][0]
                j.ld_or_nnz = X0.shape[1] if self.sparse else X0.stride(0)
                j.bias_col = (d - 1) if self.sparse else -1
                for i, sg in enumerate(segs):
                    Xp, yp = self.parts[sg.partition]
                    j.d_X[i], j.d_y[i], j.coeff[i] = Xp.data_ptr(), yp.data_ptr(), float(sg.coeff)
                jobs.append(j)
        self._jobs = (_lib.WorkerJob * max(1, len(jobs)))(*jobs)
        self.n_jobs = len(jobs)
        ld_max = max([_pad32(d)] + [self.parts[p][0].stride(0) for p in needed if not self.sparse])
        sb = max(L.eh_grad_dense_scratch_bytes(int(ld_max)), L.eh_grad_ell_scratch_bytes(int(d)))
        if self.mc and self.sparse:
            sb = max(sb, L.eh_grad_ell_mc_scratch_bytes(int(d)))
        elif self.mc:
            sb = max(sb, L.eh_grad_mc_scratch_bytes(int(spec.rows_per_part), int(d), int(self.model.n_out)))
        self.scratch = gops.Scratch(sb, dev)
        self.timeout_ns = int(120e9)
        # ---- optional NVLS multicast parameter buffer (torch symmetric memory provides the mapping)
        self.beta_mc_ptr = self.beta_local_ptr = 0
        if cfg.beta_push == "multicast" and ctx.world > 1:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            self._symm_beta = symm.empty(2 * int(lay.d_pad), dtype=torch.float32, device=dev)
            self._symm_handle = symm.rendezvous(self._symm_beta, dist.group.WORLD.group_name)
            mc = getattr(self._symm_handle, "multicast_ptr", 0)
            if not mc:
                raise RuntimeError("--beta-push=multicast: this system exposes no NVLS multicast mapping")
            self.beta_mc_ptr = int(mc)
            self.beta_local_ptr = int(self._symm_handle.buffer_ptrs[ctx.rank])
            self._symm_beta.zero_()
        elif cfg.beta_push not in ("p2p", "multicast"):
            raise ValueError("beta_push must be p2p or multicast")
        # ---- decoder state (root)
        self.dec = None
        if ctx.is_root:
            self._setup_decoder()
        plan = _lib.RoundPlan()
        plan.n_jobs = self.n_jobs
        plan.abandon = int(bool(cfg.abandon))
        plan.jobs = C.cast(self._jobs, C.POINTER(_lib.WorkerJob))
        plan.d_ws_local, plan.d_ws_root = self.ws.local, self.ws.root
        plan.d_scratch, plan.scratch_bytes = self.scratch.ptr(), self.scratch.nbytes
        plan.d_beta_local = self.beta_local_ptr or None
        plan.decoder = C.pointer(self.dec) if self.dec is not None else None
        plan.timeout_ns = self.timeout_ns
        plan.lay = lay
        self._dec_stream = None
        if self.dec is not None and cfg.decoder == "concurrent":
            _, hi = torch.cuda.Stream.priority_range()          # (least, greatest) = (0, -N)
            self._dec_stream = torch.cuda.Stream(device=dev, priority=hi)
            self._ev_fork = torch.cuda.Event(enable_timing=False)
            self._ev_join = torch.cuda.Event(enable_timing=False)
            # the events are created lazily by torch: record once so that the handles exist
            self._ev_fork.record()
            self._ev_join.record()
            plan.dec_stream = self._dec_stream.cuda_stream
            plan.ev_fork, plan.ev_join = self._ev_fork.cuda_event, self._ev_join.cuda_event
        self.plan = plan
        # ---- ranks outside the root's node: NCCL bridge between the workspace blocks (parallel/bridge.py)
        self.bridge = None
        if n_remote:
            from .parallel.bridge import NcclBridge
            self.bridge = NcclBridge(self)

    def _device_partition(self, p: int):
        """(X or ELL idx, y) of partition ``p`` on this GPU (fp32, bf16 when ``--dtype=bf16``, int32 ELL)."""
        X, y = self._device_partition_fp32(p)
        if getattr(self, "bf16", False):
            X = X.to(self.torch.bfloat16)
        return X, y

    def _synth_partition(self, rows: int, p: int):
        """Partition ``p`` (or the test set) generated on this GPU: counter-based, so a partition replicated on
        several GPUs is bit-identical everywhere.  Dense GMM (generate_data.py distribution) or one-hot ELL of the
        named dataset's shape; labels +-1, real targets, or class ids for a multi-column model."""
        from .ops import synth
        cfg, dev = self.cfg, self.ctx.device
        if self.mc and self.model.name != "softmax":
            raise ValueError("on-device synthetic labels exist for logistic, linear and softmax models")
        if self.mc and self.model.n_out > 8:
            raise ValueError("on-device synthetic multi-class labels support up to 8 classes; pass a dataset")
        task = 2 if self.mc else (1 if self.model.name == "linear" else 0)
        classes = self.model.n_out if self.mc else 0
        if self.synth_shape is not None:
            if not hasattr(self, "_t_edges"):
                self._t_edges = self.torch.from_numpy(
                    synth.onehot_edges(cfg.n_cols, self.synth_shape["nnz"], cfg.seed or 0)).to(dev)
            return synth.synth_onehot(rows, cfg.n_cols, self.synth_shape["nnz"], cfg.seed or 0, p, task, classes, dev,
                                      edges=self._t_edges)
        return synth.synth_gmm(rows, cfg.n_cols, cfg.seed or 0, p, task, dev, classes=classes)

    def _device_partition_fp32(self, p: int):
        torch = self.torch
        dev = self.ctx.device
        cfg, spec = self.cfg, self.spec
        if self.data is None:
            if not cfg.synthetic:
                raise ValueError("no dataset given and cfg.synthetic is off")
            return self._synth_partition(spec.rows_per_part, p)
        y = torch.as_tensor(np.ascontiguousarray(self.data.labels(p)), dtype=torch.float32).to(dev)
        if self.sparse:
            idx = torch.from_numpy(np.ascontiguousarray(self.data.ell(p))).to(dev)
            return idx, y
        Xh = np.asarray(self.data.partition(p), dtype=np.float32)
        ld = _pad32(cfg.n_cols)
        X = torch.zeros((Xh.shape[0], ld), dtype=torch.float32, device=dev)
        X[:, :cfg.n_cols] = torch.from_numpy(Xh).to(dev)
        return X, y

    def _setup_decoder(self) -> None:
        torch = self.torch
        from .ops import _lib, codegen
        cfg, spec, dev = self.cfg, self.spec, self.ctx.device
        n, d, R = spec.n, self.d_eff, cfg.rounds
        t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a), dtype=dt).to(dev)
        self.t_group = t(spec.group_of, torch.int32)
        self.t_B = t(spec.B, torch.float64) if spec.B is not None else None
        self.t_A = self.t_binom = None
        use_table = (spec.decode == DECODE_LSTSQ and cfg.decode != "solve"
                     and comb(n, spec.s) <= DECODE_TABLE_LIMIT)
        if cfg.decode == "table" and spec.decode == DECODE_LSTSQ and not use_table:
            raise ValueError("decode table too large; use --decode=solve")
        if use_table:
            self.t_A = codegen.decode_table(self.t_B, spec.s)
            self.t_binom = t(codes.binomial_table(n), torch.int64)
        dl = np.where(np.isfinite(self.delays), np.minimum(self.delays, 1e6) * 1e9, 0).astype(np.uint64)
        dl[~np.isfinite(self.delays)] = np.uint64(NEVER_NS)
        self.t_delay = torch.from_numpy(dl.view(np.int64)).to(dev)
        self.t_eta = t(cfg.learning_rates(), torch.float64)
        self.t_beta64 = t(self._pack(self.beta0) if self.mc else self.beta0, torch.float64)
        self.t_u64 = t(self._pack(self.u0) if self.mc else self.u0, torch.float64)
        self.t_betaset = torch.zeros((R, d), dtype=torch.float64, device=dev)
        self.t_timeset = torch.zeros(R, dtype=torch.float64, device=dev)
        self.t_wts = torch.full((R, n), -1.0, dtype=torch.float64, device=dev)
        self.t_weights = torch.zeros((R, n), dtype=torch.float64, device=dev)
        self.t_ncol = torch.zeros(R, dtype=torch.int32, device=dev)
        self.t_tstart = torch.zeros(1, dtype=torch.int64, device=dev)
        self.t_done = torch.full((R, n), -1.0, dtype=torch.float64, device=dev)
        off = [self.ws.clock_offset_ns[self.ctx.owner(w)] for w in range(n)]
        self.t_clock_off = t(np.asarray(off, dtype=np.int64), torch.int64)
        self.t_decscratch = torch.zeros((_lib.lib().eh_decoder_scratch_bytes() + 7) // 8, dtype=torch.int64, device=dev)
        D = _lib.DecoderDesc()
        D.n_workers, D.n_msgs, D.n_groups, D.s = n, spec.n_msgs, spec.n_groups, spec.s
        D.need_private, D.need_count, D.need_groups, D.decode = (spec.need_private, spec.need_count,
                                                                 spec.need_groups, spec.decode)
        D.d, D.d_pad = d, self.ws.lay.d_pad
        D.update_agd = int(cfg.update_rule == "AGD")
        dom = self.ctx.domain_ranks                    # the decoder pushes parameters to the GPUs of its own node
        D.n_peers = len(dom)
        D.d_group_of = self.t_group.data_ptr()
        D.d_B = self.t_B.data_ptr() if self.t_B is not None else None
        D.d_A = self.t_A.data_ptr() if self.t_A is not None else None
        D.d_binom = self.t_binom.data_ptr() if self.t_binom is not None else None
        D.d_delay_ns, D.d_eta = self.t_delay.data_ptr(), self.t_eta.data_ptr()
        D.d_beta64, D.d_u64 = self.t_beta64.data_ptr(), self.t_u64.data_ptr()
        D.d_ws_local = self.ws.local
        for i, r in enumerate(dom):
            D.d_ws_peer[i] = self.ws.peers[r]
        D.d_betaset, D.d_timeset = self.t_betaset.data_ptr(), self.t_timeset.data_ptr()
        D.d_worker_timeset, D.d_weights = self.t_wts.data_ptr(), self.t_weights.data_ptr()
        D.d_ncollected, D.d_tstart = self.t_ncol.data_ptr(), self.t_tstart.data_ptr()
        D.d_scratch = self.t_decscratch.data_ptr()
        D.d_beta_mc = self.beta_mc_ptr or None
        D.alpha, D.n_rows, D.lr_scale = float(cfg.alpha), float(cfg.n_rows), float(spec.lr_scale)
        D.timeout_ns = self.timeout_ns
        D.lay = self.ws.lay
        D.d_clock_off = self.t_clock_off.data_ptr() if self.ctx.world > 1 else None
        D.d_done = self.t_done.data_ptr()
        D.dec_ctas = self.dec_ctas
        self.dec = D

    # ------------------------------------------------------------------ running
    def launches_per_round(self) -> int:
        """Kernels this rank launches per round: gradient kernels (dense messages of one shape are
        batched 8 messages / 32 partitions per launch by ``eh_run_rounds``) + the decoder on the root."""
        if not self.on_gpu:
            return 0
        k = 0
        if self.mc and self.sparse:                               # batched 16 messages / 64 partitions per launch
            nb = segs = 0
            for q in range(self.n_jobs):
                ns = self._jobs[q].n_segs
                if nb and (nb == 16 or segs + ns > 64):
                    k += 1
                    nb = segs = 0
                nb += 1
                segs += ns
            k += 1 if nb else 0
        elif self.mc:
            # one launch per message for the one-pass kernels, else wait + 2 per partition + ship + flag
            one_pass = self.bf16 or _pad32(self.cfg.n_cols) <= 1024
            k = sum(1 if one_pass else 2 * self._jobs[q].n_segs + 3 for q in range(self.n_jobs))
        elif self.sparse:                                          # batched 16 messages / 64 partitions per launch
            nb = segs = 0
            for q in range(self.n_jobs):
                ns = self._jobs[q].n_segs
                if nb and (nb == 16 or segs + ns > 64):
                    k += 1
                    nb = segs = 0
                nb += 1
                segs += ns
            k += 1 if nb else 0
        else:
            nb = segs = 0
            big_prev = False
            for q in range(self.n_jobs):
                j = self._jobs[q]
                ns = j.n_segs
                big = ns * j.rows_per_seg * j.ld_or_nnz * 4 > (16 << 20)       # same rule as csrc/engine.cu
                if nb and (big or big_prev or nb == 8 or segs + ns > 32):
                    k += 1
                    nb = segs = 0
                nb += 1
                segs += ns
                big_prev = big
            k += 1 if nb else 0
        return k + (1 if self.ctx.is_root else 0)

    def prepare(self) -> None:
        """Reset flags and publish the initial parameters (untimed set-up of a run)."""
        if not self.on_gpu:
            return
        from .ops import _lib
        self.ws.reset()
        if self.cfg.debug_poison and self.ctx.is_root:
            self._poison_slots()
        self.scratch.zero_()
        if self.dec is not None:
            self.t_decscratch.zero_()
        self.torch.cuda.synchronize()
        self.ctx.barrier()
        if self.dec is not None:
            _lib.check(_lib.lib().eh_decoder_publish(C.byref(self.dec), self.first_round, _lib.stream_ptr()),
                       "eh_decoder_publish")
        if self.bridge is not None:
            self.bridge.publish(self.first_round)
        self._cursor = self.first_round

    def _poison_slots(self) -> None:
        """Debug aid (compute-sanitizer is not available on every pool): NaN-fill all gradient slots
        so that a decoder read of a slot no worker wrote in this run poisons the iterates, which
        :meth:`finish` then reports."""
        lay = self.ws.lay
        n = int(lay.n_workers) * int(lay.n_msgs) * 2 * int(lay.d_pad)
        self._tensor_from_ptr(self.ws.local + int(lay.off_slots), n).fill_(float("nan"))

    def _tensor_from_ptr(self, ptr: int, n_floats: int):
        """A float32 torch view of raw device memory owned by the workspace."""
        class _Holder:
            pass
        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}
        return self.torch.as_tensor(h, device=self.ctx.device)

    def enqueue(self, n_rounds: int) -> None:
        """Enqueue ``n_rounds`` further rounds on the current stream (no synchronisation)."""
        from .ops import _lib
        r0 = self._cursor - self.first_round
        if self.cfg.cuda_graph:
            # one graph for all n_rounds rounds, launched on a side stream (capture needs a real stream)
            torch = self.torch
            if not hasattr(self, "_graph_stream"):
                self._graph_stream = torch.cuda.Stream()
                self._graphs = []
            self._graph_stream.wait_stream(torch.cuda.current_stream())
            exec_ = C.c_void_p()
            _lib.check(_lib.lib().eh_run_rounds_graph(C.byref(self.plan), self._cursor, n_rounds, r0,
                                                      self._graph_stream.cuda_stream, C.byref(exec_)),
                       "eh_run_rounds_graph")
            self._graphs.append(exec_)
            torch.cuda.current_stream().wait_stream(self._graph_stream)
        elif self.bridge is not None:
            # several nodes: the transfers of a round are enqueued between the rounds' kernels (still no host wait)
            for it in range(n_rounds):
                self.bridge.before_round(self._cursor + it)
                _lib.check(_lib.lib().eh_run_rounds(C.byref(self.plan), self._cursor + it, 1, r0 + it,
                                                    _lib.stream_ptr()), "eh_run_rounds")
                self.bridge.after_round(self._cursor + it)
        else:
            _lib.check(_lib.lib().eh_run_rounds(C.byref(self.plan), self._cursor, n_rounds, r0, _lib.stream_ptr()),
                       "eh_run_rounds")
        self._cursor += n_rounds
        self.launches += n_rounds * self.launches_per_round()

    def step(self, delay_row: Optional[np.ndarray] = None, eta: Optional[float] = None):
        """One round through the public streaming API (used by ``bench.py`` for the end-to-end
        number).  The root copies this round's host inputs (the per-worker delay row and the
        learning rate) from pinned memory to the device tables, every rank enqueues one round,
        and the root reads the round's result -- the new iterate and the round time -- back
        into pinned host memory.  Returns ``(beta[d] float64, seconds)`` on the root, else None.
        Call :meth:`prepare` once before the first step."""
        torch = self.torch
        r = self._cursor - self.first_round
        if self.dec is not None:
            if not hasattr(self, "_pin_in"):
                n, d = self.spec.n, self.d_eff
                self._pin_in = torch.empty(n + 1, dtype=torch.int64).pin_memory()
                self._pin_out = torch.empty(d + 1, dtype=torch.float64).pin_memory()
                self.h2d_bytes = (n + 1) * 8
                self.d2h_bytes = (d + 1) * 8
            n = self.spec.n
            row = self.delays[r] if delay_row is None else np.asarray(delay_row, dtype=np.float64)
            ns = np.where(np.isfinite(row), np.minimum(row, 1e6) * 1e9, 0).astype(np.uint64)
            ns[~np.isfinite(row)] = np.uint64(NEVER_NS)
            self._pin_in[:n] = torch.from_numpy(ns.view(np.int64))
            e = float(self.cfg.learning_rates()[self._cursor] if eta is None else eta)
            self._pin_in[n:] = torch.from_numpy(np.array([e], dtype=np.float64).view(np.int64))
            self.t_delay[r].copy_(self._pin_in[:n], non_blocking=True)
            self.t_eta[r:r + 1].view(torch.int64).copy_(self._pin_in[n:], non_blocking=True)
        self.enqueue(1)
        if self.dec is not None:
            d = self.d_eff
            self._pin_out[:d].copy_(self.t_betaset[r], non_blocking=True)
            self._pin_out[d:].copy_(self.t_timeset[r:r + 1], non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return self._pin_out[:d].numpy(), float(self._pin_out[d])
        return None

    def finish(self) -> None:
        if getattr(self, "bridge", None) is not None:
            self.bridge.drain()
        self.torch.cuda.synchronize()
        for g in getattr(self, "_graphs", []):
            from .ops import _lib
            _lib.lib().eh_graph_destroy(g)
        if hasattr(self, "_graphs"):
            self._graphs = []
        err = self.ws.error()
        errs = self.ctx.all_gather_object(err)
        if any(errs):
            raise RuntimeError(f"device watchdog fired (per-rank error words {errs}); "
                               "a worker died or a stop rule became unsatisfiable")
        if self.cfg.debug_poison and self.dec is not None and not bool(self.torch.isfinite(self.t_beta64).all()):
            raise RuntimeError("debug_poison: the decoder consumed a slot that no worker wrote (stale read)")

    def run(self, rounds: Optional[int] = None) -> Optional[oracle.RunResult]:
        """Run ``rounds`` rounds (default: all remaining).  Returns the result on the root."""
        cfg = self.cfg
        R = (cfg.rounds - self.first_round) if rounds is None else rounds
        t0 = time.time()
        if not self.on_gpu:
            res = None
            if self.ctx.is_root:
                res = oracle.run_rounds(cfg, self.spec, self.data, self.delays[:R], beta0=self.beta0, u0=self.u0,
                                        first_round=self.first_round, rounds=R, model=self.model)
                res.wall_time = time.time() - t0
                if cfg.checkpoint_every > 0 and R > 0:
                    # same file at the same cadence as the device path: the state after the last full chunk
                    k = R if R % cfg.checkpoint_every == 0 else R - R % cfg.checkpoint_every
                    if k > 0:
                        self.result = res if k == R else oracle.RunResult(
                            betaset=res.betaset[:k], timeset=res.timeset[:k], worker_timeset=res.worker_timeset[:k],
                            weights=res.weights[:k], n_collected=res.n_collected[:k],
                            u=self._u_after(res, k), beta=res.betaset[k - 1])
                        self.save_checkpoint(self._ckpt_path(), self.first_round + k)
            self.result = res
            self.ctx.barrier()
            return res
        with _nvtx(self, "eh.prepare"):
            self.prepare()
        chunk = cfg.checkpoint_every if cfg.checkpoint_every > 0 else R
        done = 0
        while done < R:
            k = min(chunk, R - done)
            with _nvtx(self, f"eh.rounds[{self.first_round + done}:{self.first_round + done + k}]"):
                self.enqueue(k)
                self.finish()
            done += k
            if cfg.checkpoint_every > 0 and self.ctx.is_root:
                self.save_checkpoint(self._ckpt_path(), self.first_round + done)
        self.ctx.barrier()
        res = self._collect(R) if self.ctx.is_root else None
        if res is not None:
            res.wall_time = time.time() - t0
        self.result = res
        return res

    def _u_after(self, res: oracle.RunResult, k: int) -> np.ndarray:
        """AGD momentum state after ``k`` rounds of ``res`` (recomputed from the stored iterates:
        ``u_{i+1} = beta_i + (beta_{i+1} - beta_i) / theta_i``, naive.py:119-123)."""
        if self.cfg.update_rule != "AGD" or k >= len(res.betaset):
            return res.u
        i = self.first_round + k - 1
        prev = res.betaset[k - 2] if k >= 2 else self.beta0
        theta = 2.0 / (i + 2.0)
        return prev + (res.betaset[k - 1] - prev) * (1.0 / theta)

    def _collect(self, R: int) -> oracle.RunResult:
        cpu = lambda t: t[:R].detach().cpu().numpy()
        betaset, u, beta = cpu(self.t_betaset), self.t_u64.cpu().numpy(), self.t_beta64.cpu().numpy()
        if self.mc:
            betaset, u, beta = self._unpack(betaset), self._unpack(u), self._unpack(beta)
        return oracle.RunResult(
            betaset=betaset, timeset=cpu(self.t_timeset), worker_timeset=cpu(self.t_wts),
            weights=cpu(self.t_weights), n_collected=cpu(self.t_ncol).astype(np.int64), u=u, beta=beta,
            done=cpu(self.t_done))

    # ------------------------------------------------------------------ evaluation
    def evaluate(self, res: Optional[oracle.RunResult] = None):
        """Train/test loss (or MSE) and test AUC of every stored iterate (root only)."""
        from . import results
        res = res or self.result
        if not self.ctx.is_root or res is None:
            return None
        if not self.on_gpu:
            return results.evaluate_host(self.model, self.data, res.betaset, self._n_eval_parts())
        with _nvtx(self, "eh.evaluate"):
            return self._evaluate_gpu(res.betaset)

    def _n_eval_parts(self):
        """Q1: the reference evaluates on files 1..n_procs-2 only; None = all partitions."""
        return max(1, self.spec.n - 1) if self.cfg.compat else None

    def _evaluate_gpu_mc(self, betaset: np.ndarray):
        """Cross-entropy of every iterate: the forward tensor-core product + loss epilogue."""
        from . import results
        from .ops import mc
        torch = self.torch
        dev, cfg, spec = self.ctx.device, self.cfg, self.spec
        R, d, C_ = betaset.shape
        n_parts = self._n_eval_parts() or spec.n_partitions
        Xt, yt = self._device_test()
        if self.sparse:
            return self._evaluate_gpu_mc_ell(betaset, n_parts, Xt, yt)
        # sums stay on the device; one host read per quantity at the end
        tr = torch.zeros(R, dtype=torch.float64, device=dev)
        te = torch.zeros(R, dtype=torch.float64, device=dev)
        acc = torch.zeros(R, dtype=torch.float64, device=dev)
        from .ops import metrics as mops
        Wall = torch.from_numpy(pack_mc(betaset).astype(np.float32)).to(dev)            # [R, CN * ldw]
        WT = torch.zeros((R, C_, _pad32(d)), dtype=torch.float32, device=dev)           # class-major rows for eval_tc
        WT[:, :, :d] = torch.from_numpy(np.ascontiguousarray(np.swapaxes(betaset, 1, 2))).to(dev)
        rows = 0
        parts = []
        for p in range(n_parts):
            Xp, yp = self.parts[p] if p in self.parts else self._device_partition(p)
            parts.append((Xp.float() if Xp.dtype != torch.float32 else Xp, yp))      # the loss-only kernels read fp32 rows
            rows += Xp.shape[0]
        for i in range(R):
            for Xp, yp in parts:
                tr[i] += mc.loss_mc(Xp, yp, Wall[i], d, C_, self.model.device_id)
            te[i] = mc.loss_mc(Xt, yt, Wall[i], d, C_, self.model.device_id)
            # test accuracy: margins of all classes through the tensor-core evaluation kernel, argmax per row
            _, M = mops.eval_tc(Xt, WT[i], None, d=d, want_margins=True)
            acc[i] = (M[:, :C_].argmax(dim=1) == yt.long()).double().mean()
        tr, te, acc = (tr / rows).cpu().numpy(), (te / Xt.shape[0]).cpu().numpy(), acc.cpu().numpy()
        return results.Evaluation(tr, te, None, acc)

    def _evaluate_gpu_mc_ell(self, betaset: np.ndarray, n_parts: int, Xt, yt):
        """One-hot features, C-column model: the ELL margin kernel evaluates the C columns of up to 32 / C iterates
        per pass over the indices; cross-entropy and accuracy are reductions of the logits on the device, one host
        read per quantity at the end."""
        from . import results
        from .ops import metrics as mops
        torch = self.torch
        dev = self.ctx.device
        R, d, C_ = betaset.shape
        per = max(1, 32 // C_)
        tr = torch.zeros(R, dtype=torch.float64, device=dev)
        te = torch.zeros(R, dtype=torch.float64, device=dev)
        acc = torch.zeros(R, dtype=torch.float64, device=dev)
        rows_tr = 0

        def stats(idx, y, Bt, n_it):
            P = mops.margins_ell(idx, Bt)[:, :n_it * C_].reshape(idx.shape[0], n_it, C_).double()
            yl = y.long()
            ce = torch.logsumexp(P, dim=2) - P.gather(2, yl[:, None, None].expand(-1, n_it, 1)).squeeze(2)
            hit = (P.argmax(dim=2) == yl[:, None]).double()
            return ce.sum(dim=0), hit.sum(dim=0)

        for i0 in range(0, R, per):
            n_it = min(per, R - i0)
            Bt = torch.zeros((n_it * C_, _pad32(d)), dtype=torch.float32, device=dev)
            Wb = torch.from_numpy(np.ascontiguousarray(np.swapaxes(betaset[i0:i0 + n_it], 1, 2))).to(dev)   # [n_it, C, d]
            Bt[:, :d] = Wb.reshape(n_it * C_, d).float()
            for p in range(n_parts):
                Xp, yp = self.parts[p] if p in self.parts else self._device_partition(p)
                ce, _ = stats(Xp, yp, Bt, n_it)
                tr[i0:i0 + n_it] += ce
                if i0 == 0:
                    rows_tr += Xp.shape[0]
            ce, hit = stats(Xt, yt, Bt, n_it)
            te[i0:i0 + n_it] = ce / Xt.shape[0]
            acc[i0:i0 + n_it] = hit / Xt.shape[0]
        return results.Evaluation((tr / rows_tr).cpu().numpy(), te.cpu().numpy(), None, acc.cpu().numpy())

    def _evaluate_gpu(self, betaset: np.ndarray):
        from . import results
        from .ops import metrics as mops
        if self.mc:
            return self._evaluate_gpu_mc(betaset)
        torch = self.torch
        dev = self.ctx.device
        cfg, spec = self.cfg, self.spec
        R, d = betaset.shape
        Bt = torch.zeros((R, _pad32(d)), dtype=torch.float32, device=dev)
        Bt[:, :d] = torch.from_numpy(betaset).to(dev)
        kind = 1 if self.model.name == "linear" else 0
        n_parts = self._n_eval_parts() or spec.n_partitions                                # Q1
        tr = torch.zeros(R, dtype=torch.float64, device=dev)
        rows = 0
        # dense data: one tcgen05/TMEM pass over X evaluates all R iterates with the loss fused in
        # the epilogue (csrc/eval_tc.cu); --eval-precision=fp32 selects the exact CUDA-core path
        use_tc = (not self.sparse) and cfg.eval_precision == "tf32"
        for p in range(n_parts):
            Xp, yp = self.parts[p] if p in self.parts else self._device_partition(p)
            if Xp.dtype == torch.bfloat16:
                Xp = Xp.float()                     # evaluation kernels take fp32 rows (exact widening)
            if use_tc:
                mops.eval_tc(Xp, Bt, yp, kind=kind, d=d, loss_sums=tr)
            else:
                P = mops.margins_ell(Xp, Bt) if self.sparse else mops.margins_dense(Xp, Bt, d)
                tr += mops.loss_columns(P, yp, kind) * Xp.shape[0]
                del P
            rows += Xp.shape[0]
        tr /= rows
        Xt, yt = self._device_test()
        if use_tc:
            te, Pt = mops.eval_tc(Xt, Bt, yt, kind=kind, d=d, want_margins=True)
            te = te / Xt.shape[0]
        else:
            Pt = mops.margins_ell(Xt, Bt) if self.sparse else mops.margins_dense(Xt, Bt, d)
            te = mops.loss_columns(Pt, yt, kind)
        auc = None
        if self.model.has_auc:
            Pcol = Pt.t().contiguous()
            auc = mops.auc_batch(Pcol[:R], yt).cpu().numpy()            # all iterates: one sort, two scans
        acc = None
        if self.model.name != "linear":
            pred = torch.where(Pt[:, :R] > 0, 1.0, -1.0)
            acc = (pred == yt[:, None]).float().mean(dim=0).cpu().numpy().astype(np.float64)
        return results.Evaluation(tr.cpu().numpy(), te.cpu().numpy(), auc, acc)

    def _device_test(self):
        torch = self.torch
        dev = self.ctx.device
        cfg = self.cfg
        if self.data is None:
            n_test = max(1, int(0.2 * cfg.n_rows))                         # generate_data.py:40
            return self._synth_partition(n_test, (1 << 40))
        yt = torch.as_tensor(np.ascontiguousarray(self.data.y_test), dtype=torch.float32).to(dev)
        if self.sparse:
            return torch.from_numpy(np.ascontiguousarray(self.data.ell(test=True))).to(dev), yt
        Xh = np.asarray(self.data.X_test, dtype=np.float32)
        X = torch.zeros((Xh.shape[0], _pad32(cfg.n_cols)), dtype=torch.float32, device=dev)
        X[:, :cfg.n_cols] = torch.from_numpy(Xh).to(dev)
        return X, yt

    # ------------------------------------------------------------------ checkpoint / resume
    def _ckpt_path(self) -> str:
        from .utils import io
        return io.ensure_dir(self.cfg.results_dir) + f"{self.spec.prefix}_checkpoint.npz"

    def save_checkpoint(self, path: str, next_round: int) -> None:
        """Persist ``(next round, beta, u, B)``; the delay stream is a pure function of the round
        index so no RNG state is needed (delays.py)."""
        if self.on_gpu:
            beta, u = self.t_beta64.cpu().numpy(), self.t_u64.cpu().numpy()
            if self.mc:
                beta, u = self._unpack(beta), self._unpack(u)
        else:
            beta, u = self.result.beta, self.result.u
        np.savez(path, next_round=next_round, beta=beta, u=u,
                 B=self.spec.B if self.spec.B is not None else np.zeros((0, 0)),
                 scheme=self.spec.name, update_rule=self.cfg.update_rule)

    def _load_checkpoint(self, path: str) -> None:
        ck = np.load(path, allow_pickle=False)
        if str(ck["scheme"]) != self.spec.name:
            raise ValueError(f"checkpoint is for scheme {ck['scheme']}, not {self.spec.name}")
        self.first_round = int(ck["next_round"])
        self.beta0, self.u0 = ck["beta"], ck["u"]
        if ck["B"].size:
            self.spec = build_scheme(self.cfg, ck["B"])

    # ------------------------------------------------------------------ observability
    def write_metrics_jsonl(self, path: str, res: oracle.RunResult, ev=None) -> None:
        with open(path, "w") as f:
            for r in range(len(res.timeset)):
                rec = {"round": self.first_round + r, "time_s": float(res.timeset[r]),
                       "n_collected": int(res.n_collected[r]),
                       "arrived_mask": int(sum(1 << w for w in np.flatnonzero(res.worker_timeset[r] >= 0))),
                       "weights": [float(x) for x in res.weights[r]]}
                if ev is not None:
                    rec["train_loss"] = float(ev.training_loss[r])
                    rec["test_loss"] = float(ev.testing_loss[r])
                    if ev.auc is not None:
                        rec["auc"] = float(ev.auc[r])
                    if getattr(ev, "accuracy", None) is not None:
                        rec["test_accuracy"] = float(ev.accuracy[r])
                f.write(json.dumps(rec) + "\n")

    def close(self) -> None:
        if self.on_gpu and getattr(self, "ws", None) is not None:
            self.torch.cuda.synchronize()
            self.bridge = None                 # its tensors alias the workspace block
            self.ws.close()
            self.ws = None
