#!/usr/bin/env python
"""The benchmark configurations named in BASELINE.json, measured with device timers.

    python benchmarks/configs.py [--only exact,sweep,amazon,multi] [--out FILE] [--quick]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/configs.py ...

  exact   exact cyclic-MDS gradient coding, dense synthetic logistic regression, 8 workers, s = 2,
          product vs the NCCL + cuBLAS baseline (B2), rounds/s
  sweep   approximate FRC coding (ErasureHead) vs uncoded vs ignore-stragglers vs exact schemes under
          exponential injected delays, n_collect sweep: mean round time and time-to-target-loss
  amazon  Amazon-Employee-Access-shaped one-hot logistic regression (26215 x 241915, 44 ones/row),
          AGD, partial_coded vs partial_replication
  multi   Covertype-shaped multi-class softmax (7 classes) and KC-Housing-shaped linear regression
          (17290 x 27654 one-hot), replication vs coded

Every number: CUDA events / device %globaltimer round times, max over ranks, >= 3 warm-up rounds.
One JSON line per measurement; rank 0 prints.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from erasurehead_b200 import Engine, RunConfig, triple_from_scheme  # noqa: E402
from erasurehead_b200.data import generate_gmm, generate_onehot  # noqa: E402
from erasurehead_b200.delays import expected_kth_wait  # noqa: E402
from erasurehead_b200.parallel.context import init_distributed  # noqa: E402

OUT = []


def emit(ctx, rec):
    if ctx.is_root:
        print(json.dumps(rec), flush=True)
        OUT.append(rec)


def mkcfg(scheme, n, s, rows, d, k=0, P=0, rounds=40, add_delay=1, **kw):
    ic, pp, cv = triple_from_scheme(scheme, P)
    return RunConfig(n + 1, rows, d, "", 0, "x", ic, s, pp, cv, k, add_delay, kw.pop("rule", "AGD"), rounds=rounds,
                     quiet=True, **kw)


def timed_rounds(ctx, eng, warm, steps):
    eng.prepare()
    eng.enqueue(warm)
    eng.finish()
    ctx.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.enqueue(steps); e1.record()
    eng.finish()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=ctx.device)
    if ctx.world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    return float(ms.item()) / steps


def time_to_target(timeset, losses, target):
    hit = np.flatnonzero(losses <= target)
    if hit.size == 0:
        return None, None
    r = int(hit[0])
    return float(np.sum(timeset[:r + 1])), r + 1


def bench_exact(ctx, a):
    from erasurehead_b200.parallel.nccl_baseline import NcclBaselineEngine
    n, s, d = 8, 2, 4096
    rows_pp = a.rows_exact
    for impl in ("ours", "nccl"):
        cfg = mkcfg("coded", n, s, n * rows_pp, d, rounds=40, synthetic=True, seed=1, delay_scale=a.delay_scale)
        eng = (Engine if impl == "ours" else NcclBaselineEngine)(cfg, None, ctx=ctx)
        ms = timed_rounds(ctx, eng, 3, 10)
        per_gpu_bytes = (n / ctx.world) * (s + 1) * rows_pp * d * 4
        emit(ctx, {"config": "exact_cyclic_mds", "impl": impl, "gpus": ctx.world, "n_workers": n, "s": s,
                   "rows_per_partition": rows_pp, "d": d, "ms_per_round": ms, "rounds_per_s": 1e3 / ms,
                   "delay": f"exp({a.delay_scale}s)", "expected_delay_wait_ms": expected_kth_wait(n, n - s, a.delay_scale) * 1e3,
                   "hbm_GBps_per_gpu_incl_wait": per_gpu_bytes / ms / 1e6})
        eng.close()
        del eng
        torch.cuda.empty_cache()


def bench_sweep(ctx, a):
    n, d, rows_pp, R = 8, 1024, a.rows_sweep, a.rounds
    data = None
    runs = [("naive", 0, 0), ("avoidstragg", 1, 0), ("avoidstragg", 2, 0), ("replication", 1, 0), ("coded", 1, 0),
            ("coded", 2, 0)] + [("approx", 1, k) for k in (4, 5, 6, 7)] + [("approx", 3, 2), ("approx", 3, 4)]
    results = []
    for scheme, s, k in runs:
        cfg = mkcfg(scheme, n, s, n * rows_pp, d, k=k, rounds=R, synthetic=True, seed=1, delay_scale=a.delay_scale,
                    eta0=a.eta0)
        eng = Engine(cfg, None, ctx=ctx)
        eng.beta0 = np.zeros(d)                      # same start for every scheme
        if eng.dec is not None:
            eng.t_beta64.zero_()
        res = eng.run()
        ev = eng.evaluate(res)
        if ctx.is_root:
            results.append((scheme, s, k, res, ev))
        eng.close()
        del eng
        torch.cuda.empty_cache()
    if ctx.is_root:
        base = [r for r in results if r[0] == "naive"][0]
        target = float(base[4].training_loss[int(0.6 * R)])
        for scheme, s, k, res, ev in results:
            ttt, nr = time_to_target(res.timeset, ev.training_loss, target)
            emit(ctx, {"config": "delay_sweep", "scheme": scheme, "s": s, "n_collect": k, "gpus": ctx.world,
                       "n_workers": n, "rows_per_partition": rows_pp, "d": d, "delay": f"exp({a.delay_scale}s)",
                       "mean_round_ms": float(res.timeset.mean() * 1e3), "target_train_loss": target,
                       "time_to_target_s": ttt, "rounds_to_target": nr,
                       "final_train_loss": float(ev.training_loss[-1]), "final_auc": float(ev.auc[-1]),
                       "mean_collected": float(res.n_collected.mean())})


def bench_sweep_nccl(ctx, a):
    """Round time of the NCCL + cuBLAS arm (B2) for the sweep's schemes: a gather is a collective, so every
    message is waited for whatever the stop rule says; with the rounds-to-target of the product run (the
    iterates are the same) this gives B2's time to target."""
    from erasurehead_b200.parallel.nccl_baseline import NcclBaselineEngine
    n, d, rows_pp = 8, 1024, a.rows_sweep
    for scheme, s, k in (("naive", 0, 0), ("coded", 2, 0), ("approx", 1, 4), ("approx", 3, 2)):
        cfg = mkcfg(scheme, n, s, n * rows_pp, d, k=k, rounds=40, synthetic=True, seed=1, delay_scale=a.delay_scale,
                    eta0=a.eta0)
        eng = NcclBaselineEngine(cfg, None, ctx=ctx)
        ms = timed_rounds(ctx, eng, 3, 30)
        emit(ctx, {"config": "delay_sweep", "impl": "nccl", "scheme": scheme, "s": s, "n_collect": k, "gpus": ctx.world,
                   "n_workers": n, "rows_per_partition": rows_pp, "d": d, "delay": f"exp({a.delay_scale}s)",
                   "mean_round_ms": ms})
        eng.close()
        del eng
        torch.cuda.empty_cache()


def bench_amazon(ctx, a):
    n, s, P = 8, 1, 3
    rows, d, nnz = 26215, 241915, 44
    for scheme in ("partial_coded", "partial_replication"):
        cfg = mkcfg(scheme, n, s, rows, d, P=P, rounds=a.rounds, seed=1, delay_scale=a.delay_scale)
        cfg.is_real, cfg.dataset = 1, "amazon-dataset"
        data = generate_onehot(rows, d, nnz, cfg.n_partitions, seed=2, name="amazon-shaped")
        eng = Engine(cfg, data, ctx=ctx)
        res = eng.run()
        ev = eng.evaluate(res)
        if ctx.is_root:
            emit(ctx, {"config": "amazon_shaped_sparse", "scheme": scheme, "s": s, "partitions_per_worker": P,
                       "gpus": ctx.world, "rows": rows, "d": d, "nnz_per_row": nnz, "update": "AGD",
                       "delay": f"exp({a.delay_scale}s)", "mean_round_ms": float(res.timeset.mean() * 1e3),
                       "total_s": float(res.timeset.sum()), "final_train_loss": float(ev.training_loss[-1]),
                       "final_auc": float(ev.auc[-1])})
        eng.close()
    # compute-only cost of one round (no injected delay)
    cfg = mkcfg("partial_replication", n, s, rows, d, P=P, rounds=60, seed=1, add_delay=0)
    cfg.is_real, cfg.dataset = 1, "amazon-dataset"
    eng = Engine(cfg, data, ctx=ctx)
    ms = timed_rounds(ctx, eng, 5, 50)
    emit(ctx, {"config": "amazon_shaped_sparse", "scheme": "partial_replication", "delay": "none", "gpus": ctx.world,
               "ms_per_round_no_delay": ms, "message_bytes": d * 4})
    eng.close()


def bench_multi(ctx, a):
    n = 8
    # Covertype-shaped multi-class: 54 numeric features, 7 classes
    rows, d, C_ = 8 * (396112 // 8 // 8), 54, 7          # 1/8 of the rows keeps host generation short
    data = generate_gmm(rows, d, n, seed=3, task="softmax", classes=C_)
    for scheme, s in (("replication", 1), ("coded", 1)):
        cfg = mkcfg(scheme, n, s, rows, d, rounds=a.rounds, seed=1, delay_scale=a.delay_scale, model="softmax",
                    classes=C_, eta0=0.5)
        eng = Engine(cfg, data, ctx=ctx)
        res = eng.run()
        ev = eng.evaluate(res)
        emit(ctx, {"config": "covtype_shaped_multiclass", "scheme": scheme, "s": s, "gpus": ctx.world, "rows": rows,
                   "d": d, "classes": C_, "mean_round_ms": float(res.timeset.mean() * 1e3) if res else None,
                   "first_train_loss": float(ev.training_loss[0]) if ev else None,
                   "final_train_loss": float(ev.training_loss[-1]) if ev else None})
        eng.close()
    # KC-Housing-shaped linear regression (one-hot, 19 ones/row)
    rows, d, nnz = 17290, 27654, 19
    for scheme, s in (("replication", 1), ("naive", 0), ("approx", 1)):
        cfg = mkcfg(scheme, n, s, rows, d, k=6, rounds=a.rounds, seed=1, delay_scale=a.delay_scale,
                    eta0=0.1, lr_schedule="expdecay")
        cfg.is_real, cfg.dataset = 1, "kc_house_data"
        data = generate_onehot(rows, d, nnz, n, seed=4, task="linear", name="kc-shaped")
        eng = Engine(cfg, data, ctx=ctx)
        res = eng.run()
        ev = eng.evaluate(res)
        emit(ctx, {"config": "kc_house_shaped_linear", "scheme": scheme, "s": s, "gpus": ctx.world, "rows": rows, "d": d,
                   "model": eng.model.name, "mean_round_ms": float(res.timeset.mean() * 1e3) if res else None,
                   "first_train_mse": float(ev.training_loss[0]) if ev else None,
                   "final_train_mse": float(ev.training_loss[-1]) if ev else None})
        eng.close()


def bench_latency(ctx, a):
    """Protocol cost of one round: tiny data, no injected delay (launch + flags + NVLink ship + decode + push)."""
    for scheme, s, k in (("approx", 1, 4), ("coded", 2, 0), ("naive", 0, 0)):
        cfg = mkcfg(scheme, 8, s, 8 * 256, 4096, k=k, rounds=260, synthetic=True, seed=1, add_delay=0,
                    beta_push=a.beta_push, cuda_graph=a.cuda_graph)
        eng = Engine(cfg, None, ctx=ctx)
        ms = timed_rounds(ctx, eng, 20, 200)
        emit(ctx, {"config": "round_latency", "scheme": scheme, "gpus": ctx.world, "n_workers": 8, "d": 4096,
                   "rows_per_partition": 256, "us_per_round": ms * 1e3, "message_bytes": 4096 * 4,
                   "beta_push": a.beta_push, "cuda_graph": a.cuda_graph,
                   "kernels_per_round_rank0": eng.launches_per_round()})
        eng.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="exact,sweep,amazon,multi")
    ap.add_argument("--out", default="")
    ap.add_argument("--delay-scale", type=float, default=0.01)
    ap.add_argument("--rounds", type=int, default=60)
    ap.add_argument("--rows-exact", type=int, default=131072)
    ap.add_argument("--rows-sweep", type=int, default=32768)
    ap.add_argument("--eta0", type=float, default=10.0)
    ap.add_argument("--beta-push", default="p2p")
    ap.add_argument("--cuda-graph", action="store_true")
    a = ap.parse_args()
    ctx = init_distributed()
    which = a.only.split(",")
    if "exact" in which:
        bench_exact(ctx, a)
    if "sweep" in which:
        bench_sweep(ctx, a)
    if "sweep_nccl" in which:
        bench_sweep_nccl(ctx, a)
    if "amazon" in which:
        bench_amazon(ctx, a)
    if "multi" in which:
        bench_multi(ctx, a)
    if "latency" in which:
        bench_latency(ctx, a)
    if a.out and ctx.is_root:
        json.dump(OUT, open(a.out, "w"), indent=1)
    if ctx.world > 1:
        torch.distributed.destroy_process_group()
