#!/usr/bin/env python
"""Headline benchmark (driver contract in the task statement).

Metric (BASELINE.json): iterations/second -- reported as whole-job training samples/second --
of approximate gradient coding (ErasureHead / AGC) with injected stragglers on dense synthetic
logistic regression, d = 4096 fp32.  Weak scaling: every GPU hosts two logical workers
(n = 2N workers, s = 1 so FRC groups of two spanning different GPUs when N >= 2, n_collect = N),
each worker streams s+1 = 2 partitions of ROWS x 4096 fp32, so every GPU reads 4 x ROWS x 16 KiB
from HBM per round (>> L2, no flush needed).  A step is one full round: parameter broadcast ->
fused gradient+encode+ship kernels -> decoder (arrival polling with the injected per-message
exponential delay, stop rule, decode, AGD update, parameter push).

    python bench.py [--gpus N --steps K --warmup W] [--impl ours|reference|nccl]

N > 1: launched by ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``.
Timing: CUDA events on every rank around exactly K rounds, max over ranks.

Besides the headline the JSON line carries (each skippable, see the --no-* flags):

  baseline_nccl   the torch NCCL + cuBLAS arm (B2) on the same configuration, same K / W, in this process;
                  ``vs_baseline`` = headline value / its value (BASELINE.md has no published number)
  time_to_target  BASELINE.json's metric: seconds until the training loss reaches the loss uncoded GD has
                  after 60 % of the rounds -- uncoded, AGC and exact cyclic-MDS coding, product and B2
  parity          a small run of AGC and exact coding through the Engine on these N GPUs against the
                  float64 oracle: max |iterate error| and equality of the per-round arrival sets
  matched         the headline at the reference arm's size (--ref-rows rows per partition) without delays
                  and with the reference's Exp(0.5 s) delays: the same-configuration counterpart of
                  ``bench.py --impl reference``
  abandon         the headline with --abandon: workers whose round has been decided stop streaming (a straggler's
                  compute is skipped, not only its message); fewer rows are read per round, so this is reported
                  next to the headline, never as the headline
  covtype         AGC at the reference launcher's real-data configuration (30 workers, n_collect 15,
                  one-hot 396112 x 15509, AGD, no injected delay) -- same data generator and seed as
                  the reference arm's ``covtype`` block
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import numpy as np  # noqa: E402

D = 4096
ROWS_DEFAULT = 524288            # rows per partition: 8 GiB fp32 per partition
DELAY_SCALE = 2e-3               # mean of the exponential message delay (s); reference model, ms scale


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl"])
    ap.add_argument("--rows", type=int, default=ROWS_DEFAULT, help="rows per partition")
    ap.add_argument("--cols", type=int, default=D)
    ap.add_argument("--workers-per-gpu", type=int, default=2)
    ap.add_argument("--stragglers", type=int, default=1)
    ap.add_argument("--scheme", default="approx")
    ap.add_argument("--delay-scale", type=float, default=DELAY_SCALE)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--beta-push", default="p2p", choices=["p2p", "multicast"])
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"], help="feature storage (headline: fp32)")
    ap.add_argument("--ref-rows", type=int, default=1024,
                    help="rows per partition of the reference arm and of the repo arm's `matched` block")
    ap.add_argument("--no-covtype", action="store_true", help="skip the Covertype-shaped comparison block")
    ap.add_argument("--no-matched", action="store_true", help="skip the runs at the reference arm's size")
    ap.add_argument("--no-nccl", action="store_true", help="skip the in-process NCCL + cuBLAS arm (vs_baseline)")
    ap.add_argument("--no-ttt", action="store_true", help="skip the time-to-target-loss block")
    ap.add_argument("--no-parity", action="store_true", help="skip the engine-vs-oracle parity block")
    ap.add_argument("--quick", action="store_true", help="headline only (all the --no-* flags)")
    ap.add_argument("--decoder", default="concurrent", choices=["concurrent", "inline"])
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (profiling recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


def reference_arm(args):
    """Run the UNMODIFIED reference (baseline/_ref) through its own CLI on the box's CPUs.  Under
    torchrun only rank 0 works (the reference starts its own 2N+1 processes); the others exit."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    from baseline.reference_arm import run_reference
    print(json.dumps(run_reference(args)), flush=True)
    return 0


def timed(ctx, fn_step, steps):
    import torch
    ctx.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn_step(steps)
    e1.record()
    torch.cuda.synchronize(); ctx.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=ctx.device)
    if ctx.world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    return float(ms.item())


def release(eng):
    import gc
    import torch
    eng.close()
    del eng
    gc.collect()
    torch.cuda.empty_cache()


def headline_cfg(args, N, rows, rounds, delay_scale, add_delay=1, scheme=None, **kw):
    from erasurehead_b200 import RunConfig, triple_from_scheme
    n = args.workers_per_gpu * N
    s = args.stragglers if n % (args.stragglers + 1) == 0 and n > args.stragglers else 0
    k = max(1, n // 2)
    scheme = scheme or args.scheme
    if scheme == "naive":
        s = 0
    ic, pp, cv = triple_from_scheme(scheme)
    cfg = RunConfig(n + 1, n * rows, args.cols, "", 0, "x", ic, s, pp, cv, k, add_delay, "AGD", rounds=rounds,
                    synthetic=True, seed=1, delay_scale=delay_scale, quiet=True, dtype=args.dtype,
                    beta_push=args.beta_push, decoder=args.decoder, **kw)
    return cfg, n, s, k


def measure(ctx, eng, K, W, e2e=True):
    """Device-timed K rounds (enqueued natively, one sync) and, optionally, K rounds through step()."""
    eng.prepare()
    eng.enqueue(W)
    eng.finish()
    l0 = eng.launches
    ms = timed(ctx, lambda kk: eng.enqueue(kk), K)
    launches = eng.launches - l0
    eng.finish()
    ms_e2e = None
    if e2e:
        for _ in range(W):
            eng.step()

        def run_steps(kk):
            for _ in range(kk):
                eng.step()
        ms_e2e = timed(ctx, run_steps, K)
        eng.finish()
    return ms, ms_e2e, launches


def block_matched(ctx, args, N, K, W):
    """The headline at the reference arm's size: no delays, and the reference's stock Exp(0.5 s)."""
    from erasurehead_b200 import Engine
    out = {"rows_per_partition": args.ref_rows, "note": "same scheme / n / s / n_collect / AGD / d as `--impl reference` "
           "at the reference's rows per partition; device-timed K rounds, max over ranks"}
    for name, add_delay, scale in (("no_delay", 0, 0.5), ("injected_delay", 1, 0.5)):
        cfg, n, s, k = headline_cfg(args, N, args.ref_rows, 2 * (W + K) + 2, scale, add_delay)
        eng = Engine(cfg, None, ctx=ctx)
        ms, ms_e2e, _ = measure(ctx, eng, K, W)
        out[name] = {"value": cfg.n_rows * K / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / K,
                     "e2e_value": cfg.n_rows * K / (ms_e2e * 1e-3), "e2e_ms_per_step": ms_e2e / K,
                     "delay": "none" if not add_delay else "exp(scale=0.5s) seeded per round (reference's table)"}
        release(eng)
    return out


def block_time_to_target(ctx, args, N, with_nccl):
    """BASELINE.json's metric.  Delay-dominated regime like the reference's experiments: 1 GiB partitions
    (0.16 ms of streaming each), Exp(2 ms) message delays, 60 rounds from a zero start, eta = 10 (main.py:37)."""
    import torch
    from erasurehead_b200 import Engine
    from erasurehead_b200.parallel.nccl_baseline import NcclBaselineEngine
    R, rows = 60, 65536
    runs, evaluator = [], None
    for scheme in ("naive", "approx", "coded"):
        cfg, n, s, k = headline_cfg(args, N, rows, R, args.delay_scale, 1, scheme=scheme, init="zeros")
        eng = Engine(cfg, None, ctx=ctx)
        res = eng.run()
        ev = eng.evaluate(res)
        rec = {"scheme": scheme, "n_workers": n, "s": s, "n_collect": k if scheme == "approx" else None}
        if ctx.is_root:
            rec.update(loss=ev.training_loss, ts=res.timeset, mean_round_ms=float(res.timeset.mean() * 1e3))
        if with_nccl:
            b = NcclBaselineEngine(cfg, None, ctx=ctx)
            bs, ts = b.run(R)
            if ctx.is_root:
                evb = eng.evaluate(type(res)(betaset=bs, timeset=ts, worker_timeset=res.worker_timeset,
                                             weights=res.weights, n_collected=res.n_collected))
                rec.update(nccl_loss=evb.training_loss, nccl_ts=ts, nccl_mean_round_ms=float(ts.mean() * 1e3))
            b.close()
            del b
        runs.append(rec)
        release(eng)
    if not ctx.is_root:
        return None
    target = float(runs[0]["loss"][int(0.6 * R)])

    def ttt(ts, loss):
        hit = np.flatnonzero(loss <= target)
        return (float(np.sum(ts[:hit[0] + 1])), int(hit[0]) + 1) if hit.size else (None, None)
    out = {"target_train_loss": target, "rounds": R, "rows_per_partition": rows, "cols": args.cols,
           "delay": f"exp(scale={args.delay_scale}s) seeded per round", "target": "training loss of uncoded GD after 60% of the rounds",
           "schemes": {}}
    for rec in runs:
        t, r = ttt(rec["ts"], rec["loss"])
        e = {"time_to_target_s": t, "rounds_to_target": r, "mean_round_ms": rec["mean_round_ms"], "s": rec["s"],
             "n_collect": rec["n_collect"], "final_train_loss": float(rec["loss"][-1])}
        if "nccl_ts" in rec:
            tb, rb = ttt(rec["nccl_ts"], rec["nccl_loss"])
            e.update(nccl_time_to_target_s=tb, nccl_rounds_to_target=rb, nccl_mean_round_ms=rec["nccl_mean_round_ms"],
                     speedup_vs_nccl=(tb / t) if (t and tb) else None)
        out["schemes"]["uncoded" if rec["scheme"] == "naive" else {"approx": "agc", "coded": "exact_gc"}[rec["scheme"]]] = e
    return out


def block_parity(ctx, args, N):
    """A small AGC and exact-coding run through the Engine on these N GPUs vs the float64 oracle."""
    from erasurehead_b200 import Engine, RunConfig, build_scheme, oracle, triple_from_scheme
    from erasurehead_b200.data import generate_gmm
    n = args.workers_per_gpu * N
    rows, d, R = n * 240, 96, 8
    out = {"ok": True, "rounds": R, "n_workers": n, "rows": rows, "cols": d, "schemes": {}}
    for scheme, s, k in (("approx", 1, max(1, n // 2)), ("coded", 1, 0)):
        ic, pp, cv = triple_from_scheme(scheme)
        cfg = RunConfig(n + 1, rows, d, "", 0, "x", ic, s, pp, cv, k, 1, "AGD", rounds=R, seed=3, quiet=True,
                        decoder=args.decoder)
        data = generate_gmm(rows, d, cfg.n_partitions, seed=2)
        rng = np.random.RandomState(11)
        dl = np.zeros((R, n))
        for r in range(R):
            dl[r, rng.permutation(n)] = 1.5e-3 * (1 + np.arange(n))       # distinct arrival order every round
        eng = Engine(cfg, data, ctx=ctx, delays=dl)
        res = eng.run()
        if ctx.is_root:
            ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0)
            err = float(np.abs(res.betaset - ref.betaset).max())
            scale = float(max(1.0, np.abs(ref.betaset).max()))
            same = bool(np.array_equal(res.worker_timeset >= 0, ref.worker_timeset >= 0)
                        and np.array_equal(res.n_collected, ref.n_collected))
            ok = bool(err <= 5e-4 * scale and same)
            out["schemes"][scheme] = {"max_abs_iterate_error": err, "iterate_scale": scale, "arrival_sets_equal": same, "ok": ok}
            out["ok"] = out["ok"] and ok
        release(eng)
    return out if ctx.is_root else None


def block_covtype(ctx, args, K, W):
    """AGC at the reference launcher's configuration on the Covertype shape (see baseline/reference_arm.py)."""
    from baseline.reference_arm import COVTYPE as c
    from erasurehead_b200 import Engine, RunConfig
    from erasurehead_b200.data.shapes import generate_onehot
    t0 = time.time()
    data = generate_onehot(c["n_rows"], c["n_cols"], c["nnz"], c["n_procs"] - 1, seed=1, name="covtype")
    t_gen = time.time() - t0
    cfg = RunConfig(c["n_procs"], c["n_rows"], c["n_cols"], "", 1, "covtype", 1, c["s"], 0, 3, c["k"], 0, "AGD",
                    rounds=2 * (W + K) + 2, seed=1, quiet=True, decoder=args.decoder)
    eng = Engine(cfg, data, ctx=ctx)
    ms, ms_e2e, _ = measure(ctx, eng, K, W)
    out = {"value": c["n_rows"] * K / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / K,
           "e2e_value": c["n_rows"] * K / (ms_e2e * 1e-3), "e2e_ms_per_step": ms_e2e / K, "host_data_gen_s": t_gen,
           "config": dict(c, scheme="approx", update_rule="AGD", add_delay=0,
                          data="synthetic one-hot of the Covertype shape (same generator and seed as the reference arm)")}
    release(eng)
    return out


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)
    if args.quick:
        args.no_matched = args.no_nccl = args.no_ttt = args.no_parity = args.no_covtype = True
    import torch
    from erasurehead_b200 import Engine
    from erasurehead_b200.parallel.context import init_distributed

    ctx = init_distributed()
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device", "impl": args.impl}))
        return 1
    N = ctx.world
    K, W = args.steps, max(3, args.warmup)
    cfg, n, s, k = headline_cfg(args, N, args.rows, 2 * (W + K) + 2, args.delay_scale)
    n_rows = cfg.n_rows
    if args.impl == "nccl":
        from erasurehead_b200.parallel.nccl_baseline import NcclBaselineEngine as Eng
    else:
        Eng = Engine
    eng = Eng(cfg, None, ctx=ctx)
    dev = ctx.device
    sampler = ClockSampler(dev.index) if ctx.is_root else None
    ms, ms_e2e, launches = measure(ctx, eng, K, W, e2e=not args.no_e2e)
    clocks = sampler.stop() if sampler else None
    e2e = None
    if ms_e2e is not None and ctx.is_root:
        e2e = {"value": n_rows * K / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e / K,
               "h2d_bytes_per_step": int(eng.h2d_bytes), "d2h_bytes_per_step": int(eng.d2h_bytes),
               "note": "full-batch GD: the dataset is device-resident (loaded once, as the reference does "
                       "before its loop); per-step host inputs are the delay row + learning rate, the "
                       "per-step result is the new iterate + round time"}
    clock_sync = None
    if args.impl == "ours" and ctx.is_root and N > 1:
        clock_sync = {"peer_minus_root_ns": eng.ws.clock_offset_ns, "best_round_trip_ns": eng.ws.clock_rtt_ns}
    release(eng)
    lt = torch.tensor([launches], dtype=torch.int64, device=dev)
    if N > 1:
        torch.distributed.all_reduce(lt)
    value = n_rows * K / (ms * 1e-3)
    extra = {}

    def guarded(name, fn):
        try:
            r = fn()
        except Exception as e:                      # a side block must never cost the headline
            r = {"error": f"{type(e).__name__}: {str(e)[-300:]}"}
        if r is not None:
            extra[name] = r

    if args.impl == "ours":
        if not args.no_nccl:
            def nccl():
                from erasurehead_b200.parallel.nccl_baseline import NcclBaselineEngine
                b = NcclBaselineEngine(cfg, None, ctx=ctx)
                bms, _, _ = measure(ctx, b, K, W, e2e=False)
                release(b)
                return {"value": n_rows * K / (bms * 1e-3), "unit": "samples/s", "ms_per_step": bms / K,
                        "what": "torch NCCL broadcast/gather + cuBLAS GEMVs, same configuration, steps and warm-up"}
            guarded("baseline_nccl", nccl)
        if not args.no_ttt:
            guarded("time_to_target", lambda: block_time_to_target(ctx, args, N, with_nccl=not args.no_nccl))
        if not args.no_parity:
            guarded("parity", lambda: block_parity(ctx, args, N))
        if not args.no_matched:
            guarded("matched", lambda: block_matched(ctx, args, N, K, W))
        if not args.no_matched:
            def abandon():
                c2, _, _, _ = headline_cfg(args, N, args.rows, 2 * (W + K) + 2, args.delay_scale, abandon=True)
                e2 = Engine(c2, None, ctx=ctx)
                ams, _, _ = measure(ctx, e2, K, W, e2e=False)
                release(e2)
                return {"ms_per_step": ams / K, "value": n_rows * K / (ams * 1e-3), "unit": "samples/s (rows of decided "
                        "rounds that were not read still count: compare ms_per_step, not this, with the headline)"}
            guarded("abandon", abandon)
        if not args.no_covtype:
            guarded("covtype", lambda: block_covtype(ctx, args, K, W))
    if ctx.is_root:
        hbm_bytes = args.workers_per_gpu * (s + 1) * args.rows * args.cols * (2 if args.dtype == "bf16" else 4)
        vs_baseline = None
        if isinstance(extra.get("baseline_nccl"), dict) and extra["baseline_nccl"].get("value"):
            vs_baseline = value / extra["baseline_nccl"]["value"]
        out = {
            "metric": "agc_train_samples_per_sec_injected_stragglers", "value": value, "unit": "samples/s",
            "n_gpus": N, "steps": K, "warmup": W, "ms_per_step": ms / K, "iters_per_sec": K / (ms * 1e-3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": vs_baseline,
            "vs_baseline_note": "BASELINE.md publishes no number; denominator = baseline_nccl (B2) measured in this process",
            "dtype": args.dtype,
            "data": "synthetic (device-generated GMM, generate_data.py distribution), random-init weights",
            "impl": args.impl,
            "config": {"model": "logistic_regression", "scheme": cfg.scheme, "n_workers": n, "n_stragglers": s,
                       "n_collect": k, "update_rule": "AGD", "rows_per_partition": args.rows, "cols": args.cols,
                       "global_batch": n_rows, "seq_len": None, "parallelism": f"coded-dp{N}",
                       "delay": f"exp(scale={args.delay_scale}s) seeded per round",
                       "decoder": args.decoder,
                       "l2": "inputs >> L2 (per-GPU stream per round = %.1f GiB), no flush" % (hbm_bytes / 2**30)},
            "hbm_gbs_per_gpu": hbm_bytes / (ms / K * 1e-3) / 1e9,
            "gpu_launches": int(lt.item()), "clocks": clocks, "e2e": e2e, "clock_sync": clock_sync,
        }
        t2t = extra.get("time_to_target")
        if isinstance(t2t, dict) and "schemes" in t2t:
            out["time_to_target_s"] = {kk: v["time_to_target_s"] for kk, v in t2t["schemes"].items()}
        out.update(extra)
        print(json.dumps(out), flush=True)
    if N > 1:
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
