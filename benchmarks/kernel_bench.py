#!/usr/bin/env python
"""Kernel-level timings against the measured roofline (MEASURED_PEAKS.json).

    python benchmarks/kernel_bench.py [--rows R --cols D --iters I] [--only grad|decoder|eval]

Each kernel is timed with CUDA events after warm-up; inputs are larger than L2 (or L2 is
flushed between iterations for the small ones).  Prints one JSON line per kernel.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np  # noqa: E402
import torch  # noqa: E402


def peaks():
    p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
    try:
        return json.load(open(p)), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def time_fn(fn, iters, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


def bench_grad(args, pk, src):
    from erasurehead_b200.ops import grad, synth
    out = []
    for d in args.cols:
        rows = args.rows if args.rows else max(4096, int(6e9 // (4 * d)))
        X, y = synth.synth_gmm(rows, d, 1, 0)
        beta = torch.randn(d, device="cuda") * 0.01
        sc = grad.dense_scratch(X.stride(0), "cuda")
        o = torch.empty(d, device="cuda")
        med, mn = time_fn(lambda: grad.grad_dense(X, y, beta, d=d, scratch=sc, out=o), args.iters)
        byt = rows * X.stride(0) * 4
        # torch two-pass baseline (two cuBLAS GEMVs + elementwise)
        def two_pass():
            p = X[:, :d] @ beta
            r = -y / (torch.exp(p * y) + 1)
            return X[:, :d].t() @ r
        med2, _ = time_fn(two_pass, max(3, args.iters // 2))
        rec = {"kernel": "grad_dense", "rows": rows, "d": d, "ms": med, "ms_min": mn, "GBps": byt / med / 1e6,
               "frac_of_hbm_" + src: byt / med / 1e6 / pk["hbm_gbs"], "torch_two_pass_ms": med2,
               "speedup_vs_two_pass": med2 / med}
        print(json.dumps(rec)); out.append(rec)
        if d >= 2048:
            Xb = X.to(torch.bfloat16)
            medb, mnb = time_fn(lambda: grad.grad_dense(Xb, y, beta, d=d, scratch=sc, out=o), args.iters)
            bytb = rows * Xb.stride(0) * 2
            recb = {"kernel": "grad_dense(bf16 rows)", "rows": rows, "d": d, "ms": medb, "ms_min": mnb,
                    "GBps": bytb / medb / 1e6, "frac_of_hbm_" + src: bytb / medb / 1e6 / pk["hbm_gbs"],
                    "rows_per_s_vs_fp32": med / medb}
            print(json.dumps(recb)); out.append(recb)
            del Xb
        del X, y
    return out


def bench_decoder(args):
    """Latency of a full round with zero delay and a tiny problem (launch + flags + decode)."""
    from erasurehead_b200 import Engine, RunConfig, triple_from_scheme
    out = []
    for scheme, s in (("approx", 1), ("coded", 2), ("naive", 0)):
        ic, pp, cv = triple_from_scheme(scheme)
        cfg = RunConfig(9, 8 * 256, 4096, "", 0, "x", ic, s, pp, cv, 4, 0, "AGD", rounds=220, synthetic=True, seed=1)
        eng = Engine(cfg, None)
        eng.prepare(); eng.enqueue(20); eng.finish()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng.enqueue(200); e1.record(); eng.finish()
        us = e0.elapsed_time(e1) / 200 * 1e3
        dec = float(eng.t_timeset[20:220].mean().item()) * 1e6
        rec = {"kernel": "round_latency", "scheme": scheme, "n_workers": 8, "d": 4096, "us_per_round": us,
               "kernels_per_round": eng.launches_per_round(), "decoder_timeset_us": dec}
        print(json.dumps(rec)); out.append(rec)
        eng.close()
    return out


def bench_eval(args, pk, src):
    from erasurehead_b200.ops import metrics as mops, synth
    rows, d, R = 262144, 4096, 100
    X, y = synth.synth_gmm(rows, d, 1, 0)
    Bt = torch.randn(R, d, device="cuda") * 0.01
    med, _ = time_fn(lambda: mops.margins_dense(X, Bt, d), 5)
    med2, _ = time_fn(lambda: X @ Bt.t(), 5)
    fl = 2.0 * rows * d * R
    rec = {"kernel": "margins_dense", "rows": rows, "d": d, "R": R, "ms": med, "TFLOPs": fl / med / 1e9,
           "cublas_fp32_ms": med2}
    print(json.dumps(rec))
    torch.backends.cuda.matmul.allow_tf32 = True
    med3, _ = time_fn(lambda: X @ Bt.t(), 5)
    torch.backends.cuda.matmul.allow_tf32 = False
    ls = torch.zeros(R, dtype=torch.float64, device="cuda")
    medt, mint = time_fn(lambda: mops.eval_tc(X, Bt, y, kind=0, d=d, loss_sums=ls), 10)
    byt = rows * X.stride(0) * 4
    rec2 = {"kernel": "eval_tc(tcgen05 tf32 + fused loss)", "rows": rows, "d": d, "R": R, "ms": medt, "ms_min": mint,
            "TFLOPs": fl / medt / 1e9, "GBps": byt / medt / 1e6, "frac_of_hbm_" + src: byt / medt / 1e6 / pk["hbm_gbs"],
            "cublas_tf32_gemm_only_ms": med3, "cublas_fp32_gemm_only_ms": med2}
    print(json.dumps(rec2))
    return [rec, rec2]


def bench_ell(args, pk, src):
    """One-hot gradient at the reference's real-dataset shapes (synthetic data of that shape)."""
    from erasurehead_b200.data import generate_onehot, DATASET_SHAPES
    from erasurehead_b200.ops import grad
    out = []
    for name, sh in DATASET_SHAPES.items():
        ds = generate_onehot(sh["n_rows"], sh["n_cols"], sh["nnz"], 1, seed=1, task=sh["task"])
        idx = torch.from_numpy(ds.ell(0)).cuda()
        y = torch.tensor(ds.labels(0), dtype=torch.float32, device="cuda")
        d = sh["n_cols"]
        beta = torch.randn(d, device="cuda") * 0.1
        sc = grad.ell_scratch(d, "cuda")
        o = torch.empty(d, device="cuda")
        model = 1 if sh["task"] == "linear" else 0
        med, mn = time_fn(lambda: grad.grad_ell(idx, y, beta, model=model, bias_col=d - 1, scratch=sc, out=o), args.iters)
        Xs = torch.sparse_csr_tensor(torch.arange(0, idx.numel() + 1, idx.shape[1], device="cuda", dtype=torch.int64),
                                     idx.reshape(-1).long(), torch.ones(idx.numel(), device="cuda"), size=(idx.shape[0], d))
        def torch_ref():
            p = torch.mv(Xs, beta)
            r = -y / (torch.exp(p * y) + 1) if model == 0 else -2 * (y - p)
            return torch.mv(Xs.t(), r) if False else torch.sparse.mm(Xs.t(), r[:, None])
        try:
            med2, _ = time_fn(torch_ref, 3)
        except Exception:
            med2 = None
        byt = idx.numel() * 4 + idx.shape[0] * 4 + 3 * d * 4
        rec = {"kernel": "grad_ell", "shape": name, "rows": idx.shape[0], "d": d, "nnz": idx.shape[1], "ms": med,
               "ms_min": mn, "index_GBps": byt / med / 1e6, "torch_sparse_csr_ms": med2}
        print(json.dumps(rec)); out.append(rec)
    return out


def bench_mc(args, pk, src):
    """Multi-column (softmax) gradient: two tensor-core passes over X."""
    from erasurehead_b200.ops import mc, synth
    out = []
    shapes = ((262144, 4096, 16), (1048576, 1024, 10), (2097152, 64, 7), (4194304, 256, 16), (8388608, 128, 10))
    if args.mc_shapes:
        shapes = tuple(tuple(int(v) for v in sh.split(",")) for sh in args.mc_shapes)
    for rows, d, C_ in shapes:
        X, _ = synth.synth_gmm(rows, d, 1, 0)
        if args.mc_bf16:
            X = X.to(torch.bfloat16)
        y = torch.randint(0, C_, (rows,), device="cuda").float()
        Wp = mc.pack_params(torch.randn(d, C_, device="cuda") * 0.01)
        G = torch.empty_like(Wp)
        scr = mc.scratch(rows, d, C_, "cuda")
        job = mc.make_job([X], [y], Wp.data_ptr(), G.data_ptr(), d, C_, [1.0], mc.MODEL_SOFTMAX)
        med, mn = time_fn(lambda: mc.launch(job, scr), args.iters)
        one_pass = os.environ.get("EH_MC_TWOPASS") is None and (d <= 1024 or (d <= 4096 and os.environ.get("EH_MC_ONEPASS") == "cluster"))
        Xd = X[:, :d].float() if args.mc_bf16 else X[:, :d]
        W = torch.randn(d, C_, device="cuda") * 0.01
        def torch_ref():
            S = torch.softmax(Xd @ W, 1)
            S[torch.arange(rows, device="cuda"), y.long()] -= 1
            return Xd.t() @ S
        med2, _ = time_fn(torch_ref, max(3, args.iters // 2))
        byt = rows * X.stride(0) * (2 if args.mc_bf16 else 4)
        if args.mc_bf16:
            one_pass = True
        rec = {"kernel": "grad_mc(bf16 rows: tcgen05 fwd + tcgen05 bwd, one tile)" if args.mc_bf16 else "grad_mc(one pass: mma.sync fwd + tcgen05 bwd, %s)" % os.environ.get("EH_MC_ONEPASS", "auto") if one_pass else "grad_mc(tcgen05 fwd+bwd, softmax)",
               "rows": rows, "d": d, "C": C_, "ms": med, "ms_min": mn, "cs_forced": os.environ.get("EH_MC_CS"),
               "X_passes": 1 if one_pass else 2, "GBps_algorithmic(1 pass)": byt / med / 1e6,
               "frac_of_hbm_" + src + "_one_pass_roofline": byt / med / 1e6 / pk["hbm_gbs"],
               "torch_cublas_fp32_ms": med2, "speedup_vs_torch": med2 / med}
        print(json.dumps(rec)); out.append(rec)
        del X
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--cols", type=int, nargs="+", default=[4096, 1024, 2048, 8192, 512])
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="")
    ap.add_argument("--mc-bf16", action="store_true", help="bf16 feature rows for --only mc")
    ap.add_argument("--mc-shapes", nargs="*", default=[], help="rows,d,C triples for --only mc")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    pk, src = peaks()
    recs = []
    if args.only in ("", "grad"):
        recs += bench_grad(args, pk, src)
    if args.only in ("", "decoder"):
        recs += bench_decoder(args)
    if args.only in ("", "eval"):
        recs += bench_eval(args, pk, src)
    if args.only in ("", "ell"):
        recs += bench_ell(args, pk, src)
    if args.only in ("", "mc"):
        recs += bench_mc(args, pk, src)
    if args.out:
        json.dump(recs, open(args.out, "w"), indent=1)
