#!/usr/bin/env python
"""Loss-versus-time curves of every scheme under the reference's straggler model (the experiment
behind the reference's claim, README.md:7).  Writes one CSV per scheme: round, cumulative time,
train loss, test loss, AUC.

    python examples/compare_schemes.py --out curves/ [--delay-scale 0.5] [--rounds 100]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np  # noqa: E402

from erasurehead_b200 import Engine, RunConfig, triple_from_scheme  # noqa: E402
from erasurehead_b200.data import generate_gmm  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="curves")
ap.add_argument("--workers", type=int, default=8)
ap.add_argument("--rows", type=int, default=8 * 16384)
ap.add_argument("--cols", type=int, default=512)
ap.add_argument("--rounds", type=int, default=100)
ap.add_argument("--delay-scale", type=float, default=0.01, help="mean message delay in seconds (reference: 0.5)")
a = ap.parse_args()

runs = [("naive", 0, 0), ("avoidstragg", 2, 0), ("coded", 2, 0), ("replication", 1, 0), ("approx", 1, a.workers // 2),
        ("approx", 1, 3 * a.workers // 4)]
data = generate_gmm(a.rows, a.cols, a.workers, seed=1)
os.makedirs(a.out, exist_ok=True)
for scheme, s, k in runs:
    ic, pp, cv = triple_from_scheme(scheme)
    cfg = RunConfig(a.workers + 1, a.rows, a.cols, "", 0, "x", ic, s, pp, cv, k, 1, "AGD", rounds=a.rounds,
                    delay_scale=a.delay_scale, seed=0, quiet=True)
    eng = Engine(cfg, data)
    eng.beta0[:] = 0
    if getattr(eng, "dec", None) is not None:
        eng.t_beta64.zero_()
    res = eng.run()
    ev = eng.evaluate(res)
    if eng.ctx.is_root:
        t = np.cumsum(res.timeset)
        name = f"{scheme}_s{s}" + (f"_k{k}" if scheme == "approx" else "")
        np.savetxt(os.path.join(a.out, name + ".csv"),
                   np.c_[np.arange(a.rounds), t, ev.training_loss, ev.testing_loss, ev.auc],
                   delimiter=",", header="round,time_s,train_loss,test_loss,auc", comments="")
        print(f"{name:24s} total {t[-1]:.3f} s, final train loss {ev.training_loss[-1]:.4f}")
    eng.close()
