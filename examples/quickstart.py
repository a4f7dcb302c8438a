#!/usr/bin/env python
"""Quick start: approximate gradient coding (ErasureHead) on synthetic data through the Python API.

    python examples/quickstart.py                      # one GPU (all 8 workers on it) or CPU oracle
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/quickstart.py
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

from erasurehead_b200 import Engine, RunConfig, triple_from_scheme  # noqa: E402
from erasurehead_b200.data import generate_gmm  # noqa: E402

n_workers, s, n_collect = 8, 1, 6
is_coded, partitions, coded_ver = triple_from_scheme("approx")
cfg = RunConfig(n_procs=n_workers + 1, n_rows=8 * 8192, n_cols=256, input_dir="", is_real=0, dataset="x",
                is_coded=is_coded, n_stragglers=s, partitions=partitions, coded_ver=coded_ver,
                num_collect=n_collect, add_delay=1, update_rule="AGD",
                rounds=50, delay_scale=2e-3, seed=0, quiet=True)
data = generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=1)      # or load_dataset(cfg.data_dir, ...)
eng = Engine(cfg, data)
res = eng.run()                       # RunResult on rank 0: betaset, timeset, worker_timeset, weights
ev = eng.evaluate(res)
if eng.ctx.is_root:
    print(f"device: {eng.ctx.device}, GPUs: {eng.ctx.world}, scheme: {eng.spec.name}")
    print(f"mean round time {res.timeset.mean() * 1e3:.2f} ms, collected {res.n_collected.mean():.1f} of {n_workers} messages/round")
    print(f"train loss {ev.training_loss[0]:.3f} -> {ev.training_loss[-1]:.3f}, test AUC {ev.auc[-1]:.3f}")
eng.close()
