#!/usr/bin/env python
"""Multi-class (softmax) regression on one-hot data of the Covertype SHAPE, generated on the GPUs.

The reference's Covertype pipeline one-hot encodes 54 categorical columns + a bias into 15 509 features
(arrange_real_data.py:157-178) and binarises the labels; here the same shape carries 7 classes.  No data
files are needed: `synthetic=True` with `is_real=1` generates the partitions on the device (csrc/synth.cu),
bit-identical on every GPU that holds a replica.  Without a GPU the float64 oracle runs a smaller problem.

    python examples/onehot_multiclass.py
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/onehot_multiclass.py
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch  # noqa: E402

from erasurehead_b200 import Engine, RunConfig, triple_from_scheme  # noqa: E402

on_gpu = torch.cuda.is_available()
n_workers, s = 30, 2                                     # the reference launcher's cluster: ten FRC groups of three
rows = 396112 if on_gpu else 30 * 400
is_coded, partitions, coded_ver = triple_from_scheme("replication")
cfg = RunConfig(n_procs=n_workers + 1, n_rows=rows, n_cols=15509 if on_gpu else 600, input_dir="", is_real=1,
                dataset="covtype", is_coded=is_coded, n_stragglers=s, partitions=partitions, coded_ver=coded_ver,
                num_collect=0, add_delay=1, update_rule="AGD", rounds=30, delay_scale=5e-3, seed=1, quiet=True,
                synthetic=True, synthetic_nnz=55 if on_gpu else 12, model="softmax", classes=7, init="zeros", eta0=0.3)
eng = Engine(cfg, None)
res = eng.run()
ev = eng.evaluate(res)
if eng.ctx.is_root:
    print(f"device: {eng.ctx.device}, GPUs: {eng.ctx.world}, scheme: {eng.spec.name}, rows x features x classes: "
          f"{cfg.n_rows} x {cfg.n_cols} x 7")
    print(f"mean round time {res.timeset.mean() * 1e3:.2f} ms (Exp(5 ms) message delays), "
          f"collected {res.n_collected.mean():.1f} of {n_workers} messages per round")
    print(f"cross-entropy {ev.training_loss[0]:.3f} -> {ev.training_loss[-1]:.3f}, test accuracy {ev.accuracy[-1]:.3f}")
eng.close()
