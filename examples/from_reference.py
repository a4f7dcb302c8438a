#!/usr/bin/env python
"""A script written against the reference, switched over by changing its imports.

The reference's main.py builds ``params`` and calls one driver (main.py:48-51, 62-92).  The same
calls work here through ``erasurehead_b200.compat`` (same function names and argument lists):

    python examples/from_reference.py                                   # one GPU or, without one, the CPU model
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/from_reference.py
"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np  # noqa: E402

# was:  from approximate_coding import approx_logistic_regression;  from coded import coded_logistic_regression;  from util import *
from erasurehead_b200.compat import approx_logistic_regression, coded_logistic_regression, getB, load_data  # noqa: E402
from erasurehead_b200.data import generate_gmm_to_disk  # noqa: E402  (python -m erasurehead_b200.data.generate = generate_data.py)

n_procs, n_rows, n_cols, n_stragglers, num_collect = 9, 8 * 960, 64, 1, 6
rank = int(os.environ.get("RANK", "0"))
root = os.path.join(tempfile.gettempdir(), "eh_from_reference")
data_dir = f"{root}/artificial-data/{n_rows}x{n_cols}/{n_procs - 1}/"
if rank == 0 and not os.path.exists(data_dir + "done"):
    generate_gmm_to_disk(n_procs, n_rows, n_cols, root, n_stragglers, 0, 0, seed=0)
    open(data_dir + "done", "w").close()
while not os.path.exists(data_dir + "done"):          # the other ranks wait for the files
    import time
    time.sleep(0.2)

# ---- the reference's "modifiable parameters" block (main.py:31-51)
num_itrs = 40
alpha = 1.0 / n_rows
learning_rate_schedule = 1.0 * np.ones(num_itrs)
params = [num_itrs, alpha, learning_rate_schedule]

print_ = print if rank == 0 else (lambda *a, **k: None)
print_("B of the cyclic code, first row:", np.round(getB(n_procs - 1, n_stragglers)[0], 3))
out = approx_logistic_regression(n_procs, n_rows, n_cols, data_dir, n_stragglers, 0, params, num_collect, 1, "AGD",
                                 delay_scale=2e-3, quiet=True)
if out:
    print_("approximate coding: train loss", np.round(load_data(data_dir + "results/replication_acc_1_training_loss.dat")[[0, -1]], 3),
           "mean round time %.2f ms" % (out["result"].timeset.mean() * 1e3))
out = coded_logistic_regression(n_procs, n_rows, n_cols, data_dir, n_stragglers, 0, params, 1, "AGD", delay_scale=2e-3, quiet=True)
if out:
    print_("exact coding      : train loss", np.round(out["evaluation"].training_loss[[0, -1]], 3),
           "mean round time %.2f ms" % (out["result"].timeset.mean() * 1e3))
