#!/usr/bin/env python
"""Baseline B1 vs the product on the SAME data files and CLI arguments.

Runs (a) the unmodified reference (baseline/_ref via py2run.py, its own drivers, CPU, float64) over the
mpi4py subset and (b) this framework (`main.py`, GPU if present) for a small dense AGC / coded
configuration that the reference can load (its loader is np.loadtxt), and prints one JSON line
per arm with the mean per-iteration time from the `*_timeset.dat` files both write.

    python baseline/run_b1.py [--rows 7680 --cols 256 --workers 8 --delay 0]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=7680)
    ap.add_argument("--cols", type=int, default=256)
    ap.add_argument("--workers", type=int, default=8)
    ap.add_argument("--stragglers", type=int, default=1)
    ap.add_argument("--collect", type=int, default=6)
    ap.add_argument("--delay", type=int, default=0)
    ap.add_argument("--gpus", type=int, default=1)
    a = ap.parse_args()
    import numpy as np
    from erasurehead_b200.data import generate_gmm_to_disk
    from erasurehead_b200.utils import io
    tmp = tempfile.mkdtemp(prefix="eh_b1_")
    n_procs = a.workers + 1
    out = generate_gmm_to_disk(n_procs, a.rows, a.cols, tmp, a.stragglers, 0, 0, seed=1)
    args = [str(n_procs), str(a.rows), str(a.cols), tmp + "/", "0", "x", "1", str(a.stragglers), "0", "3",
            str(a.collect), str(a.delay), "AGD"]
    res = {}
    # (a) patched reference
    t0 = time.time()
    p = subprocess.run([sys.executable, os.path.join(HERE, "mpirun_shim.py"), "-np", str(n_procs), "main.py"] + args,
                       capture_output=True, text=True)
    wall_ref = time.time() - t0
    if p.returncode == 0:
        ts = io.load_data(out + "results/replication_acc_%d_timeset.dat" % a.stragglers)
        tl = io.load_data(out + "results/replication_acc_%d_training_loss.dat" % a.stragglers)
        res["reference_patched_cpu"] = {"ms_per_iter": float(np.mean(ts) * 1e3), "iters": len(ts), "wall_s": wall_ref,
                                        "final_train_loss": float(tl[-1])}
    else:
        res["reference_patched_cpu"] = {"error": (p.stdout + p.stderr)[-400:]}
    # (b) this framework on the same files
    t0 = time.time()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "main.py")] + args + ["--quiet"]
    p = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
    wall = time.time() - t0
    if p.returncode == 0:
        pre = out + "results/approx_acc_%d_%d" % (a.stragglers, a.collect)
        ts = np.load(pre + "_timeset.npy")
        tl = np.load(pre + "_training_loss.npy")
        res["erasurehead_b200"] = {"ms_per_iter": float(np.mean(ts) * 1e3), "iters": len(ts), "wall_s": wall,
                                   "final_train_loss": float(tl[-1])}
    else:
        res["erasurehead_b200"] = {"error": (p.stdout + p.stderr)[-400:]}
    res["config"] = vars(a)
    if "ms_per_iter" in res.get("reference_patched_cpu", {}) and "ms_per_iter" in res.get("erasurehead_b200", {}):
        res["speedup"] = res["reference_patched_cpu"]["ms_per_iter"] / res["erasurehead_b200"]["ms_per_iter"]
    print(json.dumps(res))


if __name__ == "__main__":
    main()
