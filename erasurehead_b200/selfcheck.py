"""``python -m erasurehead_b200.selfcheck``: is this installation sound?

Runs every scheme on a small synthetic problem through the engine of this machine (the CUDA kernels when a
GPU is present, launch with torchrun for several GPUs; otherwise the float64 CPU path) and compares iterates,
collected-worker sets (and, for information, round times) with the float64 oracle, the way ``tests/mp_engine_check.py`` does.
Prints one line per scheme and exits 1 on any disagreement.  Nothing is written to disk.
"""
from __future__ import annotations

import sys

import numpy as np

from . import Engine, RunConfig, build_scheme, oracle, triple_from_scheme
from .data import generate_gmm, generate_onehot
from .parallel.context import init_distributed

CASES = [
    # name, scheme, kwargs
    ("uncoded", "naive", dict(s=0)),
    ("cyclic MDS", "coded", dict(s=2, seed=3)),
    ("FRC replication", "replication", dict(s=1)),
    ("ignore stragglers", "avoidstragg", dict(s=2)),
    ("approximate (AGC)", "approx", dict(s=1, k=5)),
    ("partial coded", "partial_coded", dict(n=4, s=1, P=4, rows=12 * 160, seed=2)),
    ("partial replication", "partial_replication", dict(n=4, s=1, P=4, rows=12 * 160)),
    ("AGC on one-hot data", "approx", dict(s=1, k=6, rows=8 * 120, d=700, onehot=True)),
    ("AGC, linear regression", "approx", dict(s=1, k=6, model="linear", eta0=0.05)),
]


def spaced_delays(rounds: int, n: int, seed: int = 0, gap: float = 3e-3) -> np.ndarray:
    """A different arrival order every round, arrivals ``gap`` apart: robust against timing noise."""
    rng = np.random.RandomState(seed)
    T = np.zeros((rounds, n))
    for r in range(rounds):
        T[r, rng.permutation(n)] = gap * (1 + np.arange(n))
    return T


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    device = "cpu" if "--cpu" in argv else None
    ctx = init_distributed(device)
    say = print if ctx.is_root else (lambda *a, **k: None)
    say(f"erasurehead_b200 selfcheck: device {ctx.device}, {ctx.world} process(es)"
        + (f", ranks {ctx.remote_ranks} on other nodes" if ctx.remote_ranks else ""))
    bad = 0
    for name, scheme, kw in CASES:
        kw = dict(kw)
        n, s, P, k = kw.pop("n", 8), kw.pop("s", 1), kw.pop("P", 0), kw.pop("k", 0)
        rows, d, onehot = kw.pop("rows", 1920), kw.pop("d", 48), kw.pop("onehot", False)
        ic, pp, cv = triple_from_scheme(scheme, P)
        cfg = RunConfig(n + 1, rows, d, "", int(onehot), "x", ic, s, pp, cv, k, 1, "AGD", rounds=6,
                        gpus=(-1 if ctx.device.type == "cpu" else 0), quiet=True, **kw)
        task = "linear" if cfg.model_name == "linear" else "logistic"
        data = (generate_onehot(rows, d, 13, cfg.n_partitions, seed=2, task=task) if onehot
                else generate_gmm(rows, d, cfg.n_partitions, seed=1, task=task))
        dl = spaced_delays(cfg.rounds, n, seed=5)
        eng = Engine(cfg, data, ctx=ctx, delays=dl)
        res = eng.run()
        if ctx.is_root:
            ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0, model=eng.model)
            err = np.abs(res.betaset - ref.betaset).max() / max(1.0, np.abs(ref.betaset).max())
            same_set = np.array_equal(res.worker_timeset >= 0, ref.worker_timeset >= 0)
            t_err = np.abs(res.timeset - ref.timeset).max()
            # the round time is informational: the first rounds of a process include one-time module loads
            ok = err <= 1e-3 and same_set and np.array_equal(res.n_collected, ref.n_collected)
            say(f"  {name:26s} iterates {err:8.1e}  collected sets {'same' if same_set else 'DIFFER'}  "
                f"round time off by {t_err * 1e3:6.3f} ms  {'ok' if ok else 'MISMATCH'}")
            bad += 0 if ok else 1
        eng.close()
        ctx.barrier()
    bad = ctx.broadcast_object(bad)
    say("all schemes agree with the float64 oracle" if not bad else f"{bad} scheme(s) disagree")
    if ctx.world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
