"""Command-line driver with the reference's 13-positional contract (main.py:20-94).

    python main.py n_procs n_rows n_cols input_dir is_real dataset is_coded n_stragglers \
                   partitions coded_ver num_collect add_delay update_rule [--key=value ...]

Launch one process per GPU with ``torchrun`` (``tools/launch.py`` wraps it); without a process
group the run uses the single visible GPU (all workers on it) or, without CUDA, the CPU oracle.
Unlike ``mpirun -np n_procs`` the number of processes is the number of GPUs, not ``n_procs``:
the ``n_procs - 1`` logical workers are spread over them (main.py:55-57 checked
``size == n_procs``; here any ``1 <= gpus <= n_procs - 1`` works).
"""
from __future__ import annotations

import sys
import time
from typing import Optional, Sequence

from . import results
from .config import RunConfig, parse_argv
from .data import load_dataset
from .engine import Engine
from .parallel.context import init_distributed

BANNERS = {   # start-of-run lines of the reference drivers
    "naive": "---- Starting Naive Iterations ----",
    "coded": "---- Starting Coded Iterations for {s} stragglers ----",
    "replication": "---- Starting Replication Iterations for {s} stragglers ----",
    "avoidstragg": "---- Starting AvoidStragg Iterations for {s} stragglers ----",
    "approx": "---- Starting Approx Coding Iterations for {s} stragglerssimulated delay {delay}-------",
    "partial_coded": "---- Starting Partial Coded Iterations for {s} stragglers ----",
    "partial_replication": "---- Starting Partial Replication Iterations for {s} stragglers ----",
}


def run(cfg: RunConfig, data=None, ctx=None, out=None) -> Optional[dict]:
    out = out or sys.stdout
    ctx = ctx or init_distributed("cpu" if cfg.gpus < 0 else None)
    if data is None and not cfg.synthetic:
        import numpy as np
        on_gpu = ctx.device.type == "cuda"
        # on GPUs every rank reads only the partition files of its own workers (the root one more at a time while it
        # evaluates), as float32; the CPU oracle works on all of them in float64 like the reference
        data = load_dataset(cfg.data_dir, cfg.n_partitions, bool(cfg.is_real), cfg.n_cols,
                            rows_per_part=cfg.n_rows // cfg.n_partitions, lazy=on_gpu,
                            dtype=np.float32 if on_gpu else float)
    eng = Engine(cfg, data, ctx=ctx)
    say = (lambda *a: print(*a, file=out)) if ctx.is_root and not cfg.quiet else (lambda *a: None)
    if ctx.remote_ranks:
        say(f"ranks {ctx.domain_ranks} share rank 0's node (peer memory); ranks {ctx.remote_ranks} are reached over NCCL")
    say(BANNERS[cfg.scheme].format(s=cfg.n_stragglers, delay=cfg.add_delay))
    if cfg.scheme.startswith("partial_"):
        say("Stragglers are allowed to be atmost %.2f times slower"
            % (cfg.partitions * 1.0 / max(1, cfg.partitions - cfg.n_stragglers - 1)))   # partial_coded.py:83
    t0 = time.time()
    res = eng.run()
    ctx.barrier()
    summary = None
    if ctx.is_root:
        for i in range(0, len(res.timeset), 10):
            say("\t >>> At Iteration %d" % (eng.first_round + i))                    # naive.py:92-93
        say("Total Time Elapsed: %.3f" % (time.time() - t0))                          # naive.py:156
        ev = eng.evaluate(res) if cfg.eval != "none" else None
        if ev is not None and not cfg.quiet:
            results.print_table(ev, res.timeset, file=out)
        results.write_results(cfg.results_dir, eng.spec.prefix, ev, res.timeset, res.worker_timeset,
                              betaset=res.betaset, compat=cfg.compat, scheme=cfg.scheme)
        if cfg.metrics_jsonl:
            eng.write_metrics_jsonl(cfg.metrics_jsonl, res, ev)
        say(">>> Done")
        summary = {"result": res, "evaluation": ev, "prefix": eng.spec.prefix, "results_dir": cfg.results_dir}
    ctx.barrier()
    eng.close()
    return summary


def main(argv: Optional[Sequence[str]] = None) -> int:
    cfg = parse_argv(list(sys.argv[1:] if argv is None else argv))
    run(cfg)
    return 0


if __name__ == "__main__":
    sys.exit(main())
