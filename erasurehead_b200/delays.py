"""Straggler-delay model (fault injection).

``exp`` reproduces the reference bit-for-bit: every worker executes
``np.random.seed(i); np.random.exponential(0.5, n_workers)[rank-1]`` in round ``i``
(naive.py:141-148, coded.py:187-194, replication.py:196-203, approximate_coding.py:198-205),
i.e. the legacy MT19937 stream re-seeded with the round index.  The table is produced on the
host with numpy and uploaded; the device only consumes it.

``fixed`` is the reference's commented-out variant (naive.py:143-145): ``s`` workers chosen
per round by ``np.random.choice`` sleep for a constant time.

``dead`` marks a fixed set of workers as permanently failed (delay = +inf) on top of ``exp``
(or of zero delays when ``scale == 0``); coded / approximate schemes must keep going.

Delays are a property of the *message*: a worker's gradient becomes visible to the decoder
``delay`` seconds after the worker finished computing it.  The worker itself does not sleep,
so a straggler is genuinely skipped: it starts the next round as soon as the new parameters
arrive (the reference could only cancel the stale send, README.md:122).
"""
from __future__ import annotations

from typing import Iterable, Optional

import numpy as np

__all__ = ["delay_table", "expected_kth_wait"]


def delay_table(rounds: int, n_workers: int, mode: str = "exp", scale: float = 0.5,
                n_stragglers: int = 0, fixed_sleep: float = 0.5,
                dead: Optional[Iterable[int]] = None, enabled: bool = True,
                first_round: int = 0) -> np.ndarray:
    """``[rounds, n_workers]`` float64 seconds.  ``first_round`` offsets the per-round seed
    (used when resuming from a checkpoint)."""
    T = np.zeros((rounds, n_workers))
    if not enabled or mode == "none":
        mode = "none"
    for r in range(rounds):
        i = first_round + r
        if mode in ("exp", "dead") and scale > 0:
            T[r] = np.random.RandomState(i).exponential(scale, n_workers)
        elif mode == "fixed" and n_stragglers > 0:
            rs = np.random.RandomState(i)
            idx = rs.choice(np.arange(1, n_workers + 1), n_stragglers, replace=False)
            T[r, idx - 1] = fixed_sleep
    if dead is not None:
        for w in dead:
            T[:, int(w)] = np.inf
    return T


def expected_kth_wait(n: int, k: int, scale: float = 0.5) -> float:
    """E[k-th order statistic of n iid Exp(scale)] = scale * (H_n - H_{n-k})."""
    H = lambda m: sum(1.0 / j for j in range(1, m + 1))
    return scale * (H(n) - H(n - k))
