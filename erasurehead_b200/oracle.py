"""Event-driven float64 golden model of every scheme (test oracle T0).

This restates the *mathematics* of the reference drivers (one function for all seven
schemes instead of ten clones): worker gradients (naive.py:137-139), encode folded into the
residual numerator (coded.py:92-95,184), the per-scheme stop rule and decode,
and the GD / non-textbook Nesterov update (naive.py:112-124).  Message arrival is simulated:
worker ``w``'s message of round ``i`` reaches the master at ``compute_time + delay[i, w]``
and the master consumes arrivals in time order exactly like its ``Waitany`` loop
(approximate_coding.py:144-158).

It is also the CPU fallback of the engine (``Engine(device='cpu')``) so that every host-side
path is testable without a GPU.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import codes
from .models import Model, get_model
from .schemes import (DECODE_FIRST_PER_GROUP, DECODE_LSTSQ, DECODE_SUM, MSG_CODED, MSG_PRIVATE,
                      SchemeSpec)

__all__ = ["RunResult", "gd_update", "agd_update", "worker_messages", "run_rounds", "initial_beta"]


@dataclass
class RunResult:
    betaset: np.ndarray                 # [R, d] (or [R, d, C]) parameters *after* update i (naive.py:127)
    timeset: np.ndarray                 # [R] seconds, start of broadcast -> end of update (naive.py:95,126)
    worker_timeset: np.ndarray          # [R, n] arrival time, -1 if not collected (coded.py:140,171-173)
    weights: np.ndarray                 # [R, n] decode weight applied to each worker's coded message
    n_collected: np.ndarray             # [R] coded messages consumed before the stop rule fired
    u: Optional[np.ndarray] = None      # AGD auxiliary state after the last round
    beta: Optional[np.ndarray] = None   # parameters after the last round
    wall_time: float = 0.0


def gd_update(beta, g, eta, alpha, grad_mult):
    """``beta <- (1 - 2 alpha eta) beta - grad_mult * g``  (naive.py:115)."""
    return (1.0 - 2.0 * alpha * eta) * beta - grad_mult * g


def agd_update(beta, u, g, eta, alpha, grad_mult, i):
    """The reference's Nesterov-style rule (naive.py:118-122); gradient taken at beta, not y."""
    theta = 2.0 / (i + 2.0)
    ytemp = (1.0 - theta) * beta + theta * u
    betatemp = ytemp - grad_mult * g - (2.0 * alpha * eta) * beta
    u_new = beta + (betatemp - beta) * (1.0 / theta)
    return betatemp, u_new


def initial_beta(spec: SchemeSpec, d: int, seed: Optional[int], n_out: int = 1) -> np.ndarray:
    """``randn`` for naive/replication/approx, zeros otherwise."""
    shape = (d,) if n_out == 1 else (d, n_out)
    if spec.beta0 == "zeros":
        return np.zeros(shape)
    rng = np.random if seed is None else np.random.RandomState(seed + 7919)
    return rng.standard_normal(shape) if seed is not None else np.random.randn(*shape)


def worker_messages(model: Model, spec: SchemeSpec, data, w: int, beta: np.ndarray):
    """Gradient message(s) of worker ``w``: dict msg-kind -> vector."""
    out = {}
    rpp = spec.rows_per_part
    for kind in (MSG_PRIVATE, MSG_CODED):
        segs = [sg for sg in spec.segments[w] if sg.msg == kind]
        if not segs:
            continue
        g = None
        for sg in segs:
            X = data.partition(sg.partition)
            y = data.labels(sg.partition)
            gi = model.gradient(X, y, np.full(rpp, sg.coeff), beta)
            g = gi if g is None else g + gi
        out[kind] = np.asarray(g)
    return out


def run_rounds(cfg, spec: SchemeSpec, data, delays: np.ndarray, beta0: Optional[np.ndarray] = None,
               u0: Optional[np.ndarray] = None, first_round: int = 0, rounds: Optional[int] = None,
               compute_time: float = 0.0, model: Optional[Model] = None) -> RunResult:
    """Simulate ``rounds`` rounds starting at global round index ``first_round``."""
    model = model or get_model(cfg.model_name, cfg.classes)
    R = cfg.rounds - first_round if rounds is None else rounds
    n, d = spec.n, cfg.n_cols
    eta = cfg.learning_rates()
    alpha, N = cfg.alpha, cfg.n_rows
    beta = initial_beta(spec, d, cfg.seed, model.n_out) if beta0 is None else np.array(beta0, dtype=np.float64)
    u = np.zeros_like(beta) if u0 is None else np.array(u0, dtype=np.float64)   # naive.py:63

    betaset = np.zeros((R,) + beta.shape)
    timeset = np.zeros(R)
    wts = np.zeros((R, n))
    weights = np.zeros((R, n))
    ncol = np.zeros(R, dtype=np.int64)

    for r in range(R):
        i = first_round + r
        msgs = [worker_messages(model, spec, data, w, beta) for w in range(n)]
        events = []
        for w in range(n):
            dl = delays[r, w]
            if MSG_PRIVATE in msgs[w]:
                events.append((compute_time + dl, 0, w, MSG_PRIVATE))
            events.append((2 * compute_time + dl if MSG_PRIVATE in msgs[w] else compute_time + dl,
                           1, w, MSG_CODED))
        events.sort()
        g = np.zeros_like(beta)
        done = np.zeros(n, dtype=bool)
        grp_cov = np.zeros(spec.n_groups, dtype=bool)
        first_of_group = np.zeros(n, dtype=bool)
        n_priv = n_coded = 0
        t_last = 0.0
        wts[r, :] = -1.0
        seen_t = np.full(n, -1.0)
        for (t, _, w, kind) in events:
            if spec.stop(n_priv, n_coded, int(grp_cov.sum())):
                break
            if not np.isfinite(t):
                raise RuntimeError(
                    f"round {i}: stop rule cannot be met, too many dead workers for scheme {spec.name}")
            t_last = t
            seen_t[w] = t
            if kind == MSG_PRIVATE:
                g += msgs[w][MSG_PRIVATE]
                n_priv += 1
                continue
            n_coded += 1
            done[w] = True
            gid = spec.group_of[w]
            if not grp_cov[gid]:
                grp_cov[gid] = True
                first_of_group[w] = True
        if not spec.stop(n_priv, n_coded, int(grp_cov.sum())):
            raise RuntimeError(f"round {i}: stop rule not met after all arrivals")
        if spec.decode == DECODE_SUM:
            a = done.astype(np.float64)
        elif spec.decode == DECODE_FIRST_PER_GROUP:
            a = first_of_group.astype(np.float64)
        else:
            a = codes.decode_weights(spec.B, done)                         # coded.py:147-149
        for w in np.flatnonzero(a):
            g += a[w] * msgs[w][MSG_CODED]
        weights[r] = a
        ncol[r] = n_coded
        wts[r, done] = seen_t[done]                                        # -1 for the uncollected

        gm = eta[i] * spec.lr_scale / N
        if cfg.update_rule == "GD":
            beta = gd_update(beta, g, eta[i], alpha, gm)
        else:
            beta, u = agd_update(beta, u, g, eta[i], alpha, gm, i)
        betaset[r] = beta
        timeset[r] = t_last
    return RunResult(betaset=betaset, timeset=timeset, worker_timeset=wts, weights=weights,
                     n_collected=ncol, u=u, beta=beta)
