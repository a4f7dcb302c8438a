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

from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import codes
from .models import Model, get_model
from .schemes import DECODE_FIRST_PER_GROUP, DECODE_SUM, MSG_CODED, MSG_PRIVATE, SchemeSpec

__all__ = ["RunResult", "gd_update", "agd_update", "worker_messages", "plan_round", "run_rounds", "initial_beta"]


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
    done: Optional[np.ndarray] = None   # [R, n] seconds from round start to completion of w's coded message
                                        # (before the injected delay); GPU engine only, -1 if never seen


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
    if seed is None:
        return np.random.randn(*shape)               # unseeded like the reference (naive.py:23)
    return np.random.RandomState(seed + 7919).standard_normal(shape)


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


def carryover_delays(spec: SchemeSpec, delays: np.ndarray) -> np.ndarray:
    """Delay table of a run whose master never waits for stragglers between rounds.

    In the reference's coded driver (coded.py:137-166) the master moves on as soon as n - s messages are in,
    while a straggler is still inside its ``time.sleep``: that worker picks up the next parameters only when it
    wakes, so its next message arrives late by what was left of the old delay, on top of the new one (its
    replication / approximate drivers ``Waitall`` on every straggler after each round, replication.py:179-180,
    so there every round starts fresh -- SURVEY Q5).  This replays that backlog from the delay table and the
    stop rule alone:  ``eff[i] = backlog + delays[i]``,  ``backlog = max(0, eff[i] - t_stop[i])``,  and returns the
    effective table, which the oracle and the device decoder then consume like any other.  The compute and
    transport time of a round (milliseconds against delays of half a second) is not part of the replay."""
    delays = np.asarray(delays, dtype=np.float64)
    eff = np.empty_like(delays)
    backlog = np.zeros(delays.shape[1])
    for i in range(delays.shape[0]):
        eff[i] = backlog + delays[i]
        t_stop = plan_round(spec, eff[i], round_index=i)[2]
        backlog = np.where(np.isfinite(eff[i]), np.maximum(0.0, eff[i] - t_stop), np.inf)
    return eff


def plan_round(spec: SchemeSpec, delay_row: np.ndarray, compute_time: float = 0.0, round_index: int = 0):
    """Replay the master's Waitany loop for one round given only the message delays.

    Returns ``(a, priv, t_stop, n_coded, worker_time)``: decode weights of the coded messages,
    0/1 weights of the private messages, the time the stop rule fired, the number of coded
    messages consumed and the per-worker arrival time (-1 if the coded message was not
    collected).  Pure function of the scheme and the delay row -- shared by the oracle and the
    NCCL baseline (which has no device-side arrival detection)."""
    n = spec.n
    has_priv = spec.n_msgs > 1
    events = []
    for w in range(n):
        dl = delay_row[w]
        if has_priv:
            events.append((compute_time + dl, 0, w, MSG_PRIVATE))
        events.append(((2 * compute_time if has_priv else compute_time) + dl, 1, w, MSG_CODED))
    events.sort()
    done = np.zeros(n, dtype=bool)
    priv = np.zeros(n)
    grp_cov = np.zeros(spec.n_groups, dtype=bool)
    first_of_group = np.zeros(n, dtype=bool)
    n_priv = n_coded = 0
    t_last = 0.0
    seen_t = np.full(n, -1.0)
    order = []
    for (t, _, w, kind) in events:
        if spec.stop(n_priv, n_coded, int(grp_cov.sum())):
            break
        if not np.isfinite(t):
            raise RuntimeError(
                f"round {round_index}: stop rule cannot be met, too many dead workers for scheme {spec.name}")
        t_last = t
        seen_t[w] = t
        if kind == MSG_PRIVATE:
            priv[w] = 1.0
            n_priv += 1
            continue
        n_coded += 1
        done[w] = True
        order.append(w)
        gid = spec.group_of[w]
        if not grp_cov[gid]:
            grp_cov[gid] = True
            first_of_group[w] = True
    if not spec.stop(n_priv, n_coded, int(grp_cov.sum())):
        raise RuntimeError(f"round {round_index}: stop rule not met after all arrivals")
    if spec.decode == DECODE_SUM:
        a = done.astype(np.float64)
    elif spec.decode == DECODE_FIRST_PER_GROUP:
        a = first_of_group.astype(np.float64)
    else:
        # coded.py:147-149.  partial_coded can collect more than n-s coded messages while it waits for the
        # private parts (partial_coded.py:174); the reference then takes the minimum-norm least-squares row
        # over all of them, this framework (device and oracle alike) the exact row of the FIRST n-s arrivals.
        # Both satisfy a.B[done] = 1, so the decoded gradient is the same; only the weights differ.
        sel = np.zeros(n, dtype=bool)
        sel[order[:max(1, spec.n - spec.s)]] = True
        a = codes.decode_weights(spec.B, sel)
    wt = np.where(done, seen_t, -1.0)                                      # -1 for the uncollected
    return a, priv, t_last, n_coded, wt


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
        a, priv, t_last, n_coded, wt = plan_round(spec, delays[r], compute_time, i)
        g = np.zeros_like(beta)
        for w in np.flatnonzero(priv):
            g += msgs[w][MSG_PRIVATE]
        for w in np.flatnonzero(a):
            g += a[w] * msgs[w][MSG_CODED]
        weights[r] = a
        ncol[r] = n_coded
        wts[r] = wt

        gm = eta[i] * spec.lr_scale / N
        if cfg.update_rule == "GD":
            beta = gd_update(beta, g, eta[i], alpha, gm)
        else:
            beta, u = agd_update(beta, u, g, eta[i], alpha, gm, i)
        betaset[r] = beta
        timeset[r] = t_last
    return RunResult(betaset=betaset, timeset=timeset, worker_timeset=wts, weights=weights,
                     n_collected=ncol, u=u, beta=beta)
