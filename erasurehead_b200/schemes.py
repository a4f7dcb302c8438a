"""Scheme descriptors: every straggler-mitigation scheme of the reference as one table row.

The reference implements each scheme as a separate ~250-line driver (``src/naive.py``,
``coded.py``, ``replication.py``, ``avoidstragg.py``, ``partial_coded.py``,
``partial_replication.py``, ``approximate_coding.py``).  Here a scheme is *data*: which
partitions each worker streams, with which encode coefficient, into which message, when the
decoder may stop waiting, and how it combines what arrived.  One engine (``engine.py``), one
oracle (``oracle.py``) and one pair of device kernels consume the descriptor.

Message kinds: ``MSG_CODED`` (every scheme) and ``MSG_PRIVATE`` (partial schemes only; the
reference's tag ``2*rounds+i`` message, partial_coded.py:139-142).

Stop rule (generic form, evaluated after every arrival)::

    private_arrived >= need_private  AND
    (coded_arrived >= need_count  OR  groups_covered >= need_groups)

=====================  ============  ==========  ===========  ==================
scheme                 need_private  need_count  need_groups  decode of coded msgs
=====================  ============  ==========  ===========  ==================
naive                  0             n           inf          sum of arrived
coded (cyclic MDS)     0             n-s         inf          least-squares row
replication (FRC)      0             inf         n/(s+1)      first per group
avoidstragg            0             n-s         inf          sum of arrived
approx (AGC)           0             k           n/(s+1)      first per covered group
partial_coded          n             n-s         inf          least-squares row
partial_replication    n             inf         n/(s+1)      first per group
=====================  ============  ==========  ===========  ==================

(reference stop rules: naive.py:103, coded.py:137, replication.py:143, avoidstragg.py:106,
approximate_coding.py:144, partial_coded.py:174, partial_replication.py:166).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import codes
from .config import RunConfig

MSG_CODED = 0
MSG_PRIVATE = 1
INF = 1 << 30

DECODE_SUM = 0
DECODE_FIRST_PER_GROUP = 1
DECODE_LSTSQ = 2
DECODE_NAMES = {DECODE_SUM: "sum", DECODE_FIRST_PER_GROUP: "first_per_group", DECODE_LSTSQ: "lstsq"}


@dataclass
class Segment:
    """One partition streamed by a worker."""
    partition: int      # 0-based partition id (file ``partition+1``.dat/.npz)
    coeff: float        # encode weight folded into the residual numerator (coded.py:92-95)
    msg: int            # MSG_CODED or MSG_PRIVATE


@dataclass
class SchemeSpec:
    name: str
    n: int                       # workers
    s: int                       # designed straggler tolerance
    n_partitions: int            # partition files on disk
    rows_per_part: int
    segments: List[List[Segment]]
    n_msgs: int                  # 1, or 2 for the partial schemes
    group_of: np.ndarray         # [n] FRC group id of each worker's coded message
    n_groups: int
    need_private: int
    need_count: int
    need_groups: int
    decode: int
    lr_scale: float              # step multiplier on eta/N (avoidstragg: n/(n-s), avoidstragg.py:116)
    beta0: str                   # 'zeros' | 'randn'
    prefix: str                  # result-file prefix
    B: Optional[np.ndarray] = None
    num_collect: int = 0
    partitions_per_worker: int = 0

    def rows_of(self, w: int, msg: Optional[int] = None) -> int:
        return sum(self.rows_per_part for sg in self.segments[w] if msg is None or sg.msg == msg)

    def stop(self, n_private: int, n_coded: int, n_groups_cov: int) -> bool:
        return n_private >= self.need_private and (
            n_coded >= self.need_count or n_groups_cov >= self.need_groups)

    def redundancy(self) -> float:
        """Rows processed per round over dataset rows (s+1 for the fully coded schemes)."""
        tot = sum(len(sg) for sg in self.segments)
        return tot / float(self.n_partitions)


def build_scheme(cfg: RunConfig, B: Optional[np.ndarray] = None) -> SchemeSpec:
    name, n, s = cfg.scheme, cfg.n_workers, cfg.n_stragglers
    P = cfg.partitions
    if s < 0 or s >= n:
        raise ValueError(f"need 0 <= n_stragglers < n_workers (n={n}, s={s})")
    n_parts = cfg.n_partitions
    rows_per_part = cfg.n_rows // n_parts            # Q11: remainder rows are dropped
    if rows_per_part <= 0:
        raise ValueError("fewer rows than partitions")
    needs_frc = name in ("replication", "approx", "partial_replication")
    if needs_frc and n % (s + 1):
        # replication.py:24-26, approximate_coding.py:25-27, partial_replication.py:20-24
        raise ValueError("n_workers must be multiple of n_stragglers+1!")
    needs_B = name in ("coded", "partial_coded")
    if needs_B and B is None:
        B = codes.cyclic_code(n, s, seed=cfg.seed)
    group_of = np.arange(n)
    n_groups = n
    if needs_frc:
        group_of = np.arange(n) // (s + 1)
        n_groups = n // (s + 1)

    segs: List[List[Segment]] = []
    n_msgs = 1
    if name in ("naive", "avoidstragg"):
        segs = [[Segment(w, 1.0, MSG_CODED)] for w in range(n)]            # naive.py:29-36
    elif name == "coded":
        segs = [[Segment((w + i) % n, float(B[w, (w + i) % n]), MSG_CODED)  # coded.py:33-35,95
                 for i in range(s + 1)] for w in range(n)]
    elif name in ("replication", "approx"):
        for w in range(n):                                                  # replication.py:46-52
            a, b = divmod(w, s + 1)
            segs.append([Segment((s + 1) * a + (b + i) % (s + 1), 1.0, MSG_CODED)
                         for i in range(s + 1)])
    elif name in ("partial_coded", "partial_replication"):
        if P <= s + 0:
            raise ValueError("partitions must exceed n_stragglers")
        n_sep = P - s - 1
        n_msgs = 2
        for w in range(n):
            sg = [Segment(i + n_sep * w, 1.0, MSG_PRIVATE) for i in range(n_sep)]   # partial_coded.py:33-36
            if name == "partial_coded":
                sg += [Segment(n_sep * n + (w + j) % n, float(B[w, (w + j) % n]), MSG_CODED)
                       for j in range(s + 1)]                               # partial_coded.py:38-43,125
            else:
                a = w // (s + 1)
                sg += [Segment(n_sep * n + a * (s + 1) + j, 1.0, MSG_CODED)
                       for j in range(s + 1)]                               # partial_replication.py:44-50
            segs.append(sg)
    else:
        raise ValueError(name)

    need_private = n if n_msgs == 2 and (P - s - 1) > 0 else 0
    if n_msgs == 2 and need_private == 0:
        n_msgs = 1                         # P = s+1: no private part at all
    need_count, need_groups = INF, INF
    decode = DECODE_SUM
    lr_scale = 1.0
    if name == "naive":
        need_count = n
    elif name in ("coded", "partial_coded"):
        need_count, decode = n - s, DECODE_LSTSQ
    elif name in ("replication", "partial_replication"):
        need_groups, decode = n_groups, DECODE_FIRST_PER_GROUP
    elif name == "avoidstragg":
        need_count = n - s
        lr_scale = n / float(n - s)                                         # avoidstragg.py:116
    elif name == "approx":
        need_count, need_groups, decode = max(1, cfg.num_collect), n_groups, DECODE_FIRST_PER_GROUP

    beta0 = "randn" if name in ("naive", "replication", "approx") else "zeros"   # Q8
    if getattr(cfg, "init", "scheme") != "scheme":
        if cfg.init not in ("zeros", "randn"):
            raise ValueError("init must be scheme, zeros or randn")
        beta0 = cfg.init

    prefix = {
        "naive": "naive_acc",
        "coded": f"coded_acc_{s}",
        "replication": f"replication_acc_{s}",
        "avoidstragg": f"avoidstragg_acc_{s}",
        "approx": (f"replication_acc_{s}" if cfg.compat else f"approx_acc_{s}_{cfg.num_collect}"),  # Q2
        "partial_coded": f"partialcoded_{s}_{P}",
        "partial_replication": f"partialreplication_{s}_{P}",
    }[name]

    return SchemeSpec(
        name=name, n=n, s=s, n_partitions=n_parts, rows_per_part=rows_per_part, segments=segs,
        n_msgs=n_msgs, group_of=group_of.astype(np.int32), n_groups=n_groups,
        need_private=need_private, need_count=need_count, need_groups=need_groups, decode=decode,
        lr_scale=lr_scale, beta0=beta0, prefix=prefix, B=B, num_collect=cfg.num_collect,
        partitions_per_worker=len(segs[0]))
