"""Gradient-code construction (host side, float64).

Covers the role of the reference's ``src/util.py:64-134``:

* :func:`cyclic_code` -- the random cyclic-MDS encode matrix ``B`` (``getB``, util.py:64-83).
  Row ``i`` is supported on ``{i..i+s} mod n``, ``B[i,i] = 1`` and every row lies in the null
  space of a random ``s x n`` matrix ``H`` whose rows sum to zero, hence any ``n-s`` rows of
  ``B`` span the all-ones vector.
* :func:`decode_table` -- one decode row per straggler pattern (``getA``, util.py:85-103),
  enumerated in ``itertools.combinations`` order of the *straggler* positions.
* :func:`mask_rank` -- combinadic rank of a *completed* mask (``calculate_indexA``,
  util.py:125-134); ``decode_table(B)[mask_rank(done)]`` is the decode row for ``done``.
* :func:`decode_weights` -- per-round least-squares decode used by the drivers
  (coded.py:147-149, partial_coded.py:192-194).
* FRC helpers (:func:`frc_group`) for the replication / approximate schemes
  (replication.py:46-52, approximate_coding.py:36-49).

Unlike the reference everything takes an explicit seed so runs are reproducible
(SURVEY Q8); ``seed=None`` draws from the global numpy RNG like the reference does.
"""
from __future__ import annotations

import itertools
from math import comb
from typing import Iterable, Optional, Sequence

import numpy as np

__all__ = [
    "cyclic_code",
    "cyclic_support",
    "parity_check",
    "decode_table",
    "mask_rank",
    "unrank_mask",
    "decode_weights",
    "frc_group",
    "frc_members",
    "binomial_table",
    "unique_rows",
    "compare_masks",
    "find_mask_row",
]


def cyclic_support(n: int, s: int) -> np.ndarray:
    """``[n, s+1]`` int array: partitions touched by worker ``i`` are ``(i + j) % n``."""
    return (np.arange(n)[:, None] + np.arange(s + 1)[None, :]) % n


def _draw_H(n: int, s: int, rng) -> np.ndarray:
    Ht = rng.normal(0.0, 1.0, (s, n - 1))
    return np.hstack([Ht, -Ht.sum(axis=1, keepdims=True)])


def parity_check(n: int, s: int, seed: Optional[int] = None) -> np.ndarray:
    """The random ``s x n`` matrix ``H`` with ``H @ 1 = 0`` that defines the code."""
    rng = np.random if seed is None else np.random.RandomState(seed)
    return _draw_H(n, s, rng)


def cyclic_code(n: int, s: int, seed: Optional[int] = None, H: Optional[np.ndarray] = None) -> np.ndarray:
    """Random cyclic gradient code ``B`` (``n x n``, ``s+1`` non-zeros per row).

    ``s == 0`` degenerates to the identity (uncoded).
    """
    if not 0 <= s < n:
        raise ValueError(f"need 0 <= s < n, got n={n} s={s}")
    B = np.zeros((n, n))
    if s == 0:
        np.fill_diagonal(B, 1.0)
        return B
    if H is None:
        H = parity_check(n, s, seed)
    supp = cyclic_support(n, s)
    for i in range(n):
        head, tail = supp[i, 0], supp[i, 1:]
        B[i, head] = 1.0
        B[i, tail] = -np.linalg.solve(H[:, tail], H[:, head])
    return B


def decode_weights(B: np.ndarray, done: Sequence[bool] | np.ndarray) -> np.ndarray:
    """Length-``n`` vector ``a`` with ``a[~done] = 0`` and ``a @ B = 1`` (least squares)."""
    done = np.asarray(done, dtype=bool)
    n = B.shape[0]
    a = np.zeros(n)
    idx = np.flatnonzero(done)
    if idx.size == 0:
        return a
    a[idx] = np.linalg.lstsq(B[idx, :].T, np.ones(B.shape[1]), rcond=None)[0]
    return a


def decode_table(B: np.ndarray, s: int) -> np.ndarray:
    """``C(n, s) x n`` table of decode rows; row order = ``combinations(range(n), s)`` of the
    *missing* workers, i.e. exactly the enumeration of the reference's ``getA``."""
    n = B.shape[0]
    A = np.zeros((comb(n, s), n))
    for r, missing in enumerate(itertools.combinations(range(n), s)):
        done = np.ones(n, dtype=bool)
        done[list(missing)] = False
        A[r] = decode_weights(B, done)
    return A


def mask_rank(done: Iterable[bool]) -> int:
    """Combinadic rank of a completed-mask; matches ``calculate_indexA`` (util.py:125-134)."""
    done = list(bool(x) for x in done)
    l = len(done)
    ctr = 0
    ind = 0
    for j in range(l - 1, -1, -1):
        if done[j]:
            ctr += 1
            ind += comb(l - 1 - j, ctr)
    return ind


def unrank_mask(rank: int, n: int, s: int) -> np.ndarray:
    """Inverse of :func:`mask_rank` restricted to masks with exactly ``n - s`` ones."""
    for r, missing in enumerate(itertools.combinations(range(n), s)):
        if r == rank:
            done = np.ones(n, dtype=bool)
            done[list(missing)] = False
            return done
    raise IndexError(rank)


def binomial_table(n: int) -> np.ndarray:
    """``(n+1) x (n+2)`` int64 table of ``C(a, b)`` (0 where ``b > a``) for the device-side rank."""
    T = np.zeros((n + 1, n + 2), dtype=np.int64)
    for a in range(n + 1):
        for b in range(min(a, n + 1) + 1):
            T[a, b] = comb(a, b)
    return T


def frc_group(w: int, s: int) -> int:
    """Fractional-repetition group of worker ``w`` (groups are ``s+1`` consecutive workers)."""
    return w // (s + 1)


def frc_members(n: int, s: int) -> np.ndarray:
    """``[n_groups, s+1]`` worker ids per group; requires ``(s+1) | n``."""
    if n % (s + 1):
        raise ValueError("n_workers must be a multiple of n_stragglers+1")
    return np.arange(n).reshape(n // (s + 1), s + 1)


# ---- small helpers the reference keeps next to getA (util.py:58-62,105-123); unused by its drivers
# but part of its utility surface, so equivalents are provided.

def unique_rows(a: np.ndarray) -> np.ndarray:
    """Distinct rows of a 2-D array, sorted lexicographically (``unique_rows``, util.py:58-62)."""
    return np.unique(np.ascontiguousarray(a), axis=0)


def compare_masks(a: Sequence[bool], b: Sequence[bool]) -> int:
    """Lexicographic order used by the reference's decode-table search (``compare``, util.py:105-111):
    +1 if the first differing position is set in ``a`` only, -1 if in ``b`` only, 0 if equal."""
    for x, y in zip(a, b):
        if x and not y:
            return 1
        if y and not x:
            return -1
    return 0


def find_mask_row(table: np.ndarray, mask: Sequence[bool]) -> int:
    """Row index of ``mask`` in a completed-mask table ordered like ``decode_table`` -- the job of
    ``binary_search_row_wise`` (util.py:113-123), done with the closed-form combinadic rank."""
    r = mask_rank(mask)
    if r >= len(table) or not np.array_equal(np.asarray(table[r], dtype=bool), np.asarray(mask, dtype=bool)):
        raise KeyError("mask not in table")
    return r
