"""Code construction: properties of getB / getA / calculate_indexA (util.py:64-134)."""
import itertools

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from erasurehead_b200 import codes


@pytest.mark.parametrize("n,s", [(4, 1), (8, 1), (8, 2), (8, 3), (12, 3), (30, 3)])
def test_cyclic_code_structure(n, s):
    H = codes.parity_check(n, s, seed=3)
    B = codes.cyclic_code(n, s, H=H)
    assert np.allclose(H.sum(axis=1), 0)
    for i in range(n):
        supp = {(i + j) % n for j in range(s + 1)}
        assert set(np.flatnonzero(B[i])) <= supp
        assert B[i, i] == 1.0
    assert np.abs(H @ B.T).max() < 1e-8 * max(1.0, np.abs(B).max())


@pytest.mark.parametrize("n,s", [(4, 1), (8, 2), (10, 3)])
def test_any_n_minus_s_rows_decode(n, s):
    B = codes.cyclic_code(n, s, seed=11)
    for missing in itertools.combinations(range(n), s):
        done = np.ones(n, bool)
        done[list(missing)] = False
        a = codes.decode_weights(B, done)
        assert np.all(a[~done] == 0)
        assert np.abs(a @ B - 1).max() < 1e-6


def test_decode_table_matches_rank():
    n, s = 8, 2
    B = codes.cyclic_code(n, s, seed=5)
    A = codes.decode_table(B, s)
    assert A.shape == (28, n)
    for r, missing in enumerate(itertools.combinations(range(n), s)):
        done = np.ones(n, bool)
        done[list(missing)] = False
        assert codes.mask_rank(done) == r
        assert np.array_equal(codes.unrank_mask(r, n, s), done)
        assert np.abs(A[r] @ B - 1).max() < 1e-6


def test_s0_is_identity():
    assert np.array_equal(codes.cyclic_code(5, 0), np.eye(5))


@settings(max_examples=40, deadline=None)
@given(st.integers(2, 12).flatmap(lambda n: st.tuples(st.just(n), st.integers(0, n - 1))), st.integers(0, 10**6))
def test_mask_rank_is_a_bijection(ns, seed):
    n, s = ns
    rng = np.random.RandomState(seed)
    miss = rng.choice(n, s, replace=False)
    done = np.ones(n, bool)
    done[miss] = False
    r = codes.mask_rank(done)
    assert 0 <= r < codes.binomial_table(n)[n, s]
    assert np.array_equal(codes.unrank_mask(r, n, s), done)


def test_binomial_table():
    T = codes.binomial_table(10)
    assert T[10, 3] == 120 and T[5, 0] == 1 and T[3, 5] == 0


def test_frc_members():
    m = codes.frc_members(8, 1)
    assert m.shape == (4, 2) and codes.frc_group(5, 1) == 2
    with pytest.raises(ValueError):
        codes.frc_members(7, 1)


def test_reference_utility_helpers():
    a = np.array([[1, 0], [0, 1], [1, 0]])
    assert codes.unique_rows(a).tolist() == [[0, 1], [1, 0]]
    assert codes.compare_masks([1, 0, 1], [1, 1, 0]) == -1 and codes.compare_masks([1, 1], [1, 1]) == 0
    n, s = 6, 2
    table = np.array([codes.unrank_mask(r, n, s) for r in range(15)])
    for r in range(15):
        assert codes.find_mask_row(table, table[r]) == r
    from erasurehead_b200.data.generate import generate_random_binvec, generate_random_matrix_normal
    v = generate_random_binvec(50, np.random.RandomState(0))
    assert set(np.unique(v)) <= {-1, 1}
    m = generate_random_matrix_normal(np.ones(4), -np.ones(4), 100, 4, np.random.RandomState(0))
    assert m.shape == (100, 4)
