"""erasurehead_b200.compat: the reference's driver functions and util.py helpers under their own names."""
import os

import numpy as np
import pytest
from scipy.special import binom

from erasurehead_b200 import compat
from erasurehead_b200.data import generate_gmm_to_disk, generate_onehot


def _params(rounds=6, n_rows=480, eta=1.0):
    return [rounds, 1.0 / n_rows, eta * np.ones(rounds)]            # main.py:48-51


@pytest.mark.parametrize("name,extra,prefix", [
    ("naive_logistic_regression", (0, "GD"), "naive_acc"),
    ("coded_logistic_regression", (0, "AGD"), "coded_acc_1"),
    ("replication_logistic_regression", (1, "AGD"), "replication_acc_1"),
    ("avoidstragg_logistic_regression", (), "avoidstragg_acc_1"),
    ("approx_logistic_regression", (3, 1, "AGD"), "replication_acc_1"),            # Q2: AGC writes replication's names
])
def test_driver_functions_run_with_the_reference_argument_lists(tmp_path, name, extra, prefix):
    n_procs, rows, cols, s = 5, 480, 10, 1
    d = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, 0, 0, seed=3)
    out = getattr(compat, name)(n_procs, rows, cols, d, s, 0, _params(), *extra, gpus=-1, quiet=True, seed=1)
    assert out["result"].betaset.shape == (6, cols)
    ev = out["evaluation"]
    assert ev.training_loss[-1] < ev.training_loss[0]
    for q in ("training_loss", "testing_loss", "auc", "timeset", "worker_timeset"):
        assert os.path.exists(f"{d}results/{prefix}_{q}.dat"), q


@pytest.mark.parametrize("name", ["partial_coded_logistic_regression", "partial_replication_logistic_regression"])
def test_partial_driver_functions(tmp_path, name):
    n_procs, rows, cols, s, P = 5, 480, 10, 1, 3
    d = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, P, 1, seed=3)
    out = getattr(compat, name)(n_procs, rows, cols, d, s, P, 0, _params(), gpus=-1, quiet=True)
    assert out["result"].betaset.shape == (6, cols) and np.isfinite(out["evaluation"].training_loss).all()


@pytest.mark.parametrize("name,extra", [("naive_linear_regression", (0, "GD")), ("replication_linear_regression", (0, "GD")),
                                        ("approx_linear_regression", (4, 0, "GD"))])
def test_linear_driver_functions_on_one_hot_files(tmp_path, name, extra):
    n_procs, rows, cols, nnz = 5, 480, 40, 6
    d = str(tmp_path) + "/kc_house_data/4/"
    generate_onehot(rows, cols, nnz, n_procs - 1, seed=2, task="linear").save(d)
    out = getattr(compat, name)(n_procs, rows, cols, d, 1, 1, _params(eta=0.02), *extra, gpus=-1, quiet=True, seed=1)
    ev = out["evaluation"]
    assert ev.training_loss[-1] < ev.training_loss[0]               # mean squared error goes down


def test_every_reference_driver_has_a_counterpart():
    assert len(compat.DRIVERS) == 10 and all(callable(getattr(compat, n)) for n in compat.DRIVERS)


def test_util_helpers_under_the_reference_names(tmp_path):
    n, s = 6, 2
    B = compat.getB(n, s)
    assert B.shape == (n, n) and np.allclose(np.diag(B), 1.0)
    A = compat.getA(B, n, s)
    assert A.shape == (int(binom(n, s)), n) and np.allclose(A @ B, 1.0)
    # calculate_indexA's closed form (util.py:125-134) and the row order of getA agree
    for r in (0, 3, A.shape[0] - 1):
        mask = A[r] != 0
        l, ctr, ind = len(mask), 0, 0
        for j in range(l - 1, -1, -1):
            if mask[j]:
                ctr += 1
                ind += binom(l - 1 - j, ctr)
        assert compat.calculate_indexA(mask) == int(ind)
    assert compat.compare([1, 0, 1], [1, 1, 0]) == -1 and compat.compare([1, 1], [1, 1]) == 0
    y = np.array([1.0, -1.0, 1.0])
    p = np.array([0.3, -0.2, -0.1])
    assert np.isclose(compat.calculate_loss(y, p, 3), np.mean(np.log(1 + np.exp(-y * p))))
    assert np.isclose(compat.calculate_mse(y, p, 3), np.mean((y - p) ** 2))
    M = np.arange(12.0).reshape(3, 4)
    compat.save_matrix(M, str(tmp_path / "m.dat"))
    assert np.allclose(compat.load_data(str(tmp_path / "m.dat")), M)
    assert compat.unique_rows(np.array([[1, 2], [1, 2], [0, 5]])).tolist() == [[0, 5], [1, 2]]
    X = compat.interactionTermsAmazon(np.array([[1, 2, 3], [1, 2, 4]]), degree=2)
    assert X.shape[0] == 2 and X.shape[1] >= 1


def _reference_util():
    """The reference's util.py compiled with Python-2 division (baseline/py2run.py), as a namespace."""
    import sys
    root = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
    path = next((p for p in (os.path.join(root, "baseline", "_ref", "src", "util.py"), "/root/reference/src/util.py")
                 if os.path.exists(p)), None)
    if path is None:
        pytest.skip("reference tree not available")
    sys.path.insert(0, os.path.join(root, "baseline"))
    import py2run
    ns = {"__name__": "ref_util", "__builtins__": __builtins__}
    exec(py2run.compile_py2(open(path, "rb").read(), path), ns)
    return ns


def test_helpers_agree_with_the_reference_util_module(tmp_path):
    """Same inputs through the reference's own util.py functions and through compat's."""
    ref = _reference_util()
    rng = np.random.RandomState(0)
    for n, s in ((4, 1), (6, 2), (7, 3)):
        B = ref["getB"](n, s)                                              # the reference's own code matrix
        assert np.allclose(compat.getA(B, n, s), ref["getA"](B, n, s), atol=1e-9)
        A = compat.getA(B, n, s)
        for r in range(A.shape[0]):
            mask = A[r] != 0
            assert compat.calculate_indexA(mask) == ref["calculate_indexA"](mask) == r
        Bm = compat.getB(n, s)                                             # ours has the same structure
        assert np.array_equal(Bm != 0, B != 0) and np.allclose(ref["getA"](Bm, n, s) @ Bm, 1.0)
    y, p = rng.choice([-1.0, 1.0], 50), rng.randn(50)
    assert np.isclose(compat.calculate_loss(y, p, 50), ref["calculate_loss"](y, p, 50))
    assert np.isclose(compat.calculate_mse(y, p, 50), ref["calculate_mse"](y, p, 50))
    assert compat.compare([1, 0, 1], [1, 1, 0]) == ref["compare"]([1, 0, 1], [1, 1, 0])
    M, v = rng.randn(70, 64), rng.randn(9)                                  # 4480 values: the native writer's path
    for name, fn_a, fn_b, obj in (("m", compat.save_matrix, ref["save_matrix"], M), ("v", compat.save_vector, ref["save_vector"], v)):
        a, b = str(tmp_path / f"{name}_mine.dat"), str(tmp_path / f"{name}_ref.dat")
        fn_a(obj, a)
        fn_b(obj, b)
        assert open(a, "rb").read() == open(b, "rb").read()
        assert np.allclose(compat.load_data(b), ref["load_data"](a))
    from scipy.sparse import random as sprandom
    S = sprandom(20, 30, density=0.2, format="csr", random_state=1)
    compat.save_sparse_csr(str(tmp_path / "s_mine"), S)
    ref["save_sparse_csr"](str(tmp_path / "s_ref"), S)
    assert (ref["load_sparse_csr"](str(tmp_path / "s_mine")) != compat.load_sparse_csr(str(tmp_path / "s_ref"))).nnz == 0
