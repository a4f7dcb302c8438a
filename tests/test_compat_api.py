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
