"""Synthetic dense data: two-component GMM features with logistic labels.

Same distribution as the reference generator (generate_data.py:8-46, util.py:39-47):
``beta* in {-1,+1}^d``; each row is ``N(+-(1.5/d) beta*, (10/sqrt(d))^2 I)`` with the mixture
split drawn as ``Binomial(rows, 1/2)`` per partition; labels ``2 Bernoulli(sigmoid(x.beta*)) - 1``;
a test set of ``0.2 N`` rows.  Unlike the reference the generator is seeded.
``task='linear'`` produces real-valued targets ``x.beta*/sqrt(d) + noise`` for the
linear-regression drivers.
"""
from __future__ import annotations

import numpy as np

from .dataset import Dataset


def _gmm_block(rng, mu1, mu2, rows, d):
    c2 = rng.binomial(rows, 0.5)
    c1 = rows - c2
    mfac = 10.0 / np.sqrt(d)
    return np.concatenate((mfac * rng.standard_normal((c1, d)) + mu1,
                           mfac * rng.standard_normal((c2, d)) + mu2))


def generate_random_matrix_normal(mu1, mu2, n_rows, n_cols, rng=None):
    """Two-component GMM block with identity covariance scaled by 10/sqrt(d) (util.py:39-43)."""
    return _gmm_block(rng or np.random, np.asarray(mu1), np.asarray(mu2), n_rows, n_cols)


def generate_random_binvec(n, rng=None):
    """Random +-1 vector (util.py:46-47)."""
    return (rng or np.random).randint(2, size=n) * 2 - 1


def _labels(rng, X, beta_star, task, classes=1):
    if task == "linear":
        return X.dot(beta_star) / np.sqrt(len(beta_star)) + 0.1 * rng.standard_normal(X.shape[0])
    if task == "softmax":
        W = rng.standard_normal((len(beta_star), classes))
        return np.argmax(X.dot(W) + rng.gumbel(size=(X.shape[0], classes)), axis=1).astype(np.float64)
    prob = 1.0 / (1.0 + np.exp(-X.dot(beta_star)))
    return 2.0 * rng.binomial(1, prob) - 1.0


def generate_gmm(n_rows: int, n_cols: int, n_partitions: int, seed: int = 0,
                 task: str = "logistic", classes: int = 1, test_frac: float = 0.2) -> Dataset:
    if n_rows % n_partitions:
        raise ValueError("n_rows must be a multiple of the number of partitions")  # generate_data.py:11
    rng = np.random.RandomState(seed)
    rpp = n_rows // n_partitions
    beta_star = rng.randint(2, size=n_cols) * 2.0 - 1.0
    mu1, mu2 = (1.5 / n_cols) * beta_star, (-1.5 / n_cols) * beta_star
    if task == "softmax":
        Wrng = np.random.RandomState(seed + 1)
        W = Wrng.standard_normal((n_cols, classes))
        lab = lambda X: np.argmax(X.dot(W) + rng.gumbel(size=(X.shape[0], classes)), axis=1).astype(np.float64)
    else:
        lab = lambda X: _labels(rng, X, beta_star, task)
    parts, ys = [], []
    for _ in range(n_partitions):
        X = _gmm_block(rng, mu1, mu2, rpp, n_cols)
        parts.append(X)
        ys.append(lab(X))
    n_test = max(1, int(test_frac * n_rows))
    X_test = _gmm_block(rng, mu1, mu2, n_test, n_cols)
    y_test = lab(X_test)
    return Dataset(parts=parts, y=np.concatenate(ys), X_test=X_test, y_test=y_test,
                   n_cols=n_cols, rows_per_part=rpp, name=f"artificial-data/{n_rows}x{n_cols}")


def generate_gmm_to_disk(n_procs: int, n_rows: int, n_cols: int, output_dir: str, n_stragglers: int,
                         n_partitions: int, partial_coded: int, seed: int = 0) -> str:
    """CLI twin of ``generate_data.py`` (same 7 arguments, same directory layout :59-62)."""
    output_dir = output_dir if output_dir.endswith("/") else output_dir + "/"
    base = f"{output_dir}artificial-data/{n_rows}x{n_cols}/"
    if not partial_coded:
        parts, out = n_procs - 1, f"{base}{n_procs - 1}/"
    else:
        parts = (n_procs - 1) * (n_partitions - n_stragglers)
        out = f"{base}partial/{parts}/"
    generate_gmm(n_rows, n_cols, parts, seed=seed).save(out)
    return out


def main(argv=None) -> int:
    """CLI with the seven positionals of ``generate_data.py:50-52``."""
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) != 7:
        print("Usage: python -m erasurehead_b200.data.generate n_procs n_rows n_cols output_dir n_stragglers "
              "n_partitions partial_coded")
        return 0
    n_procs, n_rows, n_cols, out, s, P, partial = argv
    path = generate_gmm_to_disk(int(n_procs), int(n_rows), int(n_cols), out, int(s), int(P), int(partial))
    print("Data Generation Finished.", path)
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
