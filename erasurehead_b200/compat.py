"""The reference's Python entry points under their own names.

A program written against the reference calls one of its ten driver functions
(``from naive import naive_logistic_regression`` ..., dispatched by main.py:62-92) and the helpers
of ``util.py``.  This module offers both under the same names and argument lists, so such a program
switches by changing its imports::

    from erasurehead_b200.compat import approx_logistic_regression, getB, load_data

    params = [num_itrs, alpha, learning_rate_schedule]               # main.py:48-51
    approx_logistic_regression(n_procs, n_rows, n_cols, data_dir, n_stragglers, is_real, params,
                               num_collect, add_delay, update_rule)

Differences that cannot be hidden: the number of processes is the number of GPUs, not ``n_procs``
(launch with torchrun; the ``n_procs - 1`` workers are spread over the ranks), and every scheme honours
``add_delay`` / ``update_rule`` where the reference's avoidstragg and partial drivers have neither
argument (they run AGD without delays here too).  Everything else -- data directory layout, result
files under ``<data_dir>results/`` with the reference's names, printed lines -- is ``--compat``
behaviour.  Each driver returns the dict of :func:`erasurehead_b200.cli.run` on
rank 0 (``result``: iterates / round times / arrival times, ``evaluation``: loss / AUC per iterate).
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np

from . import codes
from .config import RunConfig, triple_from_scheme
from .data.arrange import interaction_terms
from .data.generate import generate_random_binvec, generate_random_matrix_normal  # noqa: F401  (util.py:39-47)
from .utils.io import load_data, load_sparse_csr, save_matrix, save_sparse_csr, save_vector  # noqa: F401  (util.py:13-36)
from .utils.metrics import logistic_loss, mse

unique_rows = codes.unique_rows                 # util.py:59-62
compare = codes.compare_masks                   # util.py:105-111
calculate_indexA = codes.mask_rank              # util.py:125-134


def getB(n_workers: int, n_stragglers: int) -> np.ndarray:
    """Cyclic gradient code (util.py:64-83): support of s + 1 consecutive columns per row, B[i, i] = 1,
    every n - s rows span the all-ones vector.  Unseeded like the reference's."""
    return codes.cyclic_code(n_workers, n_stragglers, seed=None)


def getA(B: np.ndarray, n_workers: int, n_stragglers: int) -> np.ndarray:
    """Decode table (util.py:85-103): one row per straggler pattern, ordered as ``calculate_indexA`` ranks them."""
    return codes.decode_table(np.asarray(B, dtype=np.float64), n_stragglers)


def binary_search_row_wise(Aindex, completed, st=0, nd=None) -> int:
    """Row of ``completed`` in a mask table ordered like ``getA``'s patterns (util.py:113-123)."""
    return codes.find_mask_row(np.asarray(Aindex), completed)


def calculate_loss(y, predy, n_samples) -> float:
    """Mean logistic loss (util.py:136-137)."""
    return logistic_loss(np.asarray(y), np.asarray(predy))


def calculate_mse(y, predy, n_samples) -> float:
    """Mean squared error (util.py:139-140)."""
    return mse(np.asarray(y), np.asarray(predy))


def interactionTermsAmazon(data, degree: int = 2, hash=hash):
    """Pairwise feature crosses of the Amazon columns (util.py:49-57); the crosses are coded by index arithmetic
    instead of Python's ``hash`` of the tuple, which only renames the categories."""
    return interaction_terms(np.asarray(data), degree)


# ------------------------------------------------------------------------------------------ drivers
def _drive(scheme: str, n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
           add_delay=0, update_rule="AGD", num_collect=0, n_partitions=0, model: Optional[str] = None, **overrides):
    """``input_dir`` is what main.py hands its drivers: the resolved directory of the partition files."""
    from . import cli
    from .data import load_dataset
    from .parallel.context import init_distributed
    rounds, alpha, lr = int(params[0]), float(params[1]), params[2]
    lr = np.full(rounds, float(lr)) if np.isscalar(lr) else np.asarray(lr, dtype=np.float64)
    input_dir = input_dir if input_dir.endswith("/") else input_dir + "/"
    ic, pp, cv = triple_from_scheme(scheme, int(n_partitions))
    kw = dict(rounds=rounds, alpha=alpha, lr_values=lr, compat=True, output_dir=input_dir + "results/", model=model)
    kw.update(overrides)
    cfg = RunConfig(int(n_procs), int(n_samples), int(n_features), "", int(bool(is_real_data)), "x", ic, int(n_stragglers),
                    pp, cv, int(num_collect), int(add_delay), update_rule, **kw)
    ctx = init_distributed("cpu" if cfg.gpus < 0 else None)
    on_gpu = ctx.device.type == "cuda"
    data = load_dataset(input_dir, cfg.n_partitions, bool(is_real_data), cfg.n_cols,
                        rows_per_part=cfg.n_rows // cfg.n_partitions, lazy=on_gpu, dtype=np.float32 if on_gpu else float)
    return cli.run(cfg, data=data, ctx=ctx)


def naive_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                              add_delay, update_rule, **kw):
    """naive.py:11."""
    return _drive("naive", n_procs, n_samples, n_features, input_dir, 0, is_real_data, params, add_delay, update_rule, **kw)


def naive_linear_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                            add_delay, update_rule, **kw):
    """naive.py:214."""
    return _drive("naive", n_procs, n_samples, n_features, input_dir, 0, is_real_data, params, add_delay, update_rule,
                  model="linear", **kw)


def coded_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                              add_delay, update_rule, **kw):
    """coded.py:12."""
    return _drive("coded", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params, add_delay,
                  update_rule, **kw)


def replication_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                                    add_delay, update_rule, **kw):
    """replication.py:14."""
    return _drive("replication", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                  add_delay, update_rule, **kw)


def replication_linear_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                                  add_delay, update_rule, **kw):
    """replication.py:268."""
    return _drive("replication", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                  add_delay, update_rule, model="linear", **kw)


def avoidstragg_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params, **kw):
    """avoidstragg.py:11 (no delay / update-rule arguments: AGD, no injected delay)."""
    return _drive("avoidstragg", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params, **kw)


def approx_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                               num_collect, add_delay, update_rule, **kw):
    """approximate_coding.py:14."""
    return _drive("approx", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params, add_delay,
                  update_rule, num_collect=num_collect, **kw)


def approx_linear_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                             num_collect, add_delay, update_rule, **kw):
    """approximate_coding.py:270."""
    return _drive("approx", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params, add_delay,
                  update_rule, num_collect=num_collect, model="linear", **kw)


def partial_coded_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, n_partitions,
                                      is_real_data, params, **kw):
    """partial_coded.py:12."""
    return _drive("partial_coded", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                  n_partitions=n_partitions, **kw)


def partial_replication_logistic_regression(n_procs, n_samples, n_features, input_dir, n_stragglers, n_partitions,
                                            is_real_data, params, **kw):
    """partial_replication.py:11."""
    return _drive("partial_replication", n_procs, n_samples, n_features, input_dir, n_stragglers, is_real_data, params,
                  n_partitions=n_partitions, **kw)


DRIVERS: Sequence[str] = (
    "naive_logistic_regression", "naive_linear_regression", "coded_logistic_regression",
    "replication_logistic_regression", "replication_linear_regression", "avoidstragg_logistic_regression",
    "approx_logistic_regression", "approx_linear_regression", "partial_coded_logistic_regression",
    "partial_replication_logistic_regression",
)
