"""Real-data preprocessor: CSV -> label-encode -> (pair features) -> bias -> 80/20 split ->
one-hot -> per-partition ``.npz`` files (role of ``src/arrange_real_data.py`` + ``util.py:49-55``).

    python -m erasurehead_b200.data.arrange n_procs input_dir real_dataset n_stragglers n_partitions partial_coded

(the reference's six positionals, arrange_real_data.py:23-25).  Datasets: ``amazon-dataset``
(``train.csv``: ACTION + 9 categorical columns; degree-2 interaction features except the pairs
(5,7) and (2,3), util.py:52-53), ``dna-dataset/dna`` (first 500000 lines of ``features.csv``),
``covtype`` (``covtype.csv`` / ``covtype.data[.gz]`` next to the script or sklearn's cached
copy; classes 1 vs 2 -> -1/+1), ``kc_house_data`` (``kc_house_data.csv``; price / 1e6).

Differences from the reference, all deliberate: Python 3 (``DataFrame.ix`` / ``OneHotEncoder(sparse=)``
are gone, Q10); pair features are identified by the exact pair of codes instead of Python's
process-dependent ``hash(tuple)``; the one-hot matrix is produced directly as ELL indices
(``offset[col] + code``), on the GPU when one is present (``ops.onehot``), and stored in the
reference's CSR ``.npz`` format.  Remainder rows are dropped like the reference does (:79,85).
"""
from __future__ import annotations

import itertools
import os
import sys
from typing import Optional, Tuple

import numpy as np

from .dataset import Dataset, ell_to_csr

AMAZON_SKIP_PAIRS = ((5, 7), (2, 3))          # util.py:52


def label_encode_columns(X: np.ndarray) -> np.ndarray:
    """Per-column ``LabelEncoder``: values -> dense ranks ``0..k-1`` in sorted order."""
    out = np.empty(X.shape, dtype=np.int64)
    for c in range(X.shape[1]):
        _, out[:, c] = np.unique(X[:, c], return_inverse=True)
    return out


def interaction_pairs(n_cols: int, degree: int = 2, skip=AMAZON_SKIP_PAIRS):
    """Column tuples used by ``interactionTermsAmazon`` (util.py:49-55)."""
    keep = []
    for idx in itertools.combinations(range(n_cols), degree):
        if any(a in idx and b in idx for a, b in skip):
            continue
        keep.append(idx)
    return keep


def interaction_terms(codes: np.ndarray, degree: int = 2, skip=AMAZON_SKIP_PAIRS) -> np.ndarray:
    """One integer id per distinct value-tuple of every kept column tuple."""
    cols = []
    for idx in interaction_pairs(codes.shape[1], degree, skip):
        key = np.zeros(codes.shape[0], dtype=np.int64)
        for c in idx:
            key = key * (int(codes[:, c].max()) + 1) + codes[:, c]
        cols.append(key)
    return np.stack(cols, axis=1) if cols else np.zeros((codes.shape[0], 0), dtype=np.int64)


def onehot_ell(codes_train: np.ndarray, codes_valid: np.ndarray, use_gpu: Optional[bool] = None):
    """OneHotEncoder fitted on train+valid (arrange_real_data.py:61-64) -> ELL index matrices.
    Returns ``(idx_train, idx_valid, n_cols)``."""
    both = np.vstack([codes_train, codes_valid])
    enc = label_encode_columns(both)
    sizes = enc.max(axis=0) + 1
    offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)
    n_cols = int(sizes.sum())
    if use_gpu is None:
        try:
            import torch
            use_gpu = torch.cuda.is_available()
        except Exception:
            use_gpu = False
    if use_gpu:
        from ..ops import onehot
        idx = onehot.assemble_ell(enc, offsets)
    else:
        idx = (enc + offsets[None, :]).astype(np.int32)
    nt = codes_train.shape[0]
    return idx[:nt], idx[nt:], n_cols


def split_80_20(X, y, seed: int = 0):
    from sklearn.model_selection import train_test_split
    return train_test_split(X, y, test_size=0.2, random_state=seed)           # :57


def build_onehot_dataset(trainX: np.ndarray, trainY: np.ndarray, partitions: int, pairs: bool = False,
                         encode_first: bool = True, name: str = "real", use_gpu: Optional[bool] = None) -> Dataset:
    codes = label_encode_columns(trainX) if encode_first else trainX
    if pairs:
        codes = np.hstack([codes, label_encode_columns(interaction_terms(codes))])   # :46-53
    codes = np.hstack([codes, np.zeros((codes.shape[0], 1), dtype=codes.dtype)])     # bias column :55
    X_tr, X_va, y_tr, y_va = split_80_20(codes, np.asarray(trainY, dtype=np.float64))
    idx_tr, idx_va, n_cols = onehot_ell(X_tr, X_va, use_gpu)
    rpp = idx_tr.shape[0] // partitions                                              # :79
    parts = [ell_to_csr(idx_tr[i * rpp:(i + 1) * rpp], n_cols) for i in range(partitions)]
    return Dataset(parts=parts, y=y_tr, X_test=ell_to_csr(idx_va, n_cols), y_test=y_va, n_cols=n_cols,
                   rows_per_part=rpp, name=name)


def load_raw(real_dataset: str, input_dir: str) -> Tuple[np.ndarray, np.ndarray, dict]:
    """Read the raw table of one of the four datasets -> (X, y, options)."""
    import pandas as pd
    if real_dataset == "amazon-dataset":
        df = pd.read_csv(os.path.join(input_dir, "train.csv"))
        X = df.loc[:, "RESOURCE":].values                                            # :38 (was .ix)
        y = 2 * df["ACTION"].values - 1                                              # :44
        return X, y, dict(pairs=True)
    if real_dataset == "dna-dataset/dna":
        with open(os.path.join(input_dir, "features.csv")) as fin:
            data = np.genfromtxt(itertools.islice(fin, 0, 500000, 1), delimiter=",")  # :97-98
        return data[:, 1:], data[:, 0], dict(pairs=False)
    if real_dataset == "covtype":
        X = y = None
        for fn in ("covtype.csv", "covtype.data", "covtype.data.gz"):
            p = os.path.join(input_dir, fn)
            if os.path.exists(p):
                arr = pd.read_csv(p, header=None).values
                X, y = arr[:, :-1], arr[:, -1]
                break
        if X is None:
            import sklearn.datasets as datasets
            d = datasets.fetch_covtype()                                             # :147 (needs a cached copy)
            X, y = d.data, d.target
        keep = np.where(y <= 2)[0]                                                   # :154
        X, y = X[keep], y[keep]
        return X, np.where(y == 1, -1.0, 1.0), dict(pairs=False)                      # :156
    if real_dataset == "kc_house_data":
        df = pd.read_csv(os.path.join(input_dir, "kc_house_data.csv"))
        X = df.loc[:, "bedrooms":].values                                            # :211
        y = df["price"].values / 1e6                                                 # :225-226
        return X, y, dict(pairs=False)
    raise ValueError(f"unknown dataset {real_dataset!r}")


def arrange(n_procs: int, input_dir: str, real_dataset: str, n_stragglers: int, n_partitions: int,
            partial_coded: int, use_gpu: Optional[bool] = None) -> str:
    base = os.path.join(input_dir, real_dataset) + "/"
    X, y, opt = load_raw(real_dataset, base)
    if not partial_coded:
        partitions, out = n_procs - 1, f"{base}{n_procs - 1}/"                        # :69-71
    else:
        partitions = (n_procs - 1) * (n_partitions - n_stragglers)                    # :73-74
        out = f"{base}partial/{partitions}/"
    ds = build_onehot_dataset(X, y, partitions, pairs=opt["pairs"], name=real_dataset, use_gpu=use_gpu)
    print("No. of training samples = %d, Dimension = %d" % (ds.rows_per_part * partitions, ds.n_cols))
    print("No. of testing samples = %d, Dimension = %d" % (ds.X_test.shape[0], ds.n_cols))
    ds.save(out)
    print("Data Setup Finished.")
    return out


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) != 6:
        print("Usage: python -m erasurehead_b200.data.arrange n_procs input_dir real_dataset n_stragglers "
              "n_partitions partial_coded")
        return 0
    n_procs, input_dir, real_dataset, n_stragglers, n_partitions, partial_coded = argv
    arrange(int(n_procs), input_dir if input_dir.endswith("/") else input_dir + "/", real_dataset,
            int(n_stragglers), int(n_partitions), int(partial_coded))
    return 0


if __name__ == "__main__":
    sys.exit(main())
