"""Real-data preprocessor (role of arrange_real_data.py) on synthetic CSVs of the right schema."""
import os

import numpy as np
import pandas as pd
import pytest

from erasurehead_b200.data import arrange as arr
from erasurehead_b200.data import load_dataset


def _amazon_csv(path, rows=400, seed=0):
    rng = np.random.RandomState(seed)
    cols = ["RESOURCE", "MGR_ID", "ROLE_ROLLUP_1", "ROLE_ROLLUP_2", "ROLE_DEPTNAME", "ROLE_TITLE",
            "ROLE_FAMILY_DESC", "ROLE_FAMILY", "ROLE_CODE"]
    df = pd.DataFrame({c: rng.randint(100, 100 + 5 + 3 * i, rows) for i, c in enumerate(cols)})
    df.insert(0, "ACTION", rng.randint(0, 2, rows))
    os.makedirs(path, exist_ok=True)
    df.to_csv(os.path.join(path, "train.csv"), index=False)
    return df


def test_interaction_pairs_match_reference_rule():
    pairs = arr.interaction_pairs(9)
    assert len(pairs) == 34 and (5, 7) not in pairs and (2, 3) not in pairs      # util.py:52-53
    codes = np.array([[0, 1, 2], [0, 1, 2], [1, 1, 2], [0, 0, 2]])
    t = arr.interaction_terms(codes, skip=())
    assert t.shape == (4, 3)
    assert t[0, 0] == t[1, 0] and len({t[0, 0], t[2, 0], t[3, 0]}) == 3          # same tuple <=> same id


def test_amazon_pipeline_layout_and_shapes(tmp_path):
    _amazon_csv(str(tmp_path / "amazon-dataset"))
    out = arr.arrange(5, str(tmp_path) + "/", "amazon-dataset", 1, 0, 0, use_gpu=False)
    assert out.endswith("amazon-dataset/4/")                                      # arrange_real_data.py:69-71
    ds = load_dataset(out, 4, True, 0)
    n_train = int(400 * 0.8)
    assert ds.parts[0].shape[0] == n_train // 4
    nnz = np.diff(ds.parts[0].indptr)
    assert np.all(nnz == 9 + 34 + 1)                                              # 9 raw + 34 pairs + bias = 44 ones/row
    assert np.all(ds.parts[0].data == 1.0)
    idx = ds.ell(0)
    assert np.all(idx[:, -1] == ds.parts[0].shape[1] - 1)                          # bias = last column
    assert set(np.unique(ds.y)) <= {-1.0, 1.0}                                     # 2*ACTION-1
    assert ds.X_test.shape[0] == 400 - n_train
    out2 = arr.arrange(5, str(tmp_path) + "/", "amazon-dataset", 1, 3, 1, use_gpu=False)
    assert out2.endswith("amazon-dataset/partial/8/")                              # :73-74


def test_kc_house_pipeline(tmp_path):
    rng = np.random.RandomState(1)
    rows = 300
    d = tmp_path / "kc_house_data"
    os.makedirs(d)
    df = pd.DataFrame({"id": np.arange(rows), "date": ["x"] * rows, "price": rng.uniform(1e5, 2e6, rows)})
    for c in ["bedrooms", "bathrooms", "sqft_living", "floors", "zipcode"]:
        df[c] = rng.randint(0, 12, rows)
    df.to_csv(d / "kc_house_data.csv", index=False)
    out = arr.arrange(4, str(tmp_path) + "/", "kc_house_data", 1, 0, 0, use_gpu=False)
    ds = load_dataset(out, 3, True, 0)
    assert np.all(np.diff(ds.parts[0].indptr) == 5 + 1)
    assert 0.09 < ds.y.min() and ds.y.max() < 2.1                                  # price / 1e6 (:225-226)


@pytest.mark.gpu
def test_onehot_device_kernels_match_host():
    import torch
    from erasurehead_b200.ops import onehot
    rng = np.random.RandomState(0)
    codes = rng.randint(0, 7, (500, 6)).astype(np.int64)
    sizes = codes.max(axis=0) + 1
    offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]])
    idx = onehot.assemble_ell(codes, offsets)
    assert np.array_equal(idx, (codes + offsets[None, :]).astype(np.int32))
    pairs = arr.interaction_pairs(6, skip=())
    got = onehot.pair_codes(torch.from_numpy(codes).cuda(), pairs).cpu().numpy()
    want = arr.interaction_terms(codes, skip=())
    # ids may differ, the induced partition of rows must not
    for q in range(len(pairs)):
        _, a = np.unique(got[:, q], return_inverse=True)
        _, b = np.unique(want[:, q], return_inverse=True)
        assert np.array_equal(a, b)


# ---------------------------------------------------------------------------------------------------------------
# Against the reference's own arrange_real_data.py (unmodified; baseline/py2run.py supplies Python-2 division and
# the two library interfaces it was written against, EH_REF_API_SHIMS)
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF_SRC = os.path.join(ROOT, "baseline", "_ref", "src")
needs_ref = pytest.mark.skipif(not (os.path.exists(os.path.join(REF_SRC, "arrange_real_data.py")) or os.path.exists("/root/reference/main.py")),
                               reason="reference tree not available")


def _reference_arrange(root, dataset, n_procs, s=1, P=0, partial=0):
    import subprocess
    import sys
    if not os.path.exists(os.path.join(REF_SRC, "arrange_real_data.py")):
        subprocess.check_call(["bash", os.path.join(ROOT, "baseline", "install_ref.sh")], stdout=subprocess.DEVNULL)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "py2run.py"), os.path.join(REF_SRC, "arrange_real_data.py"),
                        str(n_procs), root, dataset, str(s), str(P), str(partial)], capture_output=True, text=True, cwd=REF_SRC,
                       timeout=300, env=dict(os.environ, EH_REF_API_SHIMS="1"))
    assert p.returncode == 0 and "Data Setup Finished." in p.stdout, (p.stdout + p.stderr)[-2000:]


def _kc_csv(d, rows=300, seed=1):
    rng = np.random.RandomState(seed)
    os.makedirs(d)
    df = pd.DataFrame({"id": np.arange(rows), "date": ["x"] * rows, "price": rng.uniform(1e5, 2e6, rows)})
    for c in ["bedrooms", "bathrooms", "sqft_living", "floors", "zipcode"]:
        df[c] = rng.randint(0, 12, rows)
    df.to_csv(os.path.join(d, "kc_house_data.csv"), index=False)


@needs_ref
def test_kc_house_files_equal_the_reference_preprocessor(tmp_path):
    """Same CSV through both preprocessors: identical partitions, test matrix and labels."""
    ref_root, my_root = str(tmp_path / "ref") + "/", str(tmp_path / "mine") + "/"
    _kc_csv(ref_root + "kc_house_data")
    _kc_csv(my_root + "kc_house_data")
    _reference_arrange(ref_root, "kc_house_data", 4)
    arr.arrange(4, my_root, "kc_house_data", 1, 0, 0, use_gpu=False)
    a, b = load_dataset(ref_root + "kc_house_data/3/", 3, True, 0), load_dataset(my_root + "kc_house_data/3/", 3, True, 0)
    for k in range(3):
        assert a.parts[k].shape == b.parts[k].shape and (a.parts[k] != b.parts[k]).nnz == 0
    assert (a.X_test != b.X_test).nnz == 0
    assert np.allclose(a.y, b.y) and np.allclose(a.y_test, b.y_test)


@needs_ref
def test_amazon_files_equal_the_reference_preprocessor_up_to_pair_category_order(tmp_path):
    """The Amazon pipeline adds 34 pair features.  The reference names a pair by Python's hash of the tuple, this
    framework by the pair of codes, so the categories of a pair feature are numbered differently: the one-hot
    columns of such a block are a permutation of the reference's.  Everything invariant under that is compared:
    shapes, ones per row, the raw-feature block exactly, labels, and row-by-row agreement X X^T."""
    ref_root, my_root = str(tmp_path / "ref") + "/", str(tmp_path / "mine") + "/"
    _amazon_csv(ref_root + "amazon-dataset", rows=200)
    _amazon_csv(my_root + "amazon-dataset", rows=200)
    _reference_arrange(ref_root, "amazon-dataset", 5)
    arr.arrange(5, my_root, "amazon-dataset", 1, 0, 0, use_gpu=False)
    a, b = load_dataset(ref_root + "amazon-dataset/4/", 4, True, 0), load_dataset(my_root + "amazon-dataset/4/", 4, True, 0)
    assert np.allclose(a.y, b.y) and np.allclose(a.y_test, b.y_test)
    for k in range(4):
        A, B = a.parts[k], b.parts[k]
        assert A.shape == B.shape
        assert np.array_equal(np.diff(A.indptr), np.diff(B.indptr)) and np.all(np.diff(A.indptr) == 44)
        assert ((A @ A.T) != (B @ B.T)).nnz == 0            # which rows share which categories
    # the nine raw features come first and are numbered by sorted value on both sides: that block is identical
    raw = pd.read_csv(my_root + "amazon-dataset/train.csv").loc[:, "RESOURCE":]
    n_raw = int(sum(raw[c].nunique() for c in raw.columns))
    assert (a.parts[0][:, :n_raw] != b.parts[0][:, :n_raw]).nnz == 0
    assert (a.X_test @ a.X_test.T != b.X_test @ b.X_test.T).nnz == 0


@needs_ref
def test_covtype_files_equal_the_reference_preprocessor(tmp_path):
    """The reference takes Covertype from sklearn's cache (fetch_covtype, arrange_real_data.py:147); a small table
    of the same schema is planted in a private cache (SCIKIT_LEARN_DATA) for it and written as covtype.csv for
    this framework's reader.  Classes 1 and 2 are kept and mapped to -1 / +1 on both sides."""
    import joblib
    from sklearn.datasets._base import _pkl_filepath
    rng = np.random.RandomState(3)
    rows = 400
    X = np.hstack([rng.randint(0, 9, (rows, 10)), rng.randint(0, 2, (rows, 44))]).astype(np.float64)
    y = rng.randint(1, 8, rows).astype(np.int32)
    cache = tmp_path / "sk" / "covertype"
    os.makedirs(cache)
    joblib.dump(X, _pkl_filepath(str(cache), "samples"), compress=3)
    joblib.dump(y, _pkl_filepath(str(cache), "targets"), compress=3)
    ref_root, my_root = str(tmp_path / "ref") + "/", str(tmp_path / "mine") + "/"
    os.makedirs(ref_root + "covtype")
    os.makedirs(my_root + "covtype")
    pd.DataFrame(np.hstack([X, y[:, None]])).to_csv(my_root + "covtype/covtype.csv", header=False, index=False)
    old = os.environ.get("SCIKIT_LEARN_DATA")
    os.environ["SCIKIT_LEARN_DATA"] = str(tmp_path / "sk")
    try:
        _reference_arrange(ref_root, "covtype", 4)
    finally:
        if old is None:
            os.environ.pop("SCIKIT_LEARN_DATA")
        else:
            os.environ["SCIKIT_LEARN_DATA"] = old
    arr.arrange(4, my_root, "covtype", 1, 0, 0, use_gpu=False)
    a, b = load_dataset(ref_root + "covtype/3/", 3, True, 0), load_dataset(my_root + "covtype/3/", 3, True, 0)
    assert set(np.unique(a.y)) <= {-1.0, 1.0} and np.allclose(a.y, b.y) and np.allclose(a.y_test, b.y_test)
    for k in range(3):
        assert a.parts[k].shape == b.parts[k].shape and (a.parts[k] != b.parts[k]).nnz == 0
    assert (a.X_test != b.X_test).nnz == 0


@needs_ref
def test_dna_files_equal_the_reference_preprocessor(tmp_path):
    """dna-dataset/dna: label in the first column of features.csv, categorical features after it."""
    rng = np.random.RandomState(4)
    rows = 250
    tab = np.hstack([rng.choice([-1.0, 1.0], (rows, 1)), rng.randint(0, 4, (rows, 12)).astype(np.float64)])
    ref_root, my_root = str(tmp_path / "ref") + "/", str(tmp_path / "mine") + "/"
    for root in (ref_root, my_root):
        os.makedirs(root + "dna-dataset/dna")
        np.savetxt(root + "dna-dataset/dna/features.csv", tab, delimiter=",", fmt="%g")
    _reference_arrange(ref_root, "dna-dataset/dna", 3)
    arr.arrange(3, my_root, "dna-dataset/dna", 1, 0, 0, use_gpu=False)
    a = load_dataset(ref_root + "dna-dataset/dna/2/", 2, True, 0)
    b = load_dataset(my_root + "dna-dataset/dna/2/", 2, True, 0)
    assert np.allclose(a.y, b.y) and np.allclose(a.y_test, b.y_test)
    for k in range(2):
        assert a.parts[k].shape == b.parts[k].shape and (a.parts[k] != b.parts[k]).nnz == 0
    assert (a.X_test != b.X_test).nnz == 0
