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
