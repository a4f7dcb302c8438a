"""File formats, directory layout, result files and the CLI end to end (CPU engine)."""
import os

import numpy as np
import pytest
from scipy.sparse import csr_matrix

from erasurehead_b200 import cli
from erasurehead_b200.config import parse_argv
from erasurehead_b200.data import generate_gmm_to_disk, generate_onehot, load_dataset
from erasurehead_b200.data.dataset import csr_to_ell, ell_to_csr
from erasurehead_b200.utils import io, metrics


def test_vector_and_matrix_formats(tmp_path):
    p = str(tmp_path / "v.dat")
    io.save_vector([1.0, -1.0, 0.12345], p)
    assert open(p).read() == "1.000 \n-1.000 \n0.123 \n"          # util.py:32-36 "%5.3f" + " "
    m = np.array([[1.5, -2.25], [3.0, 4.0]])
    q = str(tmp_path / "m.dat")
    io.save_matrix(m, q)
    assert open(q).read().splitlines()[0] == "1.5 -2.25"          # util.py:26-30 str(float)
    assert np.array_equal(io.load_data(q), m)
    s = csr_matrix(np.array([[0, 1.0, 0], [1.0, 0, 1.0]]))
    io.save_sparse_csr(str(tmp_path / "s"), s)
    z = np.load(str(tmp_path / "s.npz"))
    assert set(z.files) == {"data", "indices", "indptr", "shape"}  # util.py:17-19
    assert (io.load_sparse_csr(str(tmp_path / "s")) != s).nnz == 0


def test_auc_matches_sklearn():
    from sklearn.metrics import auc, roc_curve
    rng = np.random.RandomState(0)
    for ties in (False, True):
        y = rng.choice([-1.0, 1.0], 500)
        s = rng.randn(500) + 0.5 * y
        if ties:
            s = np.round(s, 1)
        fpr, tpr, _ = roc_curve(y, s, pos_label=1)
        assert metrics.roc_auc(y, s) == pytest.approx(auc(fpr, tpr), abs=1e-12)
    assert metrics.logistic_loss([1, -1], [0.0, 0.0]) == pytest.approx(np.log(2))
    assert metrics.mse([1, 2], [1, 4]) == 2.0


def test_generate_to_disk_layout_and_reload(tmp_path):
    out = generate_gmm_to_disk(5, 400, 6, str(tmp_path), 1, 0, 0, seed=1)
    assert out.endswith("artificial-data/400x6/4/")                # generate_data.py:59-60
    assert sorted(os.listdir(out)) == ["1.dat", "2.dat", "3.dat", "4.dat", "label.dat", "label_test.dat",
                                       "test_data.dat"]
    ds = load_dataset(out, 4, False, 6)
    assert ds.parts[0].shape == (100, 6) and ds.X_test.shape == (80, 6) and set(np.unique(ds.y)) <= {-1.0, 1.0}
    out = generate_gmm_to_disk(5, 480, 6, str(tmp_path), 1, 4, 1, seed=1)
    assert out.endswith("artificial-data/480x6/partial/12/")       # generate_data.py:61-62


def test_onehot_shapes_and_ell_roundtrip():
    ds = generate_onehot(600, 500, 9, 4, seed=2)
    assert ds.is_sparse and ds.parts[0].shape == (150, 500)
    idx = ds.ell(0)
    assert idx.shape == (150, 9) and np.all(idx[:, -1] == 499)      # bias column is always on
    assert (ell_to_csr(idx, 500) != ds.parts[0]).nnz == 0
    assert np.array_equal(csr_to_ell(ds.parts[1]), ds.ell(1))


@pytest.mark.parametrize("scheme_args,prefix,files", [
    ("1 1 0 3 5 1 AGD", "approx_acc_1_5", ["training_loss", "testing_loss", "auc", "timeset", "worker_timeset"]),
    ("1 2 0 0 0 1 GD", "coded_acc_2", ["training_loss", "auc"]),
    ("0 0 0 1 0 1 GD", "naive_acc", ["timeset"]),
])
def test_cli_end_to_end_dense(tmp_path, capsys, scheme_args, prefix, files):
    generate_gmm_to_disk(9, 800, 6, str(tmp_path), 1, 0, 0, seed=1)
    argv = f"9 800 6 {tmp_path} 0 x {scheme_args} --rounds=6 --gpus=-1".split()
    out = cli.run(parse_argv(argv))
    res_dir = out["results_dir"]
    assert res_dir == f"{tmp_path}/artificial-data/800x6/8/results/"     # naive.py:200-202
    for f in files:
        assert os.path.exists(f"{res_dir}{prefix}_{f}.dat"), f
    assert np.load(f"{res_dir}{prefix}_betaset.npy").shape == (6, 6)
    wt = io.load_data(f"{res_dir}{prefix}_worker_timeset.dat")
    assert wt.shape == (6, 8)
    txt = capsys.readouterr().out
    assert "Iteration 5: Train Loss =" in txt and "AUC =" in txt and "Total Time Elapsed:" in txt
    assert ">>> At Iteration 0" in txt


def test_cli_synthetic_without_files_or_gpu(tmp_path, capsys):
    """README quick-start: --synthetic needs no data files; without a GPU the host generator stands in."""
    argv = f"9 1600 12 {tmp_path}/nowhere 0 artificial 1 1 0 3 6 1 AGD --synthetic --rounds=4 --gpus=-1 --output-dir={tmp_path}/out/".split()
    out = cli.run(parse_argv(argv))
    assert os.path.exists(f"{tmp_path}/out/approx_acc_1_6_training_loss.dat")
    assert "Iteration 3: Train Loss =" in capsys.readouterr().out
    assert out["results_dir"] == f"{tmp_path}/out/"


def test_cli_sparse_partial_and_compat_names(tmp_path):
    n, s, P = 4, 1, 3
    ds = generate_onehot(8 * 50, 120, 7, (P - s) * n, seed=3)
    d = f"{tmp_path}/amazon-dataset/partial/{(P - s) * n}/"
    ds.save(d)
    argv = f"5 400 120 {tmp_path} 1 amazon-dataset 1 {s} {P} 0 0 0 AGD --rounds=4 --gpus=-1 --compat".split()
    out = cli.run(parse_argv(argv))
    rd = out["results_dir"]
    assert rd == d + "results/"
    # Q3: partial_coded writes training loss under the partialreplication_ prefix in compat mode
    assert os.path.exists(rd + f"partialreplication_{s}_{P}_training_loss.dat")
    assert os.path.exists(rd + f"partialcoded_{s}_{P}_testing_loss.dat")


def test_checkpoint_resume_cpu(tmp_path):
    generate_gmm_to_disk(9, 800, 6, str(tmp_path), 1, 0, 0, seed=1)
    base = f"9 800 6 {tmp_path} 0 x 1 1 0 3 5 1 AGD --gpus=-1 --quiet"
    full = cli.run(parse_argv((base + " --rounds=8").split()))["result"]
    half = cli.run(parse_argv((base + " --rounds=4").split()))
    from erasurehead_b200 import Engine
    from erasurehead_b200.data import load_dataset as ld
    cfg = parse_argv((base + " --rounds=4").split())
    eng = Engine(cfg, ld(cfg.data_dir, 8, False, 6), device="cpu")
    eng.run()
    ck = str(tmp_path / "ck.npz")
    eng.save_checkpoint(ck, 4)
    cfg2 = parse_argv((base + f" --rounds=8 --resume={ck}").split())
    eng2 = Engine(cfg2, ld(cfg2.data_dir, 8, False, 6), device="cpu")
    assert eng2.first_round == 4
    res2 = eng2.run()
    assert np.allclose(res2.betaset, full.betaset[4:], atol=1e-12)
    assert np.allclose(res2.timeset, full.timeset[4:])


def test_metrics_jsonl_and_npy_outputs(tmp_path):
    import json
    generate_gmm_to_disk(9, 800, 6, str(tmp_path), 1, 0, 0, seed=1)
    mj = str(tmp_path / "m.jsonl")
    out = cli.run(parse_argv(f"9 800 6 {tmp_path} 0 x 1 1 0 3 5 1 AGD --rounds=5 --gpus=-1 --quiet --metrics-jsonl={mj}".split()))
    recs = [json.loads(l) for l in open(mj)]
    assert len(recs) == 5 and {"round", "time_s", "n_collected", "arrived_mask", "weights", "train_loss", "auc"} <= set(recs[0])
    assert bin(recs[0]["arrived_mask"]).count("1") == recs[0]["n_collected"]
    rd = out["results_dir"]
    for suffix in ("betaset", "timeset", "worker_timeset", "training_loss", "testing_loss", "auc"):
        assert os.path.exists(f"{rd}approx_acc_1_5_{suffix}.npy")


def test_predict_helpers_host():
    from erasurehead_b200 import predict
    rng = np.random.RandomState(0)
    X, B = rng.randn(20, 5), rng.randn(3, 5)
    assert np.allclose(predict.margins(X, B), X @ B.T)
    assert np.allclose(predict.predict_proba(X, B[0]), 1 / (1 + np.exp(-X @ B[0])))
    assert set(np.unique(predict.predict_label(X, B[0]))) <= {-1.0, 1.0}
    assert predict.predict_class(X, rng.randn(5, 4)).shape == (20,)


def test_cli_synthetic_onehot_shape_softmax_without_files_or_gpu(tmp_path, capsys):
    """--synthetic with is_real=1: one-hot data of a dataset shape (here a small custom one via --synthetic-nnz),
    multi-class model, straight from the 13 positionals; on a machine without a GPU the oracle runs it."""
    from erasurehead_b200 import cli
    from erasurehead_b200.config import parse_argv
    args = ["7", "1200", "400", str(tmp_path) + "/", "1", "myshape", "1", "2", "0", "3", "3", "0", "AGD",
            "--synthetic", "--synthetic-nnz=9", "--model=softmax", "--classes=4", "--rounds=6", "--gpus=-1",
            "--init=zeros", "--eta0=1.0", "--quiet", "--abandon", "--decoder=inline"]
    out = cli.run(parse_argv(args))
    ev = out["evaluation"]
    assert out["result"].betaset.shape == (6, 400, 4)
    assert np.all(np.isfinite(ev.training_loss)) and ev.training_loss[-1] < ev.training_loss[0]
    import pytest
    with pytest.raises(ValueError, match="synthetic-nnz"):
        cli.run(parse_argv(args[:5] + ["unknown-shape"] + args[6:14] + ["--gpus=-1", "--quiet"]))


def test_native_text_reader_matches_numpy_bit_for_bit(tmp_path):
    """csrc/textio.cu behind utils.io.load_data: correctly rounded decimals (17-digit shortest representations,
    half-way cases, subnormals, infinities), float32 and float64 output, vectors, and a clear error on a ragged row."""
    from erasurehead_b200.utils import io
    rng = np.random.RandomState(0)
    X = rng.standard_normal((700, 131)) * np.exp(rng.uniform(-40, 40, (700, 131)))
    X[0, :4] = [0.0, -0.0, 1e308, 5e-324]
    X[1, :3] = [123456789012345678901234.0, np.inf, -1.0]
    p = str(tmp_path / "a.dat")
    io.save_matrix(X, p)
    try:
        A = io._load_data_native(p)
    except Exception as e:           # no nvcc / no library on this machine: the numpy path is the only one
        import pytest
        pytest.skip(f"native library unavailable: {e}")
    assert np.array_equal(A, np.loadtxt(p)) and np.array_equal(A, X)
    with np.errstate(over="ignore"):
        assert np.array_equal(io._load_data_native(p, np.float32), X.astype(np.float32))
    assert np.array_equal(io.load_data(p), X)                       # public entry point takes the native path (> 64 KB)
    tricky = ["9007199254740993", "9007199254740992.5", "0.1", "1e23", "8.41e21", "2.2250738585072014e-308",
              "1.7976931348623157e308", "4.35e-17", "123456789.123456789e5", "+3.5", "-.5", "7."]
    q = str(tmp_path / "t.dat")
    open(q, "w").write(("\n".join(tricky) + "\n") * 3000)
    T = io._load_data_native(q)
    assert T.shape == (len(tricky) * 3000,) and np.array_equal(T[:len(tricky)], np.array([float(v) for v in tricky]))
    bad = str(tmp_path / "bad.dat")
    open(bad, "w").write("1 2 3\n" * 5000 + "1 2\n")
    import pytest
    with pytest.raises(Exception, match="row 5001 does not hold exactly 3"):
        io._load_data_native(bad)


def test_lazy_dataset_reads_partitions_on_demand(tmp_path):
    """load_dataset(lazy=True): a rank only reads the partition files it uses (and keeps the last two in memory);
    float32 on request; same values as the eager reader."""
    from erasurehead_b200.data import generate_gmm_to_disk, load_dataset
    from erasurehead_b200.data.shapes import generate_onehot
    out = generate_gmm_to_disk(5, 4 * 300, 40, str(tmp_path), 1, 0, 0, seed=3)
    a = load_dataset(out, 4, False, 40)
    b = load_dataset(out, 4, False, 40, lazy=True, dtype=np.float32)
    assert b.n_partitions == 4 and not b.is_sparse and b.rows_per_part == 300
    assert list(b.parts._cache) == [0]                       # is_sparse looked at partition 0 only
    assert b.partition(2).dtype == np.float32 and np.allclose(a.partition(2), b.partition(2), atol=1e-6)
    b.partition(3)
    assert sorted(b.parts._cache) == [2, 3]
    X, y = b.train_matrix(3)
    assert X.shape == (900, 40) and np.allclose(X, np.vstack(a.parts[:3]), atol=1e-6)
    generate_onehot(4 * 200, 500, 9, 4, seed=1).save(str(tmp_path / "oh") + "/")
    c = load_dataset(str(tmp_path / "oh") + "/", 4, True, 500, lazy=True)
    assert c.is_sparse and c.ell(3).shape == (200, 9) and c.labels(3).shape == (200,)


def test_cli_checkpoint_every_writes_a_resumable_file_on_cpu(tmp_path):
    """--checkpoint-every on the CPU path: the file holds the state after the last full chunk (7 rounds, every 3
    -> after round 6), and a run resumed from it reproduces the remaining iterates of an uninterrupted run."""
    generate_gmm_to_disk(9, 800, 6, str(tmp_path), 1, 0, 0, seed=1)
    base = f"9 800 6 {tmp_path} 0 x 1 1 0 3 5 1 AGD --gpus=-1 --quiet"
    full = cli.run(parse_argv((base + " --rounds=9").split()))["result"]
    first = cli.run(parse_argv((base + " --rounds=7 --checkpoint-every=3").split()))
    ck = first["results_dir"] + "approx_acc_1_5_checkpoint.npz"
    assert os.path.exists(ck) and int(np.load(ck)["next_round"]) == 6
    rest = cli.run(parse_argv((base + f" --rounds=9 --resume={ck}").split()))["result"]
    assert rest.betaset.shape[0] == 3
    assert np.allclose(rest.betaset, full.betaset[6:], atol=1e-12)
    with pytest.raises(ValueError):
        parse_argv((base + " --resume").split())


def test_native_text_writer_is_byte_identical_to_python_repr(tmp_path):
    """csrc/textio.cu eh_save_dense_text: the reference's `" ".join(str(x))` rows (util.py:26-30), i.e. Python's
    repr of every double -- shortest round-trip digits, positional for 1e-4 <= |x| < 1e16, else exponent form."""
    import ctypes as C
    try:
        from erasurehead_b200.ops import _lib
        L = _lib.lib()
        L.eh_format_repr
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"native library unavailable: {e}")
    buf = C.create_string_buffer(40)

    def nat(v):
        n = L.eh_format_repr(float(v), buf)
        return buf.raw[:n].decode()
    edge = [0.0, -0.0, 1.0, -1.0, 0.1, 1e-4, 9.999e-5, 1e-5, 1e15, 1e16, 9999999999999998.0, 1.2345e16, 123456.789, 1e22,
            1e23, 5e-324, 1.7976931348623157e308, 2.2250738585072014e-308, 1e100, 1e-100, 123456789012345.6, 100.0,
            12345678901234567.0, 2.5e-7, float("inf"), float("-inf"), float("nan"), 1e7, 65536.0, 1 / 3]
    assert [nat(v) for v in edge] == [repr(v) for v in edge]
    rng = np.random.RandomState(0)
    vals = np.frombuffer(rng.bytes(8 * 20000), dtype=np.float64)              # every magnitude and digit count
    vals = np.concatenate([vals[np.isfinite(vals)], rng.randn(20000), rng.randn(5000) * 1e-6, rng.randn(5000) * 1e18,
                           rng.randn(5000).astype(np.float32).astype(np.float64)])
    assert all(nat(v) == repr(float(v)) for v in vals)
    X = rng.randn(64, 128)
    X[0, :6] = [1e16, 1e-5, 100.0, -0.0, 123456.0, 5e-324]
    for M, name in ((X, "d"), (X.astype(np.float32), "f")):
        p, q = str(tmp_path / f"{name}_native.dat"), str(tmp_path / f"{name}_python.dat")
        io.save_matrix(M, p)                                                  # >= 4096 values: native path
        with open(q, "w") as f:
            for row in M:
                f.write(" ".join(repr(float(x)) for x in row) + "\n")
        assert open(p, "rb").read() == open(q, "rb").read()
        assert np.array_equal(io.load_data(p), M.astype(np.float64))


def test_selfcheck_runs_on_the_cpu_path(capsys):
    from erasurehead_b200 import selfcheck
    assert selfcheck.main(["--cpu"]) == 0
    out = capsys.readouterr().out
    assert out.count(" ok") == len(selfcheck.CASES) and "all schemes agree" in out
