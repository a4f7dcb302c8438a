"""Parity against the reference's own code: the unmodified copy (baseline/_ref, byte-identical; Python-2
division supplied at import time by baseline/py2run.py) runs over the in-repo mpi4py subset on the same data
files, and its result files are compared with ours.

Without injected delays the reference's arrival order is whatever the operating system makes it, so the
schemes compared are those whose trajectory does not depend on it: the exact ones (naive, cyclic-MDS coded,
FRC replication, the two partial schemes) and the approximate scheme asked to collect everything (it then stops
when every group is covered).  avoidstragg sums whichever n - s messages come first; its test pins the arrival
order through the MPI stand-in's link latencies instead.  Where the reference draws its start from numpy's unseeded global generator (naive, replication,
approximate: naive.py:23, replication.py:31, approximate_coding.py:32) the harness seeds that generator
(EH_NP_SEED, see py2run.py) with the value our ``--seed`` derives the same draw from."""
import os
import subprocess
import sys

import numpy as np
import pytest

from erasurehead_b200 import cli
from erasurehead_b200.config import parse_argv
from erasurehead_b200.data import generate_gmm_to_disk
from erasurehead_b200.utils import io

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF = "/root/reference"

pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "main.py")), reason="reference tree not mounted")


def run_reference(args, timeout=300, np_seed=None):
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "mpirun_shim.py"), "-np", args[0], "main.py"] + args
    env = dict(os.environ)
    if np_seed is not None:
        env["EH_NP_SEED"] = str(np_seed)
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    return p.stdout


@pytest.mark.parametrize("coded_ver,prefix,n_procs,s,P", [
    (1, "partialreplication_1_3", 5, 1, 3), (0, "partialcoded_1_3", 5, 1, 3),
    (1, "partialreplication_2_4", 7, 2, 4), (0, "partialcoded_2_4", 7, 2, 4),
])
def test_partial_schemes_match_the_reference_files(tmp_path, coded_ver, prefix, n_procs, s, P):
    rows, cols = 8 * 60, 10
    out = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, P, 1, seed=3)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "0", "x", "1", str(s), str(P), str(coded_ver),
            "0", "0", "AGD"]
    txt = run_reference(args)
    assert "Iteration 99: Train Loss" in txt
    ref_dir = out + "results/"
    ref_files = sorted(os.listdir(ref_dir))
    ref_test = io.load_data(ref_dir + prefix + "_testing_loss.dat")
    ref_auc = io.load_data(ref_dir + prefix + "_auc.dat")
    tl_prefix = f"partialreplication_{s}_{P}"       # Q3: partial_coded writes its training loss under this prefix
    ref_train = io.load_data(ref_dir + tl_prefix + "_training_loss.dat")
    for f in ref_files:
        os.rename(ref_dir + f, ref_dir + "ref_" + f)
    ours = cli.run(parse_argv(args + ["--gpus=-1", "--compat", "--quiet", "--seed=1"]))
    ev = ours["evaluation"]
    # the reference rounds to 3 decimals when writing
    assert np.abs(ev.training_loss - ref_train).max() <= 1.5e-3
    assert np.abs(ev.testing_loss - ref_test).max() <= 1.5e-3
    assert np.abs(ev.auc - ref_auc).max() <= 1.5e-3
    # same file names and formats
    mine = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat") and not f.startswith("ref_"))
    assert mine == [f for f in ref_files if f.endswith(".dat")]
    wt_ref = io.load_data(ref_dir + "ref_" + prefix + "_worker_timeset.dat")
    wt = io.load_data(ref_dir + prefix + "_worker_timeset.dat")
    assert wt.shape == wt_ref.shape == (100, n_procs - 1)


SEED = 5
NP_SEED = SEED + 7919          # oracle.initial_beta draws the start from RandomState(seed + 7919)


@pytest.mark.parametrize("name,triple,s,k,rule,prefix,n_procs", [
    ("naive", ("0", "0", "0"), 0, 0, "GD", "naive_acc", 5),
    ("naive", ("0", "0", "0"), 0, 0, "AGD", "naive_acc", 5),
    ("coded", ("1", "0", "0"), 1, 0, "AGD", "coded_acc_1", 5),
    ("replication", ("1", "0", "1"), 1, 0, "AGD", "replication_acc_1", 5),
    ("approx_collect_all", ("1", "0", "3"), 1, 4, "AGD", "replication_acc_1", 5),    # Q2: AGC writes replication's names
    ("coded_s2", ("1", "0", "0"), 2, 0, "GD", "coded_acc_2", 7),                      # 6 workers, any 4 decode
    ("replication_s2", ("1", "0", "1"), 2, 0, "AGD", "replication_acc_2", 7),        # two groups of three
])
def test_exact_schemes_match_the_reference_files(tmp_path, name, triple, s, k, rule, prefix, n_procs):
    rows, cols = 8 * 60, 10
    out = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, 0, 0, seed=3)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "0", "x", triple[0], str(s), triple[1], triple[2],
            str(k), "0", rule]
    txt = run_reference(args, np_seed=NP_SEED)
    assert "Iteration 99: Train Loss" in txt, txt[-1500:]
    ref_dir = out + "results/"
    ref_files = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat"))
    ref = {q: io.load_data(ref_dir + f"{prefix}_{q}.dat") for q in ("training_loss", "testing_loss", "auc")}
    for f in ref_files:
        os.rename(ref_dir + f, ref_dir + "ref_" + f)
    ours = cli.run(parse_argv(args + ["--gpus=-1", "--compat", "--quiet", f"--seed={SEED}"]))
    ev = ours["evaluation"]
    # the reference rounds to 3 decimals when writing
    assert np.abs(ev.training_loss - ref["training_loss"]).max() <= 1.5e-3
    assert np.abs(ev.testing_loss - ref["testing_loss"]).max() <= 1.5e-3
    assert np.abs(ev.auc - ref["auc"]).max() <= 1.5e-3
    mine = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat") and not f.startswith("ref_"))
    assert mine == ref_files


def test_one_hot_data_files_match_the_reference(tmp_path):
    """The real-data path: sparse .npz partitions of a one-hot dataset (arrange_real_data.py's output format),
    FRC replication, against the reference's sparse code path (load_sparse_csr + scipy products)."""
    from erasurehead_b200.data import generate_onehot
    n_procs, rows, cols, s, nnz = 5, 8 * 60, 40, 1, 6
    d = str(tmp_path) + "/covtype/" + str(n_procs - 1) + "/"
    generate_onehot(rows, cols, nnz, n_procs - 1, seed=2).save(d)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "1", "covtype", "1", str(s), "0", "1", "0", "0", "AGD"]
    txt = run_reference(args, np_seed=NP_SEED)
    assert "Iteration 99: Train Loss" in txt, txt[-1500:]
    ref_dir, prefix = d + "results/", "replication_acc_1"
    ref_files = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat"))
    ref = {q: io.load_data(ref_dir + f"{prefix}_{q}.dat") for q in ("training_loss", "testing_loss", "auc")}
    for f in ref_files:
        os.rename(ref_dir + f, ref_dir + "ref_" + f)
    ours = cli.run(parse_argv(args + ["--gpus=-1", "--compat", "--quiet", f"--seed={SEED}"]))
    ev = ours["evaluation"]
    assert np.abs(ev.training_loss - ref["training_loss"]).max() <= 1.5e-3
    assert np.abs(ev.testing_loss - ref["testing_loss"]).max() <= 1.5e-3
    assert np.abs(ev.auc - ref["auc"]).max() <= 1.5e-3
    mine = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat") and not f.startswith("ref_"))
    assert mine == ref_files


@pytest.mark.filterwarnings("ignore:overflow encountered")
def test_linear_regression_path_matches_the_reference_output(tmp_path):
    """dataset = kc_house_data selects linear regression (main.py:90-92).  The reference writes no files for it
    (naive.py:413-417), so its printed per-iteration MSE is compared.  With the stock constant step of 10 the
    iteration diverges on any one-hot data; both sides must diverge identically until float64 overflows."""
    import re
    from erasurehead_b200.data import generate_onehot
    n_procs, rows, cols, nnz = 5, 8 * 60, 40, 6
    d = str(tmp_path) + "/kc_house_data/" + str(n_procs - 1) + "/"
    generate_onehot(rows, cols, nnz, n_procs - 1, seed=2, task="linear").save(d)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "1", "kc_house_data", "0", "0", "0", "0", "0", "0", "GD"]
    txt = run_reference(args, np_seed=NP_SEED)
    pat = re.compile(r"Iteration (\d+): Train Loss = (\S+), Test Loss = (\S+),")
    ref = {int(m.group(1)): (float(m.group(2)), float(m.group(3))) for m in pat.finditer(txt)}
    assert sorted(ref) == list(range(100))
    ours = cli.run(parse_argv(args + ["--gpus=-1", "--compat", "--quiet", f"--seed={SEED}"]))
    ev = ours["evaluation"]
    finite = [i for i in range(100) if np.isfinite(ref[i][0]) and np.isfinite(ref[i][1])]
    assert len(finite) >= 20
    for i in finite:
        assert abs(ev.training_loss[i] - ref[i][0]) <= 1e-5 * abs(ref[i][0]) + 1e-6, i
        assert abs(ev.testing_loss[i] - ref[i][1]) <= 1e-5 * abs(ref[i][1]) + 1e-6, i


@pytest.mark.skipif(not os.environ.get("EH_SLOW_TESTS"), reason="two minutes of real Exp(0.5 s) sleeps each; set EH_SLOW_TESTS=1")
@pytest.mark.parametrize("name,coded_ver,k,prefix,n_procs", [
    ("approx", "3", 2, "replication_acc_1", 5),       # Q2: AGC writes replication's file names
    ("coded", "0", 0, "coded_acc_1", 5),
    ("replication", "1", 0, "replication_acc_1", 5),
    ("approx3groups", "3", 2, "replication_acc_1", 7),    # 3 groups of 2, stop after 2 arrivals: duplicates inside a group
])
def test_injected_delays_collect_the_same_workers_as_the_reference(tmp_path, name, coded_ver, k, prefix, n_procs):
    """The reference's own straggler model switched on: the seeded Exp(0.5 s) table (approximate_coding.py:198-207,
    coded.py, replication.py alike) decides who arrives first, so the set of workers the master collects in each
    round, and from it every iterate, must be the same on both sides (up to two delays closer than the
    reference's compute jitter).  The recorded outcome is in profiles/reference_parity_delays.txt."""
    rows, cols, s = 8 * 60, 10, 1
    out = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, 0, 0, seed=3)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "0", "x", "1", str(s), "0", coded_ver, str(k), "1", "AGD"]
    txt = run_reference(args, np_seed=NP_SEED, timeout=900)
    assert "Iteration 99: Train Loss" in txt, txt[-1500:]
    ref_dir = out + "results/"
    ref_wt = io.load_data(ref_dir + prefix + "_worker_timeset.dat")
    ref_train = io.load_data(ref_dir + prefix + "_training_loss.dat")
    ref_ts = io.load_data(ref_dir + prefix + "_timeset.dat")
    for f in os.listdir(ref_dir):
        os.rename(ref_dir + f, ref_dir + "ref_" + f)
    ours = cli.run(parse_argv(args + ["--gpus=-1", "--compat", "--quiet", f"--seed={SEED}"]))
    wt, tl = ours["result"].worker_timeset, ours["evaluation"].training_loss
    same = np.all((wt >= 0) == (ref_wt >= 0), axis=1)
    both = (wt >= 0) & (ref_wt >= 0)
    print(f"[{name}] rounds with the same collected set: {int(same.sum())} / {len(same)}; "
          f"max |arrival difference| = {np.abs(wt[both] - ref_wt[both]).max() * 1e3:.1f} ms; "
          f"max |train loss difference| = {np.abs(tl - ref_train).max():.1e}; "
          f"sum of round times reference / model = {ref_ts.sum():.3f} / {ours['result'].timeset.sum():.3f} s")
    assert same.mean() >= 0.97
    if same.all():
        assert np.abs(tl - ref_train).max() <= 1.5e-3
        # arrival times: the reference measures sleep + compute + transport, the model only the sleep
        assert np.abs(wt[both] - ref_wt[both]).max() < 0.05


@pytest.mark.parametrize("partial", [0, 1])
def test_files_written_by_the_reference_generator_are_read_here(tmp_path, partial):
    """The other direction of the file contract: generate_data.py of the reference (run unmodified, Python-2
    division supplied at import) writes the partitions, this framework's reader and CLI consume them."""
    from erasurehead_b200.data import load_dataset
    n_procs, rows, cols, s, P = 5, 8 * 60, 10, 1, 3
    ref_src = os.path.join(ROOT, "baseline", "_ref", "src")
    if not os.path.exists(os.path.join(ref_src, "generate_data.py")):
        subprocess.check_call(["bash", os.path.join(ROOT, "baseline", "install_ref.sh")], stdout=subprocess.DEVNULL)
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "py2run.py"), os.path.join(ref_src, "generate_data.py"),
           str(n_procs), str(rows), str(cols), str(tmp_path), str(s), str(P), str(partial)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ref_src, env=dict(os.environ, EH_NP_SEED="11"))
    assert p.returncode == 0 and "Done with Test data" in p.stdout, (p.stdout + p.stderr)[-2000:]
    triple = ("1", str(P), "1") if partial else ("1", "0", "1")
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "0", "x", triple[0], str(s), triple[1], triple[2],
            "0", "0", "AGD"]
    cfg = parse_argv(args + ["--gpus=-1", "--quiet", "--seed=1", "--rounds=30", "--init=zeros", "--eta0=0.5"])
    data = load_dataset(cfg.data_dir, cfg.n_partitions, False, cols)
    n_parts = (n_procs - 1) * (P - s) if partial else n_procs - 1
    assert cfg.n_partitions == n_parts and len(data.parts) == n_parts
    assert data.rows_per_part == rows // n_parts and data.n_cols == cols
    assert set(np.unique(data.y)) <= {-1.0, 1.0} and data.X_test.shape == (int(0.2 * rows), cols)
    ev = cli.run(cfg)["evaluation"]
    assert ev.training_loss[-1] < ev.training_loss[0] and ev.auc[-1] > 0.6


@pytest.mark.parametrize("partial,P", [(0, 0), (1, 3)])
def test_generator_files_are_byte_identical_to_the_reference_generator(tmp_path, partial, P):
    """generate_data.py draws from numpy's unseeded global generator; seeded by the harness (EH_NP_SEED) with the
    value this framework's generator is given, both write the same bytes: partitions, labels and test set."""
    import filecmp
    ref_src = os.path.join(ROOT, "baseline", "_ref", "src")
    if not os.path.exists(os.path.join(ref_src, "generate_data.py")):
        subprocess.check_call(["bash", os.path.join(ROOT, "baseline", "install_ref.sh")], stdout=subprocess.DEVNULL)
    ref_root, my_root = str(tmp_path / "ref"), str(tmp_path / "mine")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "py2run.py"), os.path.join(ref_src, "generate_data.py"),
                        "5", "480", "10", ref_root, "1", str(P), str(partial)], capture_output=True, text=True, timeout=300,
                       cwd=ref_src, env=dict(os.environ, EH_NP_SEED="3"))
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    out = generate_gmm_to_disk(5, 480, 10, my_root, 1, P, partial, seed=3)
    ref_out = out.replace(my_root, ref_root)
    files = sorted(os.listdir(out))
    assert files == sorted(os.listdir(ref_out)) and len(files) == (11 if partial else 7)
    for f in files:
        assert filecmp.cmp(out + f, ref_out + f, shallow=False), f


def test_avoidstragg_matches_the_reference_under_a_fixed_arrival_order(tmp_path):
    """The reference's avoidstragg driver has no delay model (avoidstragg.py:121-125): which n - s messages come first
    is left to the operating system.  The MPI stand-in can give every worker's link its own latency
    (EH_MPI_LINK_LATENCY_MS: rank r's messages to the master take r * L ms), which makes the master collect workers
    1 .. n - s in every round -- the set this framework's model collects when no delay is injected (ties go to the
    lower worker id).  With the arrival order pinned, the sum of the first n - s messages and the rescaled step
    (avoidstragg.py:114) must give the same iterates."""
    n_procs, rows, cols, s = 5, 8 * 60, 10, 1
    out = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, 0, 0, seed=3)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "0", "x", "1", str(s), "0", "2", "0", "0", "AGD"]
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "mpirun_shim.py"), "-np", args[0], "main.py"] + args
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, EH_MPI_LINK_LATENCY_MS="4"))
    assert p.returncode == 0 and "Iteration 99: Train Loss" in p.stdout, (p.stdout + p.stderr)[-2000:]
    ref_dir = out + "results/"
    ref_files = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat"))
    prefix = [f for f in ref_files if f.endswith("_training_loss.dat")][0][:-len("_training_loss.dat")]
    ref = {q: io.load_data(ref_dir + f"{prefix}_{q}.dat") for q in ("training_loss", "testing_loss", "auc", "worker_timeset")}
    assert np.all(ref["worker_timeset"][:, :3] >= 0) and np.all(ref["worker_timeset"][:, 3] == -1)    # workers 1..3 every round
    for f in ref_files:
        os.rename(ref_dir + f, ref_dir + "ref_" + f)
    ours = cli.run(parse_argv(args + ["--gpus=-1", "--compat", "--quiet", f"--seed={SEED}"]))
    ev = ours["evaluation"]
    assert np.array_equal(ours["result"].worker_timeset >= 0, ref["worker_timeset"] >= 0)
    assert np.abs(ev.training_loss - ref["training_loss"]).max() <= 1.5e-3
    assert np.abs(ev.testing_loss - ref["testing_loss"]).max() <= 1.5e-3
    assert np.abs(ev.auc - ref["auc"]).max() <= 1.5e-3
    mine = sorted(f for f in os.listdir(ref_dir) if f.endswith(".dat") and not f.startswith("ref_"))
    assert mine == ref_files
