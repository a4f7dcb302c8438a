"""Parity against the reference's own code: the patched copy (baseline/install_ref.sh: integer
division fixes only) runs over the in-repo mpi4py subset on the same data files, and its result
files are compared with ours.  Schemes whose trajectory does not depend on the unseeded RNG or on
arrival order are used: partial_replication and partial_coded (zero init, exact decode, AGD)."""
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


def run_reference(args, timeout=300):
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "mpirun_shim.py"), "-np", args[0], "main.py"] + args
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    return p.stdout


@pytest.mark.parametrize("coded_ver,prefix", [(1, "partialreplication_1_3"), (0, "partialcoded_1_3")])
def test_partial_schemes_match_the_reference_files(tmp_path, coded_ver, prefix):
    n_procs, rows, cols, s, P = 5, 8 * 60, 10, 1, 3
    out = generate_gmm_to_disk(n_procs, rows, cols, str(tmp_path), s, P, 1, seed=3)
    args = [str(n_procs), str(rows), str(cols), str(tmp_path) + "/", "0", "x", "1", str(s), str(P), str(coded_ver),
            "0", "0", "AGD"]
    txt = run_reference(args)
    assert "Iteration 99: Train Loss" in txt
    ref_dir = out + "results/"
    ref_files = sorted(os.listdir(ref_dir))
    ref_test = io.load_data(ref_dir + prefix + "_testing_loss.dat")
    ref_auc = io.load_data(ref_dir + prefix + "_auc.dat")
    tl_prefix = "partialreplication_1_3"            # Q3: partial_coded writes its training loss under this prefix
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
