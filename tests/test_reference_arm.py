"""The reference arm of bench.py: the unmodified reference runs through its own CLI."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
REF_OK = os.path.exists(os.path.join(ROOT, "baseline", "_ref", "main.py")) or os.path.exists("/root/reference/main.py")
pytestmark = pytest.mark.skipif(not REF_OK, reason="no reference tree (baseline/_ref or /root/reference)")


def test_installed_reference_is_byte_identical():
    if not os.path.exists("/root/reference/main.py"):
        pytest.skip("reference not mounted")
    subprocess.check_call(["bash", os.path.join(ROOT, "baseline", "install_ref.sh")], stdout=subprocess.DEVNULL)
    for rel in ["main.py"] + ["src/" + f for f in os.listdir("/root/reference/src") if f.endswith(".py")]:
        a = open(os.path.join("/root/reference", rel), "rb").read()
        b = open(os.path.join(ROOT, "baseline", "_ref", rel), "rb").read()
        assert a == b, rel


def test_py2_division_semantics():
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import numpy as np
    import py2run
    g = {}
    exec(py2run.compile_py2("a = 7 / 2\nb = 7.0 / 2\nc = n / 4\nd = x / 2\ne = 9\ne /= 2\n", "<t>"),
         {"n": np.int64(30), "x": np.array([3, 5]), "__builtins__": __builtins__}, g)
    assert g["a"] == 3 and g["b"] == 3.5 and g["c"] == 7 and list(g["d"]) == [1, 2] and g["e"] == 4


def test_reference_arm_prints_a_number(tmp_path):
    """`bench.py --impl reference` on a tiny shape (the stock 100 rounds and 0.5 s sleeps make the
    delayed run last about a minute)."""
    env = dict(os.environ, TMPDIR=str(tmp_path))
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--cols", "64",
                        "--ref-rows", "64", "--no-covtype"], capture_output=True, text=True, timeout=900, env=env)
    assert p.returncode == 0, p.stderr[-2000:]
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" not in line, line
    assert line["value"] > 0 and line["ms_per_step"] > 50          # min of two Exp(0.5 s) sleeps per round
    assert line["no_delay"]["ms_per_step"] < line["ms_per_step"]
    assert line["e2e"]["value"] > 0 and line["gpu_launches"] == 0
    assert line["config"]["n_workers"] == 2 and line["same_config"] is False
