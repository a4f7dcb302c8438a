#!/usr/bin/env python
"""`mpirun -np N python main.py ...` stand-in for baseline B1: starts N local ranks of the patched
reference (baseline/_ref_py3) with the mpi4py subset in baseline/mpi_shim on PYTHONPATH.

    python baseline/mpirun_shim.py -np 9 main.py 9 7680 128 /tmp/data/ 0 x 1 1 0 3 6 0 AGD
"""
import os
import random
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    argv = sys.argv[1:]
    assert argv[0] == "-np"
    n = int(argv[1])
    rest = argv[2:]
    ref = os.path.join(HERE, "_ref_py3")
    if not os.path.exists(os.path.join(ref, "main.py")):
        subprocess.check_call(["bash", os.path.join(HERE, "install_ref.sh")])
    port0 = 30000 + random.randint(0, 20000)
    procs = []
    for r in range(n):
        env = dict(os.environ, EH_MPI_RANK=str(r), EH_MPI_SIZE=str(n), EH_MPI_PORT0=str(port0),
                   PYTHONPATH=os.path.join(HERE, "mpi_shim") + os.pathsep + os.environ.get("PYTHONPATH", ""),
                   OMP_NUM_THREADS=os.environ.get("OMP_NUM_THREADS", "1"), PYTHONWARNINGS="ignore")
        out = None if r == 0 else subprocess.DEVNULL
        procs.append(subprocess.Popen([sys.executable] + rest, cwd=ref, env=env, stdout=out, stderr=out))
    rc = 0
    for p in procs:
        rc |= p.wait()
    return rc


if __name__ == "__main__":
    sys.exit(main())
