#!/usr/bin/env python
"""`mpirun -np N python main.py ...` stand-in for baseline B1 (no MPI runtime exists in the image).

Starts N local ranks of the UNMODIFIED reference (baseline/_ref, see install_ref.sh); each rank is
``python baseline/py2run.py main.py <args>`` (Python-2 division semantics at import time) with the
mpi4py subset of baseline/mpi_shim first on PYTHONPATH.

    python baseline/mpirun_shim.py -np 9 main.py 9 7680 128 /tmp/data/ 0 x 1 1 0 3 6 0 AGD

Rank 0's stdout/stderr pass through; the other ranks are silenced unless EH_MPI_VERBOSE is set.
"""
import os
import random
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def _free_port_range(n):
    """A base port such that port0 .. port0+n-1 can all be bound right now (one listening port per rank)."""
    import socket
    for _ in range(200):
        port0 = 30000 + random.randint(0, 20000)
        socks = []
        try:
            for r in range(n):
                sk = socket.socket()
                sk.bind(("127.0.0.1", port0 + r))
                socks.append(sk)
            return port0
        except OSError:
            continue
        finally:
            for sk in socks:
                sk.close()
    raise RuntimeError("no free range of %d local ports found" % n)


def launch(n, rest, stdout=None, env_extra=None):
    """Start the ranks; returns the list of Popen objects (rank 0 first)."""
    if not os.path.exists(os.path.join(REF, "main.py")):
        subprocess.check_call(["bash", os.path.join(HERE, "install_ref.sh")])
    port0 = _free_port_range(n)
    procs = []
    for r in range(n):
        env = dict(os.environ, EH_MPI_RANK=str(r), EH_MPI_SIZE=str(n), EH_MPI_PORT0=str(port0),
                   PYTHONPATH=os.path.join(HERE, "mpi_shim") + os.pathsep + os.environ.get("PYTHONPATH", ""),
                   OMP_NUM_THREADS=os.environ.get("OMP_NUM_THREADS", "1"), PYTHONWARNINGS="ignore",
                   CUDA_VISIBLE_DEVICES="")
        env.update(env_extra or {})
        quiet = r != 0 and not os.environ.get("EH_MPI_VERBOSE")
        out = subprocess.DEVNULL if quiet else stdout
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "py2run.py")] + rest, cwd=REF, env=env,
                                      stdout=out, stderr=subprocess.STDOUT if (out is not None and not quiet) else out))
    return procs


def main():
    argv = sys.argv[1:]
    assert argv[0] == "-np"
    n = int(argv[1])
    procs = launch(n, argv[2:])
    rc = 0
    for p in procs:
        rc |= p.wait()
    return rc


if __name__ == "__main__":
    sys.exit(main())
