"""`bench.py --impl reference`: run the unmodified reference from baseline/_ref if possible."""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def why_unavailable():
    if not os.path.isdir(REF) or not os.path.exists(os.path.join(REF, "main.py")):
        return ("reference has no setup.py/pyproject (pip install fails) and baseline/_ref is absent; "
                "it is Python-2 + mpi4py code")
    if shutil.which("mpirun") is None and shutil.which("mpiexec") is None:
        return "no MPI runtime (mpirun/mpiexec) in the image and none installable offline"
    try:
        import mpi4py  # noqa: F401
    except Exception:
        return "mpi4py is not installed and not in /opt/wheelhouse"
    if sys.version_info[0] >= 3:
        return "reference is Python-2 only (integer '/' used as index, SURVEY Q10); no python2 in the image"
    return None


def run_reference(args):
    reason = why_unavailable()
    return {"impl": "reference", "unavailable": reason or "not implemented"}
