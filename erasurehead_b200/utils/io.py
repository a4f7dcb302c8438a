"""File formats of the reference, byte-compatible.

* dense partition ``k.dat``: one row per line, space-separated ``str(float)`` (util.py:26-30),
  read back with ``np.loadtxt`` (util.py:13-15);
* sparse partition ``k.npz``: CSR arrays under ``data, indices, indptr, shape`` (util.py:17-24);
* vectors (labels, per-round losses/times): one ``"%5.3f "`` per line (util.py:32-36).
"""
from __future__ import annotations

import os

import numpy as np
from scipy.sparse import csr_matrix

__all__ = ["load_data", "save_matrix", "save_vector", "save_sparse_csr", "load_sparse_csr",
           "ensure_dir"]


def ensure_dir(path: str) -> str:
    os.makedirs(path, exist_ok=True)
    return path


def load_data(path: str, dtype=float) -> np.ndarray:
    """Dense text matrix or vector (util.py:13-15).  Files of more than a few thousand values go through the
    native multi-threaded reader of liberasurehead (csrc/textio.cu, ~50x np.loadtxt); small files, and machines
    where the library cannot be built, use np.loadtxt.  Same result either way (correctly rounded decimals)."""
    try:
        if os.path.getsize(path) >= (1 << 16):
            return _load_data_native(path, dtype)
    except (OSError, ImportError, RuntimeError):
        pass
    return np.loadtxt(path, dtype=dtype)


def _load_data_native(path: str, dtype=float) -> np.ndarray:
    import ctypes as C
    from ..ops import _lib
    L = _lib.lib()
    rows, cols = C.c_int64(0), C.c_int64(0)
    _lib.check(L.eh_text_shape(path.encode(), C.byref(rows), C.byref(cols)), "eh_text_shape")
    dt = np.dtype(dtype)
    if dt not in (np.dtype(np.float32), np.dtype(np.float64)) or rows.value == 0 or cols.value == 0:
        raise RuntimeError("native reader: unsupported dtype or empty file")
    out = np.empty((rows.value, cols.value), dtype=dt)
    _lib.check(L.eh_load_dense_text(path.encode(), rows.value, cols.value, out.ctypes.data, cols.value, dt.itemsize),
               "eh_load_dense_text")
    # np.loadtxt squeezes: a one-column file is a vector, a one-row file is a vector
    return out.reshape(-1) if 1 in out.shape else out


def save_matrix(m, path: str) -> None:
    """One row per line, space-separated ``str(float)`` (util.py:26-30).  Matrices of more than a few thousand
    values go through the native multi-threaded writer (csrc/textio.cu: shortest round-trip digits laid out by
    Python's repr rule, byte-identical output, ~100x this loop); small ones, and machines where the library
    cannot be built, use the loop."""
    a = np.asarray(m)
    if a.ndim == 2 and a.size >= (1 << 12) and a.dtype.kind == "f":
        try:
            return _save_matrix_native(a, path)
        except (OSError, ImportError, RuntimeError):
            pass
    with open(path, "w") as f:
        for row in m:
            f.write(" ".join(repr(float(x)) for x in row) + "\n")


def _save_matrix_native(a: np.ndarray, path: str) -> None:
    from ..ops import _lib
    L = _lib.lib()
    if not hasattr(L, "eh_save_dense_text"):
        raise RuntimeError("native writer not in this build")
    a = np.ascontiguousarray(a, dtype=np.float64)                 # float(x) of the loop: exact widening of float32
    _lib.check(L.eh_save_dense_text(path.encode(), a.ctypes.data, a.shape[0], a.shape[1], a.shape[1]),
               "eh_save_dense_text")


def save_vector(v, path: str) -> None:
    with open(path, "w") as f:
        for x in v:
            f.write("%5.3f" % x + " \n")


def save_sparse_csr(path: str, array) -> None:
    array = csr_matrix(array)
    np.savez(path, data=array.data, indices=array.indices, indptr=array.indptr, shape=array.shape)


def load_sparse_csr(path: str) -> csr_matrix:
    loader = np.load(path if path.endswith(".npz") else path + ".npz")
    return csr_matrix((loader["data"], loader["indices"], loader["indptr"]), shape=loader["shape"])
