"""Worker-side gradient operators (K1+K2'+K3 fused).

``grad_dense`` / ``grad_ell`` compute one encoded gradient message and write it straight into
its destination (a local tensor here; the engine passes peer-mapped slot pointers instead).
Reference math: naive.py:137-139, coded.py:92-95,184 (encode fold), naive.py:345-346 (linear).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import GradEllJob, GradJob, check, lib, stream_ptr


class Scratch:
    """Zero-initialised device scratch shared by the cooperative kernels of one stream."""

    def __init__(self, nbytes: int, device):
        self.t = torch.zeros((nbytes + 7) // 8, dtype=torch.int64, device=device)
        self.nbytes = self.t.numel() * 8

    def ptr(self) -> int:
        return self.t.data_ptr()

    def zero_(self):
        self.t.zero_()


def dense_scratch(d: int, device) -> Scratch:
    return Scratch(lib().eh_grad_dense_scratch_bytes(int(d)), device)


def ell_scratch(d: int, device) -> Scratch:
    return Scratch(lib().eh_grad_ell_scratch_bytes(int(d)), device)


def _segments(job, Xs, ys, coeff):
    n = len(Xs)
    if not (1 <= n <= _lib.EH_MAX_SEGS) or len(ys) != n or len(coeff) != n:
        raise ValueError(f"need 1..{_lib.EH_MAX_SEGS} segments with matching labels/coefficients")
    for i in range(n):
        job.d_X[i] = Xs[i].data_ptr()
        job.d_y[i] = ys[i].data_ptr()
        job.coeff[i] = float(coeff[i])
    job.n_segs = n


def make_dense_job(Xs: Sequence[torch.Tensor], ys: Sequence[torch.Tensor], beta_ptr: int, out_ptr: int, d: int,
                   coeff: Sequence[float], model: int, stamp: int = 0, flag_ptr: int = 0,
                   beta_flag_ptr: int = 0, err_ptr: int = 0, timeout_ns: int = 0) -> GradJob:
    """One message = the listed partitions (equal row counts, same row stride)."""
    X0 = Xs[0]
    for X, y in zip(Xs, ys):
        assert X.dtype in (torch.float32, torch.bfloat16) and X.dtype == X0.dtype and X.dim() == 2 and X.stride(1) == 1
        assert X.shape[0] == X0.shape[0] and X.stride(0) == X0.stride(0)
        assert y.dtype == torch.float32 and y.is_contiguous() and y.numel() == X.shape[0]
    job = GradJob()
    _segments(job, Xs, ys, coeff)
    job.d_beta, job.d_out = beta_ptr, out_ptr
    job.d_flag, job.d_beta_flag, job.d_err = flag_ptr, beta_flag_ptr, err_ptr
    job.rows_per_seg, job.d, job.ld = X0.shape[0], int(d), X0.stride(0)
    job.model, job.stamp, job.timeout_ns = int(model), int(stamp), int(timeout_ns)
    job.x_bf16 = int(X0.dtype == torch.bfloat16)
    return job


def launch_dense(job: GradJob, scratch: Scratch, stream: Optional[int] = None) -> None:
    check(lib().eh_grad_dense(C.byref(job), scratch.ptr(), scratch.nbytes,
                              stream_ptr() if stream is None else stream), "eh_grad_dense")


def launch_dense_multi(jobs: Sequence[GradJob], scratch: Scratch, stream: Optional[int] = None) -> None:
    """Up to 8 messages of one shape in a single launch (``eh_grad_dense_multi``)."""
    arr = (GradJob * len(jobs))(*jobs)
    check(lib().eh_grad_dense_multi(arr, len(jobs), scratch.ptr(), scratch.nbytes,
                                    stream_ptr() if stream is None else stream), "eh_grad_dense_multi")


def make_ell_job(idxs: Sequence[torch.Tensor], ys: Sequence[torch.Tensor], beta_ptr: int, out_ptr: int, d: int,
                 coeff: Sequence[float], model: int, bias_col: int = -1, stamp: int = 0, flag_ptr: int = 0,
                 beta_flag_ptr: int = 0, err_ptr: int = 0, timeout_ns: int = 0) -> GradEllJob:
    i0 = idxs[0]
    for idx, y in zip(idxs, ys):
        assert idx.dtype == torch.int32 and idx.dim() == 2 and idx.is_contiguous() and idx.shape == i0.shape
        assert y.dtype == torch.float32 and y.is_contiguous() and y.numel() == idx.shape[0]
    job = GradEllJob()
    n = len(idxs)
    if not (1 <= n <= _lib.EH_MAX_SEGS):
        raise ValueError("too many segments")
    for i in range(n):
        job.d_idx[i] = idxs[i].data_ptr()
        job.d_y[i] = ys[i].data_ptr()
        job.coeff[i] = float(coeff[i])
    job.n_segs = n
    job.d_beta, job.d_out = beta_ptr, out_ptr
    job.d_flag, job.d_beta_flag, job.d_err = flag_ptr, beta_flag_ptr, err_ptr
    job.rows_per_seg, job.d, job.nnz = i0.shape[0], int(d), i0.shape[1]
    job.model, job.bias_col, job.stamp, job.timeout_ns = int(model), int(bias_col), int(stamp), int(timeout_ns)
    return job


def launch_ell(job: GradEllJob, scratch: Scratch, stream: Optional[int] = None) -> None:
    check(lib().eh_grad_ell(C.byref(job), scratch.ptr(), scratch.nbytes,
                            stream_ptr() if stream is None else stream), "eh_grad_ell")


def launch_ell_multi(jobs: Sequence[GradEllJob], scratch: Scratch, stream: Optional[int] = None) -> None:
    arr = (GradEllJob * len(jobs))(*jobs)
    check(lib().eh_grad_ell_multi(arr, len(jobs), scratch.ptr(), scratch.nbytes,
                                  stream_ptr() if stream is None else stream), "eh_grad_ell_multi")


def _aslist(x):
    return list(x) if isinstance(x, (list, tuple)) else [x]


def grad_dense(X, y, beta: torch.Tensor, coeff: Sequence[float] = (1.0,), model: int = 0,
               d: Optional[int] = None, scratch: Optional[Scratch] = None,
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Standalone call: returns ``g`` (float32 ``[d]``).  ``X``/``y`` may be lists of equally
    shaped partitions (one encode coefficient each); ``X`` may have a padded row stride."""
    Xs, ys = _aslist(X), _aslist(y)
    d = int(d or Xs[0].shape[1])
    out = torch.empty(d, dtype=torch.float32, device=Xs[0].device) if out is None else out
    scratch = scratch or dense_scratch(Xs[0].stride(0), Xs[0].device)
    job = make_dense_job(Xs, ys, beta.data_ptr(), out.data_ptr(), d, list(coeff), model)
    launch_dense(job, scratch)
    return out


def grad_ell(idx, y, beta: torch.Tensor, coeff: Sequence[float] = (1.0,), model: int = 0, bias_col: int = -1,
             scratch: Optional[Scratch] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    idxs, ys = _aslist(idx), _aslist(y)
    d = beta.numel()
    out = torch.empty(d, dtype=torch.float32, device=idxs[0].device) if out is None else out
    scratch = scratch or ell_scratch(d, idxs[0].device)
    job = make_ell_job(idxs, ys, beta.data_ptr(), out.data_ptr(), d, list(coeff), model, bias_col)
    launch_ell(job, scratch)
    return out


def grad_ell_mc(idx, y, W: torch.Tensor, classes: int, coeff: Sequence[float] = (1.0,), model: int = 2) -> torch.Tensor:
    """Standalone multi-column gradient on one-hot features: ``W`` and the result are ``[d, 8]`` float32
    (column-major, ``classes <= 8`` used; csrc/grad_ell_mc.cu).  Labels: class ids (model 2) or +-1 (model 3)."""
    idxs, ys = _aslist(idx), _aslist(y)
    d = W.shape[0]
    assert W.dtype == torch.float32 and W.is_contiguous() and W.shape[1] == lib().eh_grad_ell_mc_padded_classes()
    out = torch.empty_like(W)
    sc = Scratch(lib().eh_grad_ell_mc_scratch_bytes(d), idxs[0].device)
    job = make_ell_job(idxs, ys, W.data_ptr(), out.data_ptr(), d, list(coeff), model, -1)
    _lib.check(lib().eh_grad_ell_mc_multi(C.byref(job), 1, int(classes), sc.ptr(), sc.nbytes, _lib.stream_ptr()),
               "eh_grad_ell_mc_multi")
    return out
