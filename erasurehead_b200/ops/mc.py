"""Multi-column (softmax / batched-logistic) gradient on the tensor cores (``csrc/grad_mc_tc.cu``).

Layout: parameters and gradients are class-major ``W[CN, ldw]`` (``CN`` = 16 or 32 padded
classes, ``ldw`` = features padded to 32) -- one flat vector for the decoder."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from ._lib import McJob, check, lib, stream_ptr

MODEL_SOFTMAX = 2
MODEL_MULTI_LOGISTIC = 3


def padded_classes(C_: int) -> int:
    if not 1 <= C_ <= 32:
        raise ValueError("1..32 columns supported")
    return 16 if C_ <= 16 else 32


def pad32(d: int) -> int:
    return (d + 31) // 32 * 32


def flat_size(d: int, C_: int) -> int:
    return padded_classes(C_) * pad32(d)


def pack_params(W: torch.Tensor) -> torch.Tensor:
    """``[d, C]`` -> class-major padded ``[CN, ldw]`` float32."""
    d, C_ = W.shape
    out = torch.zeros((padded_classes(C_), pad32(d)), dtype=torch.float32, device=W.device)
    out[:C_, :d] = W.t().float()
    return out


def unpack_params(Wp: torch.Tensor, d: int, C_: int) -> torch.Tensor:
    return Wp.reshape(padded_classes(C_), pad32(d))[:C_, :d].t().contiguous()


def scratch(rows_per_seg: int, d: int, C_: int, device) -> torch.Tensor:
    n = lib().eh_grad_mc_scratch_bytes(int(rows_per_seg), int(d), int(C_))
    return torch.zeros((n + 7) // 8, dtype=torch.int64, device=device)


def make_job(Xs: Sequence[torch.Tensor], ys: Sequence[torch.Tensor], W_ptr: int, out_ptr: int, d: int, C_: int,
             coeff: Sequence[float], model: int, loss_ptr: int = 0, stamp: int = 0, flag_ptr: int = 0,
             beta_flag_ptr: int = 0, err_ptr: int = 0, timeout_ns: int = 0) -> McJob:
    job = McJob()
    X0 = Xs[0]
    bf16 = X0.dtype == torch.bfloat16          # bf16 rows: all-tcgen05 one-pass kernel (csrc/grad_mc_bf16.cu)
    for i, (X, y) in enumerate(zip(Xs, ys)):
        assert X.dtype == X0.dtype and X.dtype in (torch.float32, torch.bfloat16)
        assert X.stride(1) == 1 and X.stride(0) == X0.stride(0) and X.shape[0] == X0.shape[0]
        assert y.dtype == torch.float32 and y.is_contiguous()
        job.d_X[i], job.d_y[i], job.coeff[i] = X.data_ptr(), y.data_ptr(), float(coeff[i])
    job.n_segs, job.rows_per_seg = len(Xs), X0.shape[0]
    job.d, job.ld, job.ldw, job.C, job.model = int(d), X0.stride(0), pad32(d), int(C_), int(model)
    job.d_W, job.d_out, job.d_loss = W_ptr, out_ptr, loss_ptr
    job.d_flag, job.d_beta_flag, job.d_err = flag_ptr, beta_flag_ptr, err_ptr
    job.stamp, job.timeout_ns = int(stamp), int(timeout_ns)
    job.x_bf16 = 1 if bf16 else 0
    return job


def launch(job: McJob, scr: torch.Tensor, stream: Optional[int] = None) -> None:
    check(lib().eh_grad_mc(C.byref(job), scr.data_ptr(), scr.numel() * 8, stream_ptr() if stream is None else stream),
          "eh_grad_mc")


def grad_mc(X, y, Wp: torch.Tensor, d: int, C_: int, coeff=(1.0,), model: int = MODEL_SOFTMAX,
            want_loss: bool = False):
    """Standalone: returns ``(G[CN, ldw] float32, loss_sum float64 0-d or None)``."""
    Xs = list(X) if isinstance(X, (list, tuple)) else [X]
    ys = list(y) if isinstance(y, (list, tuple)) else [y]
    G = torch.empty((padded_classes(C_), pad32(d)), dtype=torch.float32, device=Xs[0].device)
    loss = torch.zeros((), dtype=torch.float64, device=Xs[0].device) if want_loss else None
    scr = scratch(Xs[0].shape[0], d, C_, Xs[0].device)
    job = make_job(Xs, ys, Wp.data_ptr(), G.data_ptr(), d, C_, list(coeff), model,
                   loss_ptr=0 if loss is None else loss.data_ptr())
    launch(job, scr)
    return G, loss


def loss_mc(X, y, Wp: torch.Tensor, d: int, C_: int, model: int = MODEL_SOFTMAX) -> torch.Tensor:
    """Summed loss only (forward product + epilogue)."""
    Xs = list(X) if isinstance(X, (list, tuple)) else [X]
    ys = list(y) if isinstance(y, (list, tuple)) else [y]
    loss = torch.zeros((), dtype=torch.float64, device=Xs[0].device)
    scr = scratch(Xs[0].shape[0], d, C_, Xs[0].device)
    job = make_job(Xs, ys, Wp.data_ptr(), 0, d, C_, [1.0] * len(Xs), model, loss_ptr=loss.data_ptr())
    launch(job, scr)
    return loss
