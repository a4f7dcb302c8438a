"""The allocation-free gradient entry point, called through the C ABI.

``eh_grad_dense_noalloc`` exists for hosts that may not allocate device memory once they are
running (fixed-latency serving loops, processes whose HBM is fully committed to resident data,
plain C callers without an allocator): its scratch is a device array built into the library.
This test drives it with ctypes exactly as such a caller would -- inputs and outputs are the
caller's, nothing else is passed -- and checks it against the ordinary entry point
(``eh_grad_dense`` with caller-provided scratch) on two calls in a row, because the built-in
array still holds the first call's partial sums when the second call starts.  It is collected
last on purpose so that a run stopping at the first failure has already run everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _job(X, y, beta, out, d, coeff=1.0, model=0):
    from erasurehead_b200.ops import grad
    return grad.make_dense_job([X], [y], beta.data_ptr(), out.data_ptr(), d, [coeff], model)


def _both(rows, d, seed, model=0, coeff=1.0, dtype=torch.float32):
    from erasurehead_b200.ops import _lib, grad, synth
    L = _lib.lib()
    torch.manual_seed(seed)
    ld = synth.pad_cols(d)
    X = torch.zeros(rows, ld, device="cuda")
    X[:, :d] = torch.randn(rows, d, device="cuda") / d ** 0.5
    X = X.to(dtype)
    y = torch.where(torch.rand(rows, device="cuda") > 0.5, 1.0, -1.0)
    beta = torch.randn(d, device="cuda")
    ref = torch.empty(d, device="cuda")
    got = torch.full((d,), float("nan"), device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    sc = grad.dense_scratch(ld, "cuda")
    _lib.check(L.eh_grad_dense(C.byref(_job(X, y, beta, ref, d, coeff, model)), sc.ptr(), sc.nbytes, stream))
    _lib.check(L.eh_grad_dense_noalloc(C.byref(_job(X, y, beta, got, d, coeff, model)), stream))
    torch.cuda.synchronize()
    return got, ref


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    # two different problems back to back: the second starts with the first one's partials and
    # barrier history still in the built-in scratch
    for rows, d, seed, model, coeff in ((3000, 4096, 1, 0, 1.0), (1777, 1000, 2, 1, -0.75)):
        got, ref = _both(rows, d, seed, model, coeff)
        scale = max(1.0, ref.abs().max().item())
        assert torch.isfinite(got).all()
        assert (got - ref).abs().max().item() <= 2e-4 * scale
    # same problem again: identical reduction order => identical bits
    a, _ = _both(3000, 4096, 1)
    b, _ = _both(3000, 4096, 1)
    assert torch.equal(a, b)


def test_noalloc_narrow_bf16_rows_take_the_tensor_core_kernel_through_the_same_array():
    # bf16 rows with 96 <= ld <= 640 are routed to a different kernel (all-tcgen05, grad_mc_bf16.cu) that
    # keeps its partial sums and barrier words in the same built-in array; then an fp32 call right after
    for rows, d, seed, model in ((4096, 256, 5, 0), (2500, 640, 6, 1)):
        got, ref = _both(rows, d, seed, model, 1.0, torch.bfloat16)
        scale = max(1.0, ref.abs().max().item())
        assert torch.isfinite(got).all()
        assert (got - ref).abs().max().item() <= 2e-4 * scale
    got, ref = _both(3000, 4096, 7)
    assert (got - ref).abs().max().item() <= 2e-4 * max(1.0, ref.abs().max().item())


def test_noalloc_rejects_shapes_beyond_the_builtin_scratch():
    from erasurehead_b200.ops import _lib
    L = _lib.lib()
    assert L.eh_grad_dense_noalloc_max_ld() == 4096
    rows, d = 64, 8192
    X = torch.zeros(rows, d, device="cuda")
    y = torch.ones(rows, device="cuda")
    beta = torch.zeros(d, device="cuda")
    out = torch.zeros(d, device="cuda")
    rc = L.eh_grad_dense_noalloc(C.byref(_job(X, y, beta, out, d)), torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = L.eh_last_error().decode()
    assert "exceeds the built-in scratch limit" in msg and "4096" in msg
