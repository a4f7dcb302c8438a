"""Numerics of every CUDA kernel against a plain PyTorch / numpy float32-64 reference."""
import itertools

import numpy as np
import pytest
import torch

from erasurehead_b200 import codes
from erasurehead_b200.utils import metrics as hmetrics

pytestmark = pytest.mark.gpu
DEV = "cuda"


def ref_grad(X, y, beta, coeff, model):
    """fp64 torch reference of the encoded gradient (naive.py:137-139 / 345-346, coded.py:184)."""
    X, y, beta = X.double(), y.double(), beta.double()
    p = X @ beta
    r = -y / (torch.exp(p * y) + 1) if model == 0 else -2 * (y - p)
    return X.t() @ (coeff * r)


@pytest.mark.parametrize("rows,d", [(64, 16), (1000, 100), (517, 1024), (300, 1500), (1024, 4096), (96, 8000),
                                    (700, 20000), (500, 12288), (333, 32768), (150, 40000)])
@pytest.mark.parametrize("model", [0, 1])
def test_grad_dense_matches_reference(rows, d, model):
    from erasurehead_b200.ops import grad, synth
    torch.manual_seed(rows + d)
    ld = synth.pad_cols(d)
    X = torch.zeros(rows, ld, device=DEV)
    X[:, :d] = torch.randn(rows, d, device=DEV) / np.sqrt(d)
    y = torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0) if model == 0 else torch.randn(rows, device=DEV)
    beta = torch.randn(d, device=DEV)
    g = grad.grad_dense(X, y, beta, coeff=(1.7,), model=model, d=d)
    want = ref_grad(X[:, :d], y, beta, 1.7, model)
    err = (g.double() - want).abs().max().item()
    assert err <= 2e-4 * max(1.0, want.abs().max().item()), err


def test_grad_dense_segments_and_repeat_calls():
    """Several partitions with different encode weights in one message; scratch reuse across calls."""
    from erasurehead_b200.ops import grad
    torch.manual_seed(0)
    rows, d = 130, 256
    Xs = [torch.randn(rows, d, device=DEV) * 0.1 for _ in range(3)]
    ys = [torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0) for _ in range(3)]
    cf = [1.0, -2.5, 0.75]
    beta = torch.randn(d, device=DEV)
    sc = grad.dense_scratch(d, DEV)
    want = sum(ref_grad(X, y, beta, c, 0) for X, y, c in zip(Xs, ys, cf))
    for _ in range(3):
        g = grad.grad_dense(Xs, ys, beta, coeff=cf, model=0, scratch=sc)
        assert (g.double() - want).abs().max().item() <= 2e-4 * max(1.0, want.abs().max().item())
    g2 = grad.grad_dense(Xs, ys, beta, coeff=cf, model=0, scratch=sc)
    assert torch.equal(g, g2)            # fixed-order reduction => bitwise reproducible


@pytest.mark.parametrize("model", [0, 1])
def test_grad_ell_matches_reference(model):
    from erasurehead_b200.data import generate_onehot
    from erasurehead_b200.ops import grad
    ds = generate_onehot(4 * 700, 5000, 19, 4, seed=1, task="linear" if model else "logistic")
    idxs = [torch.from_numpy(ds.ell(p)).to(DEV) for p in range(2)]
    ys = [torch.tensor(ds.labels(p), dtype=torch.float32, device=DEV) for p in range(2)]
    beta = torch.randn(5000, device=DEV) * 0.3
    cf = [1.0, -0.5]
    g = grad.grad_ell(idxs, ys, beta, coeff=cf, model=model, bias_col=4999)
    want = sum(ref_grad(torch.tensor(ds.parts[p].toarray(), device=DEV), ys[p], beta, cf[p], model) for p in range(2))
    assert (g.double() - want).abs().max().item() <= 5e-4 * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("model", [0, 1])
def test_grad_ell_fixed_point_scatter_is_exactly_reproducible(model):
    """Covertype-sized feature space: the shared-memory scatter adds fixed-point integers (native
    ATOMS.ADD), so two runs agree bit for bit whatever order the warps ran in, and the result is at
    least as close to float64 as an fp32 accumulation (large and small encode weights together)."""
    from erasurehead_b200.data import generate_onehot
    from erasurehead_b200.ops import grad
    d, nnz = 15509, 55
    ds = generate_onehot(3 * 9000, d, nnz, 3, seed=4, task="linear" if model else "logistic")
    idxs = [torch.from_numpy(ds.ell(p)).to(DEV) for p in range(3)]
    ys = [torch.tensor(ds.labels(p), dtype=torch.float32, device=DEV) for p in range(3)]
    beta = torch.randn(d, device=DEV) * 0.2
    cf = [1.0, -37.5, 0.003]
    g1 = grad.grad_ell(idxs, ys, beta, coeff=cf, model=model, bias_col=d - 1)
    g2 = grad.grad_ell(idxs, ys, beta, coeff=cf, model=model, bias_col=d - 1)
    assert torch.equal(g1, g2)
    Xs = [torch.sparse_csr_tensor(torch.arange(0, i.numel() + 1, nnz, device=DEV), i.reshape(-1).long(),
                                  torch.ones(i.numel(), dtype=torch.float64, device=DEV), size=(i.shape[0], d)) for i in idxs]
    want = torch.zeros(d, dtype=torch.float64, device=DEV)
    for X, y, c in zip(Xs, ys, cf):
        pm = torch.mv(X, beta.double())
        r = -y.double() / (torch.exp(pm * y.double()) + 1) if model == 0 else -2 * (y.double() - pm)
        want += c * torch.mv(X.t().to_sparse_csr(), r)
    assert (g1.double() - want).abs().max().item() <= 2e-5 * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("d", [5000, 70000])
def test_grad_ell_multi_batches_messages_in_one_launch(d):
    """Several one-hot messages (different partition counts / coefficients) in one launch; d = 70000 does
    not fit in shared memory and takes the per-message fallback."""
    from erasurehead_b200.data import generate_onehot
    from erasurehead_b200.ops import grad
    ds = generate_onehot(5 * 600, d, 19, 5, seed=2)
    idxs = [torch.from_numpy(ds.ell(p)).to(DEV) for p in range(5)]
    ys = [torch.tensor(ds.labels(p), dtype=torch.float32, device=DEV) for p in range(5)]
    beta = torch.randn(d, device=DEV) * 0.3
    msgs = [([0], [1.0]), ([1, 2, 3], [0.5, -2.0, 1.5]), ([4, 0], [1.0, 1.0])]
    outs = [torch.full((d,), float("nan"), device=DEV) for _ in msgs]
    sc = grad.ell_scratch(d, DEV)
    jobs = [grad.make_ell_job([idxs[p] for p in ps], [ys[p] for p in ps], beta.data_ptr(), o.data_ptr(), d, cf, 0, d - 1)
            for (ps, cf), o in zip(msgs, outs)]
    grad.launch_ell_multi(jobs, sc)
    torch.cuda.synchronize()
    for (ps, cf), o in zip(msgs, outs):
        want = sum(ref_grad(torch.tensor(ds.parts[p].toarray(), device=DEV), ys[p], beta, c, 0) for p, c in zip(ps, cf))
        assert (o.double() - want).abs().max().item() <= 5e-4 * max(1.0, want.abs().max().item())


def test_margins_loss_and_auc():
    from erasurehead_b200.ops import metrics as mops
    torch.manual_seed(1)
    rows, d, R = 3000, 200, 37
    X = torch.randn(rows, d, device=DEV)
    Bt = torch.randn(R, d, device=DEV) * 0.1
    y = torch.where(torch.rand(rows, device=DEV) > 0.4, 1.0, -1.0)
    P = mops.margins_dense(X, Bt)
    want = X.double() @ Bt.double().t()
    assert (P.double() - want).abs().max().item() < 1e-3
    L = mops.loss_columns(P, y, 0).cpu().numpy()
    M = mops.loss_columns(P, y, 1).cpu().numpy()
    Pn, yn = want.cpu().numpy(), y.cpu().numpy().astype(np.float64)
    for c in (0, 5, R - 1):
        assert L[c] == pytest.approx(hmetrics.logistic_loss(yn, Pn[:, c]), rel=1e-4)
        assert M[c] == pytest.approx(hmetrics.mse(yn, Pn[:, c]), rel=1e-4)
    s = P[:, 3].contiguous()
    assert mops.auc(s, y).item() == pytest.approx(hmetrics.roc_auc(yn, s.cpu().numpy()), abs=1e-9)
    st = torch.round(s * 4) / 4                     # heavy ties
    assert mops.auc(st, y).item() == pytest.approx(hmetrics.roc_auc(yn, st.cpu().numpy()), abs=1e-9)


def test_margins_ell():
    from erasurehead_b200.data import generate_onehot
    from erasurehead_b200.ops import metrics as mops
    ds = generate_onehot(500, 900, 11, 1, seed=4)
    idx = torch.from_numpy(ds.ell(0)).to(DEV)
    Bt = torch.randn(9, 900, device=DEV)
    P = mops.margins_ell(idx, Bt)
    want = torch.tensor(ds.parts[0].toarray(), device=DEV).double() @ Bt.double().t()
    assert (P.double() - want).abs().max().item() < 1e-4


def test_synth_gmm_statistics_and_determinism():
    from erasurehead_b200.ops import synth
    rows, d = 20000, 64
    X, y = synth.synth_gmm(rows, d, seed=7, partition=3)
    X2, y2 = synth.synth_gmm(rows, d, seed=7, partition=3)
    assert torch.equal(X, X2) and torch.equal(y, y2)
    X3, _ = synth.synth_gmm(rows, d, seed=7, partition=4)
    assert not torch.equal(X, X3)
    assert set(y.unique().tolist()) == {-1.0, 1.0}
    sd = X[:, :d].std().item()
    assert sd == pytest.approx(10 / np.sqrt(d), rel=0.05)            # util.py:41 mfac
    assert abs(X[:, :d].mean().item()) < 0.05
    # labels carry signal: a few GD steps on the generated data beat chance
    from erasurehead_b200.ops import grad
    beta = torch.zeros(d, device=DEV)
    for _ in range(30):
        beta -= (1.0 / rows) * grad.grad_dense(X, y, beta, d=d)
    acc = ((X[:, :d] @ beta) * y > 0).float().mean().item()
    assert acc > 0.9


@pytest.mark.parametrize("n,s", [(8, 2), (12, 3), (30, 3)])
def test_codegen_matches_host(n, s):
    from erasurehead_b200.ops import codegen
    H = codes.parity_check(n, s, seed=2)
    B_host = codes.cyclic_code(n, s, H=H)
    B = codegen.cyclic_code(torch.from_numpy(H), n, s)
    assert np.abs(B.cpu().numpy() - B_host).max() < 1e-8 * max(1.0, np.abs(B_host).max())
    if n <= 12:
        A = codegen.decode_table(B, s).cpu().numpy()
        assert A.shape[0] == len(list(itertools.combinations(range(n), s)))
        assert np.abs(A @ B_host - 1).max() < 1e-6
        for r, missing in enumerate(itertools.combinations(range(n), s)):
            assert np.all(A[r, list(missing)] == 0)


@pytest.mark.parametrize("rows,d,R", [(128, 32, 16), (1000, 200, 37), (5000, 4096, 100), (777, 96, 150)])
@pytest.mark.parametrize("kind", [0, 1])
def test_eval_tc_tensor_core_margins_and_loss(rows, d, R, kind):
    """tcgen05/TMEM evaluation vs the fp64 reference (TF32 operands: ~1e-3 relative on margins)."""
    from erasurehead_b200.ops import metrics as mops, synth
    torch.manual_seed(rows + R)
    ld = synth.pad_cols(d)
    X = torch.zeros(rows, ld, device=DEV)
    X[:, :d] = torch.randn(rows, d, device=DEV)
    Bt = torch.zeros(R, ld, device=DEV)
    Bt[:, :d] = torch.randn(R, d, device=DEV) / np.sqrt(d)
    y = torch.where(torch.rand(rows, device=DEV) > 0.4, 1.0, -1.0) if kind == 0 else torch.randn(rows, device=DEV)
    sums, P = mops.eval_tc(X, Bt, y, kind=kind, d=d, want_margins=True)
    want = X[:, :d].double() @ Bt[:, :d].double().t()
    assert (P.double() - want).abs().max().item() < 8e-3
    if kind == 0:
        ref = torch.nn.functional.softplus(-y.double()[:, None] * want).sum(0)
    else:
        ref = ((y.double()[:, None] - want) ** 2).sum(0)
    rel = ((sums - ref).abs() / ref.abs().clamp_min(1e-9)).max().item()
    assert rel < 2e-3, rel
    # loss-only call (no margins written) gives the same sums
    sums2, P2 = mops.eval_tc(X, Bt, y, kind=kind, d=d)
    assert P2 is None and torch.allclose(sums, sums2, rtol=1e-12)


@pytest.mark.parametrize("rows,d,C,model", [(256, 128, 4, 2), (1000, 200, 7, 2), (4000, 1024, 16, 2),
                                            (777, 96, 20, 2), (1500, 300, 5, 3), (640, 4096, 16, 2)])
def test_grad_mc_tensor_core_softmax_and_batched_logistic(rows, d, C, model):
    """Both tcgen05 products (forward K-major, backward MN-major X^T) vs an fp64 torch reference."""
    from erasurehead_b200.ops import mc, synth
    torch.manual_seed(rows + d + C)
    ld = synth.pad_cols(d)
    Xs, ys = [], []
    for _ in range(2):
        X = torch.zeros(rows, ld, device=DEV)
        X[:, :d] = torch.randn(rows, d, device=DEV) / np.sqrt(d)
        Xs.append(X)
        ys.append(torch.randint(0, C, (rows,), device=DEV).float() if model == 2
                  else torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0))
    W = torch.randn(d, C, device=DEV)
    Wp = mc.pack_params(W)
    cf = [1.0, -0.6]
    G, loss = mc.grad_mc(Xs, ys, Wp, d, C, coeff=cf, model=model, want_loss=True)
    want = torch.zeros(d, C, dtype=torch.float64, device=DEV)
    want_loss = 0.0
    for X, y, c in zip(Xs, ys, cf):
        P = X[:, :d].double() @ W.double()
        if model == 2:
            S = torch.softmax(P, dim=1)
            want_loss += -torch.log(S[torch.arange(rows), y.long()]).sum().item()
            S[torch.arange(rows), y.long()] -= 1.0
            R = S
        else:
            yy = y.double()[:, None]
            want_loss += torch.nn.functional.softplus(-yy * P).sum().item()
            R = -yy / (torch.exp(P * yy) + 1)
        want += X[:, :d].double().t() @ (c * R)
    got = mc.unpack_params(G, d, C).double()
    scale = max(1.0, want.abs().max().item())
    assert (got - want).abs().max().item() <= 4e-3 * scale
    assert loss.item() == pytest.approx(want_loss, rel=2e-3)
    assert torch.all(G[C:] == 0)                       # padded classes carry no gradient


@pytest.mark.parametrize("rows,d,C,model", [(256, 128, 4, 2), (1000, 200, 7, 2), (777, 96, 20, 2), (1500, 54, 7, 2),
                                            (90000, 256, 16, 2), (70001, 120, 5, 3), (3000, 384, 3, 3),
                                            (9000, 640, 10, 2), (5000, 1024, 16, 2), (2500, 2048, 24, 2),
                                            (3333, 4096, 7, 2), (1200, 3000, 12, 3)])
@pytest.mark.parametrize("variant", ["auto", "cluster"])
def test_grad_mc_one_pass_kernels(rows, d, C, model, variant, monkeypatch):
    """One-pass kernels (mma.sync forward + tcgen05 backward from one shared-memory tile, one launch):
    the single-CTA kernel for narrow models and the thread-block-cluster kernel (column slices per
    CTA, DSMEM exchange of the partial margins) must agree with fp64 torch and the two-pass kernels."""
    from erasurehead_b200.ops import mc, synth
    if variant == "cluster":
        monkeypatch.setenv("EH_MC_ONEPASS", "cluster")
    torch.manual_seed(rows + d + C)
    ld = synth.pad_cols(d)
    Xs, ys = [], []
    for _ in range(3):
        X = torch.zeros(rows, ld, device=DEV)
        X[:, :d] = torch.randn(rows, d, device=DEV) / np.sqrt(d)
        Xs.append(X)
        ys.append(torch.randint(0, C, (rows,), device=DEV).float() if model == 2
                  else torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0))
    W = torch.randn(d, C, device=DEV)
    Wp = mc.pack_params(W)
    cf = [1.0, -0.6, 0.25]
    G, _ = mc.grad_mc(Xs, ys, Wp, d, C, coeff=cf, model=model)
    G2, _ = mc.grad_mc(Xs, ys, Wp, d, C, coeff=cf, model=model)
    if variant == "cluster" or d <= 1024:             # one-pass kernels: fixed-order reductions, bitwise repeatable
        assert torch.equal(G, G2)
    want = torch.zeros(d, C, dtype=torch.float64, device=DEV)
    for X, y, c in zip(Xs, ys, cf):
        P = X[:, :d].double() @ W.double()
        if model == 2:
            R = torch.softmax(P, dim=1)
            R[torch.arange(rows), y.long()] -= 1.0
        else:
            yy = y.double()[:, None]
            R = -yy / (torch.exp(P * yy) + 1)
        want += X[:, :d].double().t() @ (c * R)
    got = mc.unpack_params(G, d, C).double()
    scale = max(1.0, want.abs().max().item())
    assert (got - want).abs().max().item() <= 4e-3 * scale
    assert torch.all(G[C:] == 0) and torch.all(G[:, d:] == 0)
    monkeypatch.setenv("EH_MC_TWOPASS", "1")
    G3, _ = mc.grad_mc(Xs, ys, Wp, d, C, coeff=cf, model=model)
    assert (mc.unpack_params(G3, d, C).double() - got).abs().max().item() <= 4e-3 * scale


@pytest.mark.parametrize("rows,d,C,model", [(256, 128, 4, 2), (1000, 200, 7, 2), (777, 96, 20, 2), (1500, 54, 7, 2),
                                            (90000, 256, 16, 2), (70001, 120, 5, 3), (30000, 512, 10, 2),
                                            (5000, 640, 3, 3)])
def test_grad_mc_bf16_rows_all_tcgen05(rows, d, C, model):
    """bf16 feature rows: forward (K-major) and backward (MN-major) tcgen05 products from the same
    128B-swizzled TMA tile, one launch.  Reference: fp64 torch on the same bf16 values; W and the
    residual are rounded to bf16 as MMA operands, hence the looser tolerance."""
    from erasurehead_b200.ops import mc, synth
    torch.manual_seed(rows + d + C)
    ld = synth.pad_cols(d)
    Xs, ys = [], []
    for _ in range(2):
        X = torch.zeros(rows, ld, device=DEV, dtype=torch.bfloat16)
        X[:, :d] = (torch.randn(rows, d, device=DEV) / np.sqrt(d)).to(torch.bfloat16)
        Xs.append(X)
        ys.append(torch.randint(0, C, (rows,), device=DEV).float() if model == 2
                  else torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0))
    W = torch.randn(d, C, device=DEV)
    Wp = mc.pack_params(W)
    cf = [1.0, -0.6]
    G, _ = mc.grad_mc(Xs, ys, Wp, d, C, coeff=cf, model=model)
    G2, _ = mc.grad_mc(Xs, ys, Wp, d, C, coeff=cf, model=model)
    assert torch.equal(G, G2)
    want = torch.zeros(d, C, dtype=torch.float64, device=DEV)
    # up to 16 columns: parameters and residuals enter the MMAs as hi + lo bf16 pairs (16 mantissa bits)
    split = C <= 16
    Wb = W.double() if split else W.to(torch.bfloat16).double()
    for X, y, c in zip(Xs, ys, cf):
        P = X[:, :d].double() @ Wb
        if model == 2:
            R = torch.softmax(P, dim=1)
            R[torch.arange(rows), y.long()] -= 1.0
        else:
            yy = y.double()[:, None]
            R = -yy / (torch.exp(P * yy) + 1)
        want += X[:, :d].double().t() @ (c * R)
    got = mc.unpack_params(G, d, C).double()
    scale = max(1.0, want.abs().max().item())
    assert (got - want).abs().max().item() <= (5e-4 if split else 1.5e-2) * scale
    assert torch.all(G[C:] == 0) and torch.all(G[:, d:] == 0)


@pytest.mark.parametrize("d", [96, 1024, 3000])
def test_grad_dense_multi_batches_messages_in_one_launch(d):
    """Several messages (different partition counts / coefficients) share one launch."""
    from erasurehead_b200.ops import grad, synth
    torch.manual_seed(d)
    rows, ld = 333, synth.pad_cols(d)
    parts = []
    for _ in range(5):
        X = torch.zeros(rows, ld, device=DEV)
        X[:, :d] = torch.randn(rows, d, device=DEV) / np.sqrt(d)
        parts.append((X, torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0)))
    beta = torch.randn(d, device=DEV)
    msgs = [([0], [1.0]), ([1, 2, 3], [0.5, -2.0, 1.5]), ([4, 0], [1.0, 1.0]), ([2], [-1.0])]
    outs = [torch.full((d,), float("nan"), device=DEV) for _ in msgs]
    jobs = [grad.make_dense_job([parts[i][0] for i in idx], [parts[i][1] for i in idx], beta.data_ptr(),
                                o.data_ptr(), d, cf, 0) for (idx, cf), o in zip(msgs, outs)]
    sc = grad.dense_scratch(ld, DEV)
    grad.launch_dense_multi(jobs, sc)
    for (idx, cf), o in zip(msgs, outs):
        want = sum(ref_grad(parts[i][0][:, :d], parts[i][1], beta, c, 0) for i, c in zip(idx, cf))
        assert (o.double() - want).abs().max().item() <= 2e-4 * max(1.0, want.abs().max().item())
    # and the batched result equals the one-message-per-launch result bit for bit? (same tile -> CTA
    # mapping is not guaranteed, so compare within rounding only)
    single = grad.grad_dense([parts[1][0], parts[2][0], parts[3][0]], [parts[1][1], parts[2][1], parts[3][1]], beta,
                             coeff=[0.5, -2.0, 1.5], d=d, scratch=sc)
    assert torch.allclose(single, outs[1], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("rows,d", [(200, 64), (5000, 100), (3000, 250), (2000, 500), (1000, 1000), (1000, 1500),
                                    (900, 2048), (777, 4096), (300, 8000)])
@pytest.mark.parametrize("model", [0, 1])
def test_grad_dense_bf16_rows(rows, d, model):
    """bf16 feature storage (half the HBM traffic): exact against the reference on the same bf16 values."""
    from erasurehead_b200.ops import grad, synth
    torch.manual_seed(rows + d)
    ld = synth.pad_cols(d)
    Xb = torch.zeros(rows, ld, device=DEV, dtype=torch.bfloat16)
    Xb[:, :d] = (torch.randn(rows, d, device=DEV) / np.sqrt(d)).to(torch.bfloat16)
    y = torch.where(torch.rand(rows, device=DEV) > 0.5, 1.0, -1.0) if model == 0 else torch.randn(rows, device=DEV)
    beta = torch.randn(d, device=DEV)
    g = grad.grad_dense([Xb, Xb], [y, y], beta, coeff=(1.25, -0.5), model=model, d=d)
    want = ref_grad(Xb[:, :d].float(), y, beta, 0.75, model)
    assert (g.double() - want).abs().max().item() <= 2e-4 * max(1.0, want.abs().max().item())


def test_predict_margins_on_gpu_uses_tensor_core_path():
    from erasurehead_b200 import predict
    torch.manual_seed(0)
    X = torch.randn(500, 70, device=DEV)
    B = torch.randn(6, 70, device=DEV) * 0.1
    P = predict.margins(X, B)
    assert (P.double() - X.double() @ B.double().t()).abs().max().item() < 5e-3
    assert predict.predict_proba(X, B[0]).shape == (500,)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_kernels_on_a_second_device_in_one_process():
    """Per-device kernel attributes (dynamic shared memory) and the per-device scratch address of the
    no-alloc entry point: run the dense gradient on cuda:0, then on cuda:1, in the same process."""
    import ctypes as C
    from erasurehead_b200.ops import _lib, grad as gops
    L = _lib.lib()
    outs = []
    try:
        for dev in ("cuda:0", "cuda:1"):
            torch.cuda.set_device(dev)
            g = torch.Generator(device="cpu").manual_seed(3)
            X = (torch.randn(4096, 2048, generator=g) / 45.0).to(dev)
            y = torch.sign(torch.randn(4096, generator=g)).to(dev)
            beta = torch.randn(2048, generator=g).to(dev)
            a = gops.grad_dense(X, y, beta)
            b = torch.full((2048,), float("nan"), device=dev)
            job = gops.make_dense_job([X], [y], beta.data_ptr(), b.data_ptr(), 2048, [1.0], 0)
            _lib.check(L.eh_grad_dense_noalloc(C.byref(job), torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-5)
            outs.append(a.cpu())
        assert torch.allclose(outs[0], outs[1], rtol=1e-5, atol=1e-5)
    finally:
        torch.cuda.set_device(0)


@pytest.mark.parametrize("model,C_", [(2, 7), (2, 3), (3, 5)])
def test_grad_ell_mc_matches_reference(model, C_):
    """Multi-column gradient on one-hot features (column-major W[d][8]) against float64 torch."""
    from erasurehead_b200.data import generate_onehot
    from erasurehead_b200.ops import grad
    d, nnz, rpp = 6000, 23, 2500
    ds = generate_onehot(3 * rpp, d, nnz, 3, seed=6, task="softmax", classes=C_)
    idxs = [torch.from_numpy(ds.ell(p)).to(DEV) for p in range(3)]
    if model == 2:
        ys = [torch.tensor(ds.labels(p), dtype=torch.float32, device=DEV) for p in range(3)]
    else:
        ys = [torch.where(torch.rand(rpp, device=DEV) > 0.5, 1.0, -1.0) for _ in range(3)]
    cf = [1.0, -2.5, 0.25]
    W = torch.zeros(d, 8, device=DEV)
    W[:, :C_] = torch.randn(d, C_, device=DEV) * 0.3
    G = grad.grad_ell_mc(idxs, ys, W, C_, coeff=cf, model=model)
    want = torch.zeros(d, C_, dtype=torch.float64, device=DEV)
    for i, y, c in zip(idxs, ys, cf):
        X = torch.sparse_csr_tensor(torch.arange(0, i.numel() + 1, nnz, device=DEV), i.reshape(-1).long(),
                                    torch.ones(i.numel(), dtype=torch.float64, device=DEV), size=(i.shape[0], d))
        P = X @ W[:, :C_].double()
        if model == 2:
            R = torch.softmax(P, dim=1)
            R[torch.arange(len(y)), y.long()] -= 1.0
        else:
            R = -y.double()[:, None] / (torch.exp(P * y.double()[:, None]) + 1)
        want += c * (X.t().to_sparse_csr() @ R)
    assert (G[:, :C_].double() - want).abs().max().item() <= 5e-4 * max(1.0, want.abs().max().item())
    assert G[:, C_:].abs().max().item() == 0.0


def test_synth_onehot_and_multiclass_labels_on_device():
    """Device generators of the named dataset shapes (K14): fixed ones per row inside their blocks, bias last,
    reproducible per (seed, partition), label distributions sane; dense multi-class labels."""
    from erasurehead_b200.ops import synth
    d, nnz, rows = 15509, 55, 20000
    edges = synth.onehot_edges(d, nnz, 3)
    idx, y = synth.synth_onehot(rows, d, nnz, 3, 7, task=0)
    idx2, y2 = synth.synth_onehot(rows, d, nnz, 3, 7, task=0)
    assert torch.equal(idx, idx2) and torch.equal(y, y2)
    idx3, _ = synth.synth_onehot(rows, d, nnz, 3, 8, task=0)
    assert not torch.equal(idx, idx3)
    e = torch.from_numpy(edges).to(DEV)
    assert torch.all(idx[:, -1] == d - 1)
    assert torch.all(idx[:, :-1] >= e[:-1][None, :]) and torch.all(idx[:, :-1] < e[1:][None, :])
    assert set(y.unique().tolist()) == {-1.0, 1.0} and 0.1 < (y > 0).float().mean().item() < 0.9
    # heavy-tailed popularity: the most frequent level of a wide block takes a large share of the rows
    wide = int(np.argmax(np.diff(edges)))
    top = torch.bincount(idx[:, wide].long() - int(edges[wide])).max().item() / rows
    assert 0.1 < top < 0.6
    _, yl = synth.synth_onehot(rows, d, nnz, 3, 7, task=1)
    assert yl.std().item() > 0.05
    _, yc = synth.synth_onehot(rows, d, nnz, 3, 7, task=2, classes=7)
    cnt = torch.bincount(yc.long(), minlength=7)
    assert cnt.numel() == 7 and cnt.min().item() > rows // 70
    Xd, yd = synth.synth_gmm(4096, 512, seed=5, partition=1, task=2, classes=5)
    cd = torch.bincount(yd.long(), minlength=5)
    assert cd.numel() == 5 and cd.min().item() > 4096 // 50
    X0, _ = synth.synth_gmm(4096, 512, seed=5, partition=1, task=0)
    assert torch.equal(Xd, X0)              # the features do not depend on the label task


def test_auc_batch_equals_per_column_auc():
    """All iterates' ROC-AUC from one sort + two scans: identical to the per-column routine, ties included."""
    from erasurehead_b200.ops import metrics as mops
    torch.manual_seed(3)
    n, R = 20011, 13
    S = torch.randn(R, n, device=DEV)
    S[2] = torch.round(S[2] * 2) / 2          # heavy ties
    S[5] = 0.0                                # all tied -> 0.5
    S[7, ::3] = -0.0
    S[7, 1::3] = 0.0
    y = torch.where(torch.rand(n, device=DEV) > 0.3, 1.0, -1.0)
    got = mops.auc_batch(S, y)
    want = torch.stack([mops.auc(S[i], y) for i in range(R)])
    assert torch.allclose(got, want, rtol=0, atol=1e-12)
    assert abs(got[5].item() - 0.5) < 1e-12
    assert torch.allclose(mops.auc_batch(S, y, max_elems=3 * n), want, rtol=0, atol=1e-12)      # chunked path
