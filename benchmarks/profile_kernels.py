#!/usr/bin/env python
"""Launch each fused kernel a few times at a representative shape, for an ncu capture.

    python benchmarks/profile_kernels.py --which rowwarp,cluster,bf16,ell_smem,ell_global,round

Run it once plainly (must exit 0), then under
``ncu --set full --clock-control none --import-source on -k regex:<kernel> -c 1 -o gpurun_out/<name>``.
Nothing here is a timing: ncu replays kernels, numbers come from kernel_bench.py / bench.py.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch  # noqa: E402


def dense(rows, d, bf16=False, reps=2):
    from erasurehead_b200.ops import grad, synth
    X, y = synth.synth_gmm(rows, d, 1, 0)
    if bf16:
        X = X.to(torch.bfloat16)
    beta = torch.randn(d, device="cuda") * 0.01
    sc = grad.dense_scratch(X.stride(0), "cuda")
    o = torch.empty(d, device="cuda")
    for _ in range(reps):
        grad.grad_dense(X, y, beta, d=d, scratch=sc, out=o)
    torch.cuda.synchronize()
    assert torch.isfinite(o).all()


def ell(rows, d, nnz, reps=2):
    from erasurehead_b200.data import generate_onehot
    from erasurehead_b200.ops import grad
    ds = generate_onehot(rows, d, nnz, 1, seed=5, name="profile")
    idx = torch.from_numpy(ds.ell(0)).cuda()
    y = torch.tensor(ds.labels(0), dtype=torch.float32, device="cuda")
    beta = torch.randn(d, device="cuda") * 0.01
    sc = grad.ell_scratch(d, "cuda")
    o = torch.empty(d, device="cuda")
    for _ in range(reps):
        grad.grad_ell(idx, y, beta, bias_col=d - 1, scratch=sc, out=o)
    torch.cuda.synchronize()
    assert torch.isfinite(o).all()


def mc_fused(rows, d, C_, reps=2, bf16=False):
    from erasurehead_b200.ops import mc, synth
    X, _ = synth.synth_gmm(rows, d, 1, 0)
    if bf16:
        X = X.to(torch.bfloat16)
    y = torch.randint(0, C_, (rows,), device="cuda").float()
    Wp = mc.pack_params(torch.randn(d, C_, device="cuda") * 0.01)
    G = torch.empty_like(Wp)
    scr = mc.scratch(rows, d, C_, "cuda")
    job = mc.make_job([X], [y], Wp.data_ptr(), G.data_ptr(), d, C_, [1.0], mc.MODEL_SOFTMAX)
    for _ in range(reps):
        mc.launch(job, scr)
    torch.cuda.synchronize()
    assert torch.isfinite(G).all()


def ell_mc(rows, d, nnz, C_, reps=2):
    from erasurehead_b200.ops import grad, synth
    idx, y = synth.synth_onehot(rows, d, nnz, 1, 0, task=2, classes=C_)
    W = torch.zeros(d, 8, device="cuda")
    W[:, :C_] = torch.randn(d, C_, device="cuda") * 0.1
    for _ in range(reps):
        G = grad.grad_ell_mc(idx, y, W, C_)
    torch.cuda.synchronize()
    assert torch.isfinite(G).all()


def rounds():
    from erasurehead_b200 import Engine, RunConfig, triple_from_scheme
    ic, p, cv = triple_from_scheme("coded")
    cfg = RunConfig(9, 8 * 4096, 4096, "", 0, "x", ic, 2, p, cv, 0, 1, "AGD", rounds=8, synthetic=True, seed=1,
                    delay_scale=2e-4, eval="none", quiet=True)
    eng = Engine(cfg, None)
    eng.run()
    eng.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", default="rowwarp,cluster,bf16,ell_smem,ell_global,round")
    a = ap.parse_args()
    for w in a.which.split(","):
        if w == "rowwarp":
            dense(262144, 1024)
        elif w == "dense4096":
            dense(524288, 4096)             # one partition of the headline configuration
        elif w == "cluster":
            dense(16384, 16384)
        elif w == "bf16":
            dense(131072, 4096, bf16=True)
        elif w == "ell_smem":
            ell(396112, 15509, 55)
        elif w == "ell_global":
            ell(26215, 241915, 44)
        elif w == "mc_fused":
            mc_fused(2097152, 128, 10)
        elif w == "mc_fused256":
            mc_fused(1048576, 256, 16)
        elif w == "mc_bf16":
            mc_fused(4194304, 256, 16, bf16=True)
        elif w == "mc_cluster":
            mc_fused(524288, 1024, 16)
        elif w == "ell_mc":
            ell_mc(396112, 15509, 55, 7)
        elif w == "round":
            rounds()
        print("ok", w, flush=True)
