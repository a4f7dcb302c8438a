"""Time the one-hot multi-column gradient kernel at the Covertype x 7 shape (device events, median of 10)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from erasurehead_b200.ops import synth, grad
d, nnz, rows = 15509, 55, 396112
idx, y = synth.synth_onehot(rows, d, nnz, 1, 0, task=2, classes=7)
W = torch.zeros(d, 8, device="cuda"); W[:, :7] = torch.randn(d, 7, device="cuda") * 0.1
for _ in range(3):
    grad.grad_ell_mc(idx, y, W, 7)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); grad.grad_ell_mc(idx, y, W, 7); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
print("grad_ell_mc covtype x7: median %.4f ms, min %.4f ms" % (sorted(ts)[len(ts) // 2], min(ts)))
