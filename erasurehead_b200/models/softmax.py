"""Multi-class softmax regression: the multi-column extension (BASELINE.json config 5).

Parameters are a ``d x C`` matrix; labels are class ids ``0..C-1``.  The worker gradient is
``X^T diag(coeff) (softmax(X W) - onehot(y))`` -- two GEMM-shaped products per row tile, which
is what the tcgen05/TMEM kernel path exists for.  No counterpart in the reference (Covertype
is binarised there, arrange_real_data.py:154-156).
"""
import numpy as np

from .base import Model, MODEL_SOFTMAX


class SoftmaxRegression(Model):
    name = "softmax"
    device_id = MODEL_SOFTMAX
    metric = "loss"

    def __init__(self, classes: int = 2):
        self.n_out = int(classes)

    def residual(self, p, y):
        z = p - p.max(axis=1, keepdims=True)
        e = np.exp(z)
        sm = e / e.sum(axis=1, keepdims=True)
        sm[np.arange(len(y)), np.asarray(y).astype(int)] -= 1.0
        return sm

    def loss(self, y, p):
        z = p - p.max(axis=1, keepdims=True)
        lse = np.log(np.exp(z).sum(axis=1))
        return float(np.mean(lse - z[np.arange(len(y)), np.asarray(y).astype(int)]))
