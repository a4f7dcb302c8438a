"""Model interface shared by the oracle (numpy, float64) and the engine (device ids)."""
from __future__ import annotations

import numpy as np

MODEL_LOGISTIC = 0
MODEL_LINEAR = 1
MODEL_SOFTMAX = 2

_REGISTRY = {}


class Model:
    """A generalised linear model trained by (coded) full-batch gradient descent.

    ``gradient(X, y, coeff, beta)`` returns the *encoded* partial gradient
    ``X^T diag(coeff) r(X beta, y)`` including the model's sign/scale, exactly what a worker
    ships.  ``coeff`` is the per-row encode weight (1 for uncoded partitions).
    """
    name = "?"
    device_id = -1
    metric = "loss"          # name of the headline evaluation metric
    has_auc = False
    n_out = 1                # parameter columns

    def margins(self, X, beta):
        return X.dot(beta)

    def residual(self, p, y):          # pragma: no cover - interface
        raise NotImplementedError

    def gradient(self, X, y, coeff, beta):
        p = self.margins(X, beta)
        r = self.residual(p, y)
        if r.ndim == 1:
            r = r * coeff
        else:
            r = r * np.asarray(coeff).reshape(-1, 1)
        return X.T.dot(r)

    def loss(self, y, p):              # pragma: no cover - interface
        raise NotImplementedError

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        if cls.name != "?":
            _REGISTRY[cls.name] = cls


def get_model(name: str, classes: int = 1) -> Model:
    try:
        cls = _REGISTRY[name]
    except KeyError:
        raise ValueError(f"unknown model {name!r}; have {sorted(_REGISTRY)}") from None
    return cls(classes) if name == "softmax" else cls()
