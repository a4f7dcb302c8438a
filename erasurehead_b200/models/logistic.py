"""Binary logistic regression, labels in {-1,+1} (reference naive.py:137-139, util.py:136-137)."""
import numpy as np

from .base import Model, MODEL_LOGISTIC


class LogisticRegression(Model):
    name = "logistic"
    device_id = MODEL_LOGISTIC
    has_auc = True

    def residual(self, p, y):
        # g = -X^T ( y / (exp(p*y) + 1) )      (naive.py:138)
        return -y / (np.exp(p * y) + 1.0)

    def loss(self, y, p):
        # mean log(1 + exp(-y p))              (util.py:136-137); logaddexp avoids overflow
        return float(np.mean(np.logaddexp(0.0, -y * p)))
