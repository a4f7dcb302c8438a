"""Least-squares linear regression (reference naive.py:345-346, util.py:139-141)."""
import numpy as np

from .base import Model, MODEL_LINEAR


class LinearRegression(Model):
    name = "linear"
    device_id = MODEL_LINEAR
    metric = "mse"

    def residual(self, p, y):
        # g = -2 X^T (y - p)                   (naive.py:345-346)
        return -2.0 * (y - p)

    def loss(self, y, p):
        return float(np.mean((y - p) ** 2))
