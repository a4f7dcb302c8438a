"""Model families: the per-row residual, the gradient scale and the evaluation metric.

The reference has two (binary logistic regression with labels +-1, and linear regression
selected by the ``kc_house_data`` dataset name); multi-class softmax regression is the
multi-column extension named by BASELINE.json (Covertype-shaped multi-class).
"""
from .base import Model, get_model, MODEL_LOGISTIC, MODEL_LINEAR, MODEL_SOFTMAX  # noqa: F401
from .logistic import LogisticRegression  # noqa: F401
from .linear import LinearRegression  # noqa: F401
from .softmax import SoftmaxRegression  # noqa: F401
