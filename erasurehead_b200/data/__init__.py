"""Datasets: containers, synthetic generators and the real-data preprocessor."""
from .dataset import Dataset, load_dataset  # noqa: F401
from .generate import generate_gmm, generate_gmm_to_disk  # noqa: F401
from .shapes import DATASET_SHAPES, generate_onehot  # noqa: F401
