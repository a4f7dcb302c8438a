"""Device operators: thin torch-tensor wrappers over the C ABI of ``liberasurehead.so``."""
from ._lib import EhError, build, lib  # noqa: F401
