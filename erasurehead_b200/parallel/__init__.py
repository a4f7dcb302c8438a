"""Process-group plumbing and the peer-mapped workspace (one process per GPU)."""
from .context import DistContext, init_distributed  # noqa: F401
from .workspace import Workspace  # noqa: F401
