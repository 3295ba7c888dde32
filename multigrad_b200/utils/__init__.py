"""Host-side helpers: random keys, timing / NVTX, checkpoints."""
from . import rng  # noqa: F401
from . import timing  # noqa: F401
from . import checkpoint  # noqa: F401
