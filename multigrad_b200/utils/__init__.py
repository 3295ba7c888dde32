"""Host-side helpers: random keys, timing, clock sampling."""
from . import rng  # noqa: F401
