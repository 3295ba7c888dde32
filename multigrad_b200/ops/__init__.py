"""Kernel wrappers (CUDA, sm_100a) and their plain-PyTorch references."""
from . import reference  # noqa: F401
