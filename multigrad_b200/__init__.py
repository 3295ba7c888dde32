"""multigrad_b200 -- Blackwell-native data-parallel differentiable model
fitting with the capabilities and public API of AlanPearl/multigrad.

Exports mirror the reference (``multigrad/__init__.py:3-9``):
``OnePointModel, OnePointGroup, reduce_sum, split_subcomms,
split_subcomms_by_node, util`` plus the B200-native additions (``models``,
``ops``, ``parallel``).

Like the reference (which runs ``MPI_Init`` at import,
``multigrad/multigrad.py:15-21``) importing this package under ``torchrun``
(``WORLD_SIZE > 1``) initialises the process group; set
``MULTIGRAD_B200_NO_AUTO_INIT=1`` to do it yourself with
``multigrad_b200.parallel.init_distributed()``.
"""
import os as _os

from ._version import __version__  # noqa: F401
from . import util  # noqa: F401
from .parallel import comm as _comm
from .parallel.comm import Comm, init_distributed, get_world, shutdown  # noqa
from .parallel.collectives import (reduce_sum, split_subcomms,  # noqa: F401
                                   split_subcomms_by_node, scatter_nd)
from .core import OnePointModel, OnePointGroup  # noqa: F401

if (int(_os.environ.get("WORLD_SIZE", "1") or 1) > 1
        and not _os.environ.get("MULTIGRAD_B200_NO_AUTO_INIT")):
    init_distributed()


def __getattr__(name):
    # COMM / RANK / N_RANKS are resolved lazily so that a late
    # ``init_distributed()`` is honoured (reference: module globals,
    # ``multigrad.py:18-20``).
    if name == "COMM":
        return get_world()
    if name == "RANK":
        return get_world().rank
    if name == "N_RANKS":
        return get_world().size
    if name in ("models", "ops", "parallel", "utils", "experimental", "adam",
                "bfgs"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)


__all__ = [
    "OnePointModel", "OnePointGroup", "reduce_sum",
    "split_subcomms", "split_subcomms_by_node", "util",
    "scatter_nd", "Comm", "init_distributed", "get_world", "shutdown",
]
