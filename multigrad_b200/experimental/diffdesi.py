"""Catalogue index utilities (component C15).

Parity target: ``multigrad/diffdesi_experimental/util.py`` --
``find_ultimate_top_indices`` (``:18-28``), ``sort_and_reindex`` (``:31-35``),
``sort_all_by_ultimate_top_dump`` (``:4-15``).  numpy inputs run on the host;
CUDA tensors run on the device (pointer jumping is our K8 kernel, sorting is
``torch.sort`` -- off the hot path).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

__all__ = ["find_ultimate_top_indices", "sort_and_reindex",
           "sort_all_by_ultimate_top_dump"]

MAX_RECURSION = 50


def find_ultimate_top_indices(indices):
    """Follow ``indices = indices[indices]`` to the fixed point (every object
    ends up pointing at its ultimate host).  Raises ``RecursionError`` after
    50 rounds, like the reference."""
    if isinstance(indices, torch.Tensor) and indices.is_cuda:
        return _find_top_cuda(indices)
    was_tensor = isinstance(indices, torch.Tensor)
    idx = indices.cpu().numpy().copy() if was_tensor else np.array(indices)
    count = 0
    while np.any(idx != idx[idx]):
        count += 1
        if count > MAX_RECURSION:
            raise RecursionError(
                f"Host search hasn't finished after {MAX_RECURSION} steps")
        idx = idx[idx]
    return torch.as_tensor(idx) if was_tensor else idx


def _find_top_cuda(indices: torch.Tensor) -> torch.Tensor:
    from ..ops import _lib
    lib = _lib.load()
    idx = indices.detach().to(torch.int64).contiguous().clone()
    tmp = torch.empty_like(idx)
    flag = torch.zeros(1, dtype=torch.int32, device=idx.device)
    rounds = C.c_int32(0)
    with torch.cuda.device(idx.device):
        rc = lib.mgb_pointer_jump(
            idx.data_ptr(), tmp.data_ptr(), idx.numel(), MAX_RECURSION,
            C.byref(rounds), flag.data_ptr(),
            C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc == -4:
        raise RecursionError(
            f"Host search hasn't finished after {MAX_RECURSION} steps")
    _lib.check(rc, "mgb_pointer_jump")
    return idx


def _argsort(a):
    if isinstance(a, torch.Tensor):
        return torch.argsort(a, stable=True)
    return np.argsort(a, kind="stable")


def sort_and_reindex(indices, argsort=None, argsort2=None):
    """``argsort2[indices][argsort]``: re-express pointers after the catalogue
    has been permuted by ``argsort``."""
    if not isinstance(indices, torch.Tensor):
        indices = np.asarray(indices)
    argsort = _argsort(indices) if argsort is None else argsort
    argsort2 = _argsort(argsort) if argsort2 is None else argsort2
    return argsort2[indices][argsort]


def sort_all_by_ultimate_top_dump(ultimate_dump, arrays_to_sort=(),
                                  arrays_to_sort_and_reindex=()):
    top = find_ultimate_top_indices(ultimate_dump)
    argsort = _argsort(top)
    argsort2 = _argsort(argsort)
    conv = (lambda x: x) if isinstance(top, torch.Tensor) else np.asarray
    sorted_arrays = [conv(x)[argsort] for x in arrays_to_sort]
    reindexed = [sort_and_reindex(x, argsort, argsort2)
                 for x in arrays_to_sort_and_reindex]
    return sorted_arrays, reindexed
