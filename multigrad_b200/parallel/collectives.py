"""Communication primitives (layer L1 of SURVEY.md).

Public surface mirrors the reference: ``reduce_sum`` (reference
``multigrad/multigrad.py:149-183``), ``split_subcomms`` (``:88-146``),
``split_subcomms_by_node`` (``:48-85``) and ``scatter_nd``
(``multigrad/util.py:65-77``).

Behavioural differences, all deliberate:

* CUDA tensors are reduced **on the device** by our own NVLink kernel
  (``ops/allreduce.py``); the reference forces a device->host copy and a host
  MPI call (Q8).
* ``root=r``: non-root ranks get zeros instead of uninitialised memory (Q7).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from . import comm as _comm
from .comm import Comm

__all__ = ["reduce_sum", "split_subcomms", "split_subcomms_by_node",
           "scatter_nd", "group_assignment"]

_DEFAULT = object()


def _resolve(comm):
    if comm is _DEFAULT:
        return _comm.get_world()
    return comm


def reduce_sum(value, root: Optional[int] = None, comm=_DEFAULT):
    """Sum ``value`` over every rank of ``comm``.

    Parameters
    ----------
    value : scalar | numpy array | torch tensor (CPU or CUDA)
    root : int, optional
        If given only this rank receives the sum (others receive zeros).
    comm : Comm, optional
        Defaults to the world communicator; ``None`` means "no communicator"
        and returns ``value`` unchanged, as in the reference.

    Returns
    -------
    Same kind as the input: Python scalar for scalars, ndarray for ndarrays,
    tensor (same device) for tensors.
    """
    comm = _resolve(comm)
    if comm is None:
        return value
    return_to_scalar = not hasattr(value, "__len__")
    if isinstance(value, torch.Tensor):
        return _reduce_tensor(value, root, comm)
    arr = np.asarray(value)
    if comm.size == 1:
        total = arr.copy()
    elif root is None:
        total = comm.Allreduce(arr)
    else:
        total = comm.Reduce(arr, root=root)
    if return_to_scalar:
        total = total.tolist()
    return total


def _reduce_tensor(value: torch.Tensor, root, comm: Comm) -> torch.Tensor:
    if comm.size == 1:
        return value.clone()
    if value.is_cuda:
        from ..ops import allreduce as _ar
        total = _ar.all_reduce_sum(value, comm)
        if root is not None and comm.rank != root:
            total = torch.zeros_like(total)
        return total
    total = value.detach().clone().contiguous()
    if root is None:
        dist.all_reduce(total, group=comm.group)
    else:
        dist.reduce(total, dst=comm.global_rank(root), group=comm.group)
        if comm.rank != root:
            total = torch.zeros_like(total)
    return total


def group_assignment(size: int, num_groups: Optional[int] = None,
                     ranks_per_group: Optional[Sequence[int]] = None):
    """Rank -> group index table used by :func:`split_subcomms`.

    Reproduces the reference mapping (``multigrad.py:112-128``; the tables in
    SURVEY.md C4): with ``num_groups=G`` rank ``r`` takes the first label of
    the ``r``-th near-even chunk of ``repeat(arange(G), ceil(size/G))``.
    """
    main_msg = "Specify either num_groups OR ranks_per_group"
    if num_groups is not None:
        assert ranks_per_group is None, main_msg
        assert size >= num_groups, \
            "Cannot create more subcomms than there are ranks"
        num_groups = int(num_groups)
        # NB: *not* truncated to ``size`` labels -- the reference slices the
        # row axis, which is a no-op, and the surplus labels shape the chunks.
        labels = np.repeat(np.arange(num_groups),
                           math.ceil(size / num_groups))
    else:
        assert ranks_per_group is not None, main_msg
        assert sum(ranks_per_group) == size, \
            "The sum of ranks_per_group must equal comm.size"
        num_groups = len(ranks_per_group)
        labels = np.repeat(np.arange(num_groups), ranks_per_group)
    table = [int(chunk[0]) for chunk in np.array_split(labels, size)]
    return table, num_groups


def split_subcomms(num_groups: Optional[int] = None,
                   ranks_per_group: Optional[Sequence[int]] = None,
                   comm=_DEFAULT):
    """Split ``comm`` into sub-communicators (not grouped by node).

    Returns ``(sub_comm, num_groups, group_index)`` like the reference
    (``multigrad.py:88-146``).  Each sub-communicator later grows its own
    NVLink peer workspace, so a group's reductions only involve its GPUs.
    """
    comm = _resolve(comm)
    assert comm is not None, "Cannot split subcomms without a communicator"
    table, num_groups = group_assignment(comm.size, num_groups,
                                         ranks_per_group)
    mine = table[comm.rank]
    sub = comm.Split(color=mine, key=comm.rank)
    sub.Set_name(f"{comm.name}.{mine}".replace("WORLD.", ""))
    return sub, num_groups, mine


def split_subcomms_by_node(comm=_DEFAULT):
    """Split ``comm`` by host name (reference ``multigrad.py:48-85``).

    On the single 8xB200 NVSwitch box this returns one group containing every
    rank; the multi-node case works over gloo/NCCL but has no peer workspace.
    """
    comm = _resolve(comm)
    assert comm is not None, "Cannot split subcomms without a communicator"
    node = _comm.processor_name()
    nodes = sorted(set(comm.allgather(node)))
    idx = nodes.index(node)
    sub = comm.Split(color=idx, key=comm.rank)
    sub.Set_name(f"{comm.name}.{idx}".replace("WORLD.", ""))
    return sub, len(nodes), idx


def scatter_nd(array, axis: int = 0, comm=_DEFAULT, root: int = 0):
    """Scatter an n-d array from ``root``: rank ``i`` gets the ``i``-th of
    ``size`` near-even chunks along ``axis`` (reference ``util.py:65-77``).

    Accepts numpy arrays or torch tensors on the root.  Host arrays travel
    over the cold-path (gloo) point-to-point channel.  A CUDA tensor on the
    root is scattered GPU-to-GPU over NVLink through the peer workspace (K7)
    and arrives as a CUDA tensor on every rank.
    """
    comm = _resolve(comm)
    if comm is None or comm.size == 1:
        return array
    on_device = comm.bcast(
        isinstance(array, torch.Tensor) and array.is_cuda
        if comm.rank == root else None, root=root)
    if on_device:
        return _scatter_nd_device(array, axis, comm, root)
    if comm.rank == root:
        if isinstance(array, torch.Tensor):
            splits = list(torch.tensor_split(array.cpu(), comm.size, dim=axis))
        else:
            splits = np.array_split(np.asarray(array), comm.size, axis=axis)
        ans = splits[root]
        for i in range(comm.size):
            if i != root:
                comm.send(splits[i], dest=i)
        return ans
    return comm.recv(source=root)


def _scatter_nd_device(array, axis, comm: Comm, root: int) -> torch.Tensor:
    """K7: scatter a CUDA tensor living on ``root`` over NVLink.  The root
    stages windows of the (axis-major, contiguous) array in the bulk region of
    its peer workspace; every rank copies its part of the window straight from
    the root's memory (``mgb_peer_copy``, a peer-to-peer cudaMemcpy)."""
    import ctypes as C
    ws = comm.workspace()
    lib = ws.lib
    if comm.rank == root:
        arr = array.detach().movedim(axis, 0).contiguous()
        meta = (tuple(arr.shape), arr.dtype)
    else:
        arr, meta = None, None
    shape, dtype = comm.bcast(meta, root=root)
    rows = shape[0]
    row_bytes = int(np.prod(shape[1:], dtype=np.int64)) * \
        torch.empty(0, dtype=dtype).element_size()
    base, extra = divmod(rows, comm.size)       # np.array_split boundaries
    lo = comm.rank * base + min(comm.rank, extra)
    hi = lo + base + (1 if comm.rank < extra else 0)
    dev = torch.device("cuda", torch.cuda.current_device())
    out = torch.empty((hi - lo,) + tuple(shape[1:]), dtype=dtype, device=dev)
    my_lo, my_hi = lo * row_bytes, hi * row_bytes
    window = (int(ws.table.bulk_bytes) // 2) & ~1023   # lower half of A
    off = int(lib.mgb_ws_bulk_offset(comm.size, ws.table.ll_units))
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    total = rows * row_bytes
    for w0 in range(0, max(total, 1), max(window, 1)):
        w1 = min(w0 + window, total)
        if comm.rank == root and w1 > w0:
            _check(lib.mgb_peer_copy(ws.table.base[root] + off,
                                     arr.data_ptr() + w0, w1 - w0, stream))
        torch.cuda.synchronize()
        comm.Barrier()                      # window staged
        a, b = max(w0, my_lo), min(w1, my_hi)
        if b > a:
            _check(lib.mgb_peer_copy(out.data_ptr() + (a - my_lo),
                                     ws.table.base[root] + off + (a - w0),
                                     b - a, stream))
        torch.cuda.synchronize()
        comm.Barrier()                      # window consumed
    return out.movedim(0, axis) if axis else out


def _check(rc):
    if rc:
        from ..ops import _lib
        _lib.check(rc, "mgb_peer_copy")
