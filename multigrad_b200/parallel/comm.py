"""Process / communicator bootstrap (layer L0' of SURVEY.md).

The reference boots MPI at import time and exposes module globals
``COMM, RANK, N_RANKS`` (reference ``multigrad/multigrad.py:15-27``).  Here the
process model is *one process per B200* launched by ``torchrun``; the
communicator is a thin MPI-flavoured facade over ``torch.distributed`` that is
used **only for initialisation and cold-path object traffic**.  Hot-path
reductions never go through it: they are issued from inside our CUDA kernels
over the NVLink peer workspace owned by :class:`Comm` (see
``parallel/workspace.py``).

A :class:`Comm` offers the subset of the mpi4py surface that the reference and
its users touch: ``rank``/``size``/``name``, ``Get_rank``/``Get_size``,
``bcast``, ``allgather``, ``Barrier``, ``Split``, ``send``/``recv``,
``Allreduce``/``Reduce``.
"""
from __future__ import annotations

import datetime
import os
import pickle
import socket
from typing import Any, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

__all__ = [
    "Comm", "init_distributed", "get_world", "is_distributed",
    "local_device", "shutdown",
]

_WORLD: Optional["Comm"] = None
_DEVICE: Optional[torch.device] = None


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def local_device() -> torch.device:
    """Device owned by this process (``cuda:LOCAL_RANK`` or ``cpu``)."""
    global _DEVICE
    if _DEVICE is None:
        if torch.cuda.is_available():
            idx = _env_int("LOCAL_RANK", 0) % max(torch.cuda.device_count(), 1)
            torch.cuda.set_device(idx)
            _DEVICE = torch.device("cuda", idx)
        else:
            _DEVICE = torch.device("cpu")
    return _DEVICE


class Comm:
    """MPI-like communicator over a ``torch.distributed`` process group.

    ``group is None`` denotes the single-process communicator (the analogue of
    the reference's ``COMM = None`` fallback, ``multigrad.py:23-27``, except
    that it is still a usable object with ``rank == 0`` and ``size == 1``).
    """

    def __init__(self, group=None, ranks: Optional[Sequence[int]] = None,
                 name: str = "WORLD", parent: Optional["Comm"] = None):
        self._group = group
        if group is None:
            self._ranks = [0] if ranks is None else list(ranks)
            self.rank = 0
            self.size = 1
        else:
            self._ranks = (list(range(dist.get_world_size()))
                           if ranks is None else list(ranks))
            self.size = len(self._ranks)
            self.rank = self._ranks.index(dist.get_rank())
        self.name = name
        self.parent = parent
        self._workspace = None          # lazily-built NVLink peer workspace
        self._nccl_ok = (group is not None and torch.cuda.is_available())

    # -- mpi4py style accessors -------------------------------------------
    def Get_rank(self) -> int:
        return self.rank

    def Get_size(self) -> int:
        return self.size

    def Get_name(self) -> str:
        return self.name

    def Set_name(self, name: str) -> None:
        self.name = name

    @property
    def ranks(self) -> List[int]:
        """Global (world) ranks of the members, in communicator order."""
        return list(self._ranks)

    @property
    def group(self):
        return self._group

    def global_rank(self, local: int) -> int:
        return self._ranks[local]

    def __repr__(self):
        return f"Comm(name={self.name!r}, rank={self.rank}, size={self.size})"

    def __hash__(self):
        return hash((self.name, tuple(self._ranks)))

    def __eq__(self, other):
        return self is other

    # -- object (cold path) collectives -------------------------------------
    def bcast(self, obj: Any = None, root: int = 0) -> Any:
        """Broadcast a picklable object from ``root`` (communicator rank)."""
        if self.size == 1:
            return obj
        box = [obj if self.rank == root else None]
        dist.broadcast_object_list(box, src=self._ranks[root],
                                   group=self._group,
                                   device=torch.device("cpu"))
        return box[0]

    def allgather(self, obj: Any) -> List[Any]:
        if self.size == 1:
            return [obj]
        out: List[Any] = [None] * self.size
        dist.all_gather_object(out, obj, group=self._group)
        return out

    def gather(self, obj: Any, root: int = 0) -> Optional[List[Any]]:
        out = self.allgather(obj)
        return out if self.rank == root else None

    def Barrier(self) -> None:
        if self.size > 1:
            # A gloo all-reduce of one int: works for every backend mix and
            # does not touch the GPU stream.
            t = torch.zeros(1, dtype=torch.int32)
            dist.all_reduce(t, group=self._group)

    barrier = Barrier

    def send(self, obj: Any, dest: int, tag: int = 0) -> None:
        payload = np.frombuffer(pickle.dumps(obj), dtype=np.uint8).copy()
        size = torch.tensor([payload.size], dtype=torch.int64)
        dist.send(size, dst=self._ranks[dest], group=self._group, tag=tag)
        dist.send(torch.from_numpy(payload), dst=self._ranks[dest],
                  group=self._group, tag=tag)

    def recv(self, source: int, tag: int = 0) -> Any:
        size = torch.zeros(1, dtype=torch.int64)
        dist.recv(size, src=self._ranks[source], group=self._group, tag=tag)
        buf = torch.empty(int(size.item()), dtype=torch.uint8)
        dist.recv(buf, src=self._ranks[source], group=self._group, tag=tag)
        return pickle.loads(buf.numpy().tobytes())

    # -- buffer collectives (host path; device tensors use the workspace) ---
    def Allreduce(self, sendbuf, recvbuf=None, op: str = "sum"):
        """Sum a host buffer over ranks (gloo).  ``recvbuf`` is filled in
        place when given (mpi4py upper-case semantics) and also returned."""
        arr = _c_array(sendbuf)
        if self.size > 1:
            t = torch.from_numpy(arr.copy())
            if t.dtype == torch.bool:
                t = t.to(torch.int64)
            dist.all_reduce(t, op=_op(op), group=self._group)
            arr = t.numpy().astype(arr.dtype, copy=False)
        if recvbuf is not None:
            np.copyto(np.asarray(recvbuf), arr)
        return arr

    def Reduce(self, sendbuf, recvbuf=None, op: str = "sum", root: int = 0):
        arr = _c_array(sendbuf)
        if self.size > 1:
            t = torch.from_numpy(arr.copy())
            dist.reduce(t, dst=self._ranks[root], op=_op(op),
                        group=self._group)
            arr = t.numpy() if self.rank == root else np.zeros_like(arr)
        if recvbuf is not None and self.rank == root:
            np.copyto(np.asarray(recvbuf), arr)
        return arr

    # -- communicator management -------------------------------------------
    def Split(self, color: int = 0, key: int = 0) -> "Comm":
        """Partition into disjoint sub-communicators (``MPI_Comm_split``).

        Ranks with equal ``color`` land in the same group, ordered by
        ``(key, rank)``.  Implemented with ``torch.distributed.new_group`` and
        group-local synchronisation so that it also works on a sub-comm.
        """
        if self.size == 1:
            return Comm(None, ranks=self._ranks, name=f"{self.name}.{color}",
                        parent=self)
        info = self.allgather((int(color), int(key), self.rank))
        colors = sorted({c for c, _, _ in info})
        mine = None
        for c in colors:
            members = sorted([(k, r) for cc, k, r in info if cc == c])
            granks = [self._ranks[r] for _, r in members]
            if int(color) == c:
                pg = dist.new_group(ranks=granks,
                                    use_local_synchronization=True)
                mine = Comm(pg, ranks=granks, name=f"{self.name}.{c}",
                            parent=self)
        assert mine is not None
        return mine

    def Free(self) -> None:
        if self._workspace is not None:
            self._workspace.close()
            self._workspace = None
        if self._group is not None and self.parent is not None:
            try:
                dist.destroy_process_group(self._group)
            except Exception:
                pass
            self._group = None

    # -- device workspace -----------------------------------------------------
    def workspace(self):
        """NVLink peer workspace (built on first use; collective call)."""
        if self._workspace is None:
            from .workspace import PeerWorkspace
            self._workspace = PeerWorkspace(self)
        return self._workspace


def _c_array(buf) -> np.ndarray:
    """C-contiguous view/copy that keeps 0-d shapes (``ascontiguousarray``
    would promote scalars to 1-d)."""
    arr = np.asarray(buf)
    return np.ascontiguousarray(arr).reshape(arr.shape)


def _op(op):
    if isinstance(op, str):
        return {"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX,
                "min": dist.ReduceOp.MIN, "prod": dist.ReduceOp.PRODUCT}[
                    op.lower()]
    return op


def is_distributed() -> bool:
    return dist.is_available() and dist.is_initialized()


def init_distributed(backend: Optional[str] = None,
                     timeout_s: float = 600.0) -> Comm:
    """Initialise ``torch.distributed`` from the ``torchrun`` environment and
    return the world communicator.  Safe to call repeatedly.

    Backend default: ``cpu:gloo,cuda:nccl`` on a GPU box, ``gloo`` otherwise.
    The NCCL half is used by the *baseline proxies* and for bulk cold-path
    copies only; the product's hot path uses the peer workspace.
    """
    global _WORLD
    if _WORLD is not None:
        return _WORLD
    world = _env_int("WORLD_SIZE", 1)
    if world <= 1 and not is_distributed():
        _WORLD = Comm(None, name="WORLD")
        return _WORLD
    if not is_distributed():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dev = local_device()
        if backend is None:
            backend = "cpu:gloo,cuda:nccl" if dev.type == "cuda" else "gloo"
        kw = {}
        if dev.type == "cuda":
            kw["device_id"] = dev
        dist.init_process_group(
            backend=backend, rank=_env_int("RANK", 0), world_size=world,
            timeout=datetime.timedelta(seconds=timeout_s), **kw)
    _WORLD = Comm(dist.group.WORLD, name="WORLD")
    return _WORLD


def get_world() -> Comm:
    return init_distributed()


def shutdown() -> None:
    """Tear down workspaces and the process group (idempotent)."""
    global _WORLD
    if _WORLD is not None and _WORLD._workspace is not None:
        _WORLD._workspace.close()
        _WORLD._workspace = None
    if is_distributed():
        try:
            dist.destroy_process_group()
        except Exception:
            pass
    _WORLD = None


def processor_name() -> str:
    """Analogue of ``MPI.Get_processor_name`` (reference
    ``multigrad.py:65``)."""
    return os.environ.get("MULTIGRAD_B200_NODE_NAME", socket.gethostname())
