"""NVLink peer workspace of a communicator (component N1 of SURVEY.md).

Every rank ``cudaMalloc`` s one block, exports it with CUDA IPC, and maps the
blocks of all other ranks of the (sub)communicator, so that kernels can load
and store peer memory directly over NVLink/NVSwitch.  The handles travel once
over the cold-path object channel (gloo); after that no host-side message is
needed by any collective: synchronisation is by epoch-tagged flags inside the
workspace (``csrc/mgb_common.cuh``).

This replaces the communicator plumbing of the reference
(``multigrad/multigrad.py:15-27`` MPI bootstrap, ``:88-146`` ``Split``): a
sub-communicator made by ``split_subcomms`` simply builds its own workspace
over its own GPUs.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List

import torch

from ..ops import _lib

__all__ = ["PeerWorkspace"]

DEFAULT_LL_UNITS = 8192            # 8-byte slots per (parity, source)
DEFAULT_BULK_BYTES = 64 << 20      # each of the two bulk regions
DEFAULT_TIMEOUT_MS = 20000


class _Multicast:
    """Handles and the two views of a multicast-bound buffer.  The buffer is
    split in two: ``[0, alloc_bytes)`` holds the symmetric tensors handed out
    by ``PeerWorkspace.multicast_tensor``, ``[stage_off, bytes)`` is the
    staging window through which ordinary tensors are reduced -- the two never
    overlap, so a staged reduction cannot overwrite a resident tensor."""
    __slots__ = ("handle", "mem", "uc", "mc", "bytes", "alloc_bytes",
                 "stage_off", "stage_bytes")

    def __init__(self, handle, mem, uc, mc, nbytes):
        self.handle, self.mem = handle, mem
        self.uc, self.mc, self.bytes = uc, mc, nbytes
        self.alloc_bytes = (nbytes // 2) & ~4095
        self.stage_off = self.alloc_bytes
        self.stage_bytes = nbytes - self.stage_off


class PeerWorkspace:
    """Symmetric peer-mapped memory + the ``MgbPeerTable`` kernels consume."""

    def __init__(self, comm, ll_units: int = None, bulk_bytes: int = None,
                 spin_timeout_ms: int = None):
        if not torch.cuda.is_available():
            raise RuntimeError("PeerWorkspace needs CUDA")
        self.comm = comm
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device())
        ll_units = int(ll_units or os.environ.get(
            "MULTIGRAD_B200_LL_UNITS", DEFAULT_LL_UNITS))
        bulk_bytes = int(bulk_bytes if bulk_bytes is not None else
                         os.environ.get("MULTIGRAD_B200_BULK_BYTES",
                                        DEFAULT_BULK_BYTES))
        timeout = int(spin_timeout_ms or os.environ.get(
            "MULTIGRAD_B200_SPIN_TIMEOUT_MS", DEFAULT_TIMEOUT_MS))
        world = comm.size
        if world > _lib.MGB_MAX_WORLD:
            raise ValueError(f"peer workspace supports <= "
                             f"{_lib.MGB_MAX_WORLD} ranks, got {world}")
        self.bytes = int(self.lib.mgb_ws_bytes(world, ll_units, bulk_bytes))
        local = C.c_void_p()
        _lib.check(self.lib.mgb_ws_alloc(self.bytes, C.byref(local)),
                   "mgb_ws_alloc")
        self._local = local
        self._imported: List[C.c_void_p] = []
        table = _lib.MgbPeerTable()
        table.world, table.rank = world, comm.rank
        table.ll_units, table.spin_timeout_ms = ll_units, timeout
        table.bulk_bytes = bulk_bytes
        if world > 1:
            handle = C.create_string_buffer(64)
            _lib.check(self.lib.mgb_ws_export(local, handle), "mgb_ws_export")
            handles = comm.allgather(bytes(handle.raw))
            for r, h in enumerate(handles):
                if r == comm.rank:
                    table.base[r] = local.value
                    continue
                p = C.c_void_p()
                _lib.check(self.lib.mgb_ws_import(h, C.byref(p)),
                           f"mgb_ws_import(rank {r})")
                self._imported.append(p)
                table.base[r] = p.value
            comm.Barrier()      # everyone mapped everyone before first use
        else:
            table.base[0] = local.value
        self.table = table
        self._closed = False
        self._mc = False            # not set up yet (None = unavailable)
        self._mc_used = 0           # bump allocator of multicast_tensor()

    # -------------------------------------------------------------- multicast
    def multicast(self):
        """NVLink-switch multicast buffer of this communicator (built on first
        use; collective call).  ``None`` when the GPUs / driver / topology do
        not offer multicast objects, when the ranks are not on one host, or
        with ``MULTIGRAD_B200_NVLS=0``: callers then fall back to the
        peer-load kernels."""
        if self._mc is False:
            self._mc = self._setup_multicast()
        return self._mc

    def multicast_tensor(self, numel: int, dtype=torch.float32):
        """A zeroed CUDA tensor carved out of the multicast-bound buffer
        (symmetric memory: every rank must make the same sequence of calls so
        that the tensors sit at the same offsets -- checked here, this is a
        collective call).  ``reduce_sum`` /
        ``all_reduce_sum_`` on such a tensor runs entirely inside the NVLink
        switch, without staging copies.  ``None`` when multicast is
        unavailable or the buffer is exhausted."""
        mc = self.multicast()
        if mc is None:
            return None
        esz = torch.empty(0, dtype=dtype).element_size()
        nbytes = (numel * esz + 255) & ~255
        # the upper half of the bound buffer is the staging window of the
        # reductions of ordinary tensors (ops/allreduce.py): symmetric
        # tensors only ever live in the lower half
        if self._mc_used + nbytes > mc.alloc_bytes:
            return None
        offset = self._mc_used
        # every rank must be at the same offset, or the in-switch reduction
        # would add unrelated memory (and ranks would pick different kernels)
        offs = self.comm.allgather((offset, nbytes))
        if any(o != offs[0] for o in offs):
            raise RuntimeError(
                "multicast_tensor: ranks are at different offsets of the "
                f"symmetric buffer ({offs}); every rank has to make the same "
                "sequence of multicast_tensor() calls")
        ptr = mc.uc + offset
        self._mc_used += nbytes
        typestr = {torch.float32: "<f4", torch.float64: "<f8",
                   torch.int32: "<i4", torch.uint8: "|u1"}[dtype]

        class _Raw:
            __cuda_array_interface__ = {
                "shape": (numel,), "typestr": typestr, "data": (ptr, False),
                "version": 2}
        t = torch.as_tensor(_Raw(), device=self.device)
        t.zero_()
        return t

    def _setup_multicast(self):
        import socket
        lib, comm, world = self.lib, self.comm, self.comm.size
        if world < 2:
            return None
        sup = C.c_int(0)
        ok = (os.environ.get("MULTIGRAD_B200_NVLS", "1") != "0"
              and lib.mgb_mc_supported(C.byref(sup)) == 0 and sup.value == 1
              and hasattr(socket, "send_fds"))
        votes = comm.allgather((bool(ok), socket.gethostname()))
        if not all(v for v, _ in votes) or len({h for _, h in votes}) != 1:
            return None
        want = int(os.environ.get("MULTIGRAD_B200_MC_BYTES", 128 << 20))
        size = C.c_uint64(0)
        rc = lib.mgb_mc_round_size(world, want, C.byref(size))
        if not all(r == 0 for r in comm.allgather(rc)):
            return None
        nbytes = int(size.value)
        handle = C.c_uint64(0)
        # the root creates the object and hands its file descriptor to every
        # peer over an abstract unix socket (SCM_RIGHTS)
        rc = 0
        if comm.rank == 0:
            fd = C.c_int(-1)
            rc = lib.mgb_mc_create(world, nbytes, C.byref(handle),
                                   C.byref(fd))
            srv, name = None, None
            if rc == 0:
                name = "\0mgb-mc-%d-%x" % (os.getpid(), id(self))
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.bind(name)
                srv.listen(world)
            comm.bcast(name, root=0)
            if rc == 0:
                srv.settimeout(60)
                try:
                    for _ in range(world - 1):
                        conn, _a = srv.accept()
                        socket.send_fds(conn, [b"fd"], [fd.value])
                        conn.close()
                except OSError:
                    rc = -1
                srv.close()
                os.close(fd.value)
        else:
            name = comm.bcast(None, root=0)
            if name is None:
                rc = -1
            else:
                try:
                    c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                    c.settimeout(60)
                    c.connect(name)
                    _msg, fds, _f, _a = socket.recv_fds(c, 16, 1)
                    c.close()
                    rc = lib.mgb_mc_import(fds[0], C.byref(handle))
                    os.close(fds[0])
                except (OSError, IndexError):
                    rc = -1
        if not all(r == 0 for r in comm.allgather(rc)):
            return None
        rc = lib.mgb_mc_add_device(handle.value)
        if not all(r == 0 for r in comm.allgather(rc)):
            return None             # (allgather doubles as the barrier)
        uc, mc, mem = C.c_void_p(), C.c_void_p(), C.c_uint64(0)
        rc = lib.mgb_mc_bind(handle.value, nbytes, C.byref(uc), C.byref(mc),
                             C.byref(mem))
        if not all(r == 0 for r in comm.allgather(rc)):
            return None
        return _Multicast(handle.value, mem.value, uc.value, mc.value, nbytes)

    # ------------------------------------------------------------------ state
    def read_ctrl(self, idx: int) -> int:
        out = C.c_uint32()
        _lib.check(self.lib.mgb_ws_read_ctrl(
            self._local, idx, C.byref(out),
            C.c_void_p(torch.cuda.current_stream().cuda_stream)),
            "mgb_ws_read_ctrl")
        return int(out.value)

    @property
    def epoch(self) -> int:
        return self.read_ctrl(0)

    def check_error(self) -> None:
        """Raise if any in-kernel wait timed out (sticky device flag)."""
        if self.read_ctrl(1):
            raise RuntimeError(
                f"rank {self.comm.rank}: a device-side wait on a peer GPU "
                f"timed out after {self.table.spin_timeout_ms} ms "
                "(a rank died, diverged, or never launched the collective)")

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize()
            if self.comm.size > 1:
                self.comm.Barrier()
            if self._mc:
                m = self._mc
                self.lib.mgb_mc_release(m.handle, m.mem, m.uc, m.mc, m.bytes)
                self._mc = None
            for p in self._imported:
                self.lib.mgb_ws_unimport(p)
            self.lib.mgb_ws_free(self._local)
        except Exception:
            pass

    def __del__(self):
        # never run collectives from a finaliser
        pass
