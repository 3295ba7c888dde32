"""Failure detection: an in-kernel wait on a peer that never
shows up must terminate within the watchdog time and raise the sticky error
flag instead of hanging the GPU.  Injected on ONE GPU with a two-rank peer
table whose "rank 1" block is never written by anybody."""
import ctypes as C
import time

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _fake_table(lib, _lib, timeout_ms=300):
    nbytes = int(lib.mgb_ws_bytes(2, 256, 1 << 16))
    a, b = C.c_void_p(), C.c_void_p()
    _lib.check(lib.mgb_ws_alloc(nbytes, C.byref(a)))
    _lib.check(lib.mgb_ws_alloc(nbytes, C.byref(b)))
    t = _lib.MgbPeerTable()
    t.world, t.rank, t.ll_units = 2, 0, 256
    t.spin_timeout_ms, t.bulk_bytes = timeout_ms, 1 << 16
    t.base[0], t.base[1] = a.value, b.value
    return t, a, b


def _ctrl(lib, base, idx):
    out = C.c_uint32()
    assert lib.mgb_ws_read_ctrl(base, idx, C.byref(out), None) == 0
    return out.value


def test_allreduce_times_out_and_flags_error():
    from multigrad_b200.ops import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    t, a, b = _fake_table(lib, _lib)
    try:
        x = torch.ones(8, device="cuda:0")
        assert _ctrl(lib, a, 1) == 0
        t0 = time.perf_counter()
        _lib.check(lib.mgb_allreduce_sum(C.byref(t), x.data_ptr(), 8, 0, 1,
                                         None))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        assert 0.2 < dt < 5.0, dt                 # bounded by the watchdog
        assert _ctrl(lib, a, 1) == 1              # sticky error flag
        assert _ctrl(lib, a, 0) == 1              # the epoch still advanced
    finally:
        lib.mgb_ws_free(a)
        lib.mgb_ws_free(b)


def test_fused_step_times_out_and_python_raises():
    import multigrad_b200 as mg
    from multigrad_b200.models import SMFModel, smf
    from multigrad_b200.ops import _lib
    from multigrad_b200.ops.erfbin import ErfBinPlan, MODE_LOSS_GRAD
    lib = _lib.load()
    torch.cuda.set_device(0)
    t, a, b = _fake_table(lib, _lib)
    try:
        x = torch.rand(4096, device="cuda:0") + 10.0
        plan = ErfBinPlan(x, np.linspace(9, 10, 11), np.full(10, 1e-4),
                          target=np.full(10, 1e-2), peer_table=t)
        t0 = time.perf_counter()
        out = plan.evaluate(torch.tensor([-1.0, 0.3]), MODE_LOSS_GRAD)
        torch.cuda.synchronize()
        assert time.perf_counter() - t0 < 5.0
        assert _ctrl(lib, a, 1) == 1

        class FakeComm:            # the path PeerWorkspace.check_error takes
            rank, size = 0, 2
        from multigrad_b200.parallel.workspace import PeerWorkspace
        ws = PeerWorkspace.__new__(PeerWorkspace)
        ws.comm, ws.lib, ws._local, ws.table = FakeComm(), lib, a, t
        with pytest.raises(RuntimeError, match="timed out"):
            ws.check_error()
    finally:
        lib.mgb_ws_free(a)
        lib.mgb_ws_free(b)
