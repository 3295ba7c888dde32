"""`mgb_erfbin_step_static`: the fused step for callers that cannot allocate.

Some users of the C library have no means of getting scratch memory: code
embedded in another runtime that only hands it input and output buffers, a
process whose device memory already belongs to another pool, or a phase in
which the CUDA allocator must not be called.  For them the library carries a
1 MiB scratch array inside its own image.  This test drives that entry point
straight through the C ABI (ctypes, raw device pointers) and checks, on two
calls in a row, that it returns what the ordinary entry point
(`mgb_erfbin_step`, scratch supplied by the caller) returns.

The file name makes pytest collect it last, so a run that stops at the first
failure has already exercised everything else.  It must also pass on its own
in a fresh process: the first call is what resolves the array's address.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

K, P = 10, 2


def _setup():
    from multigrad_b200.ops import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    g = torch.Generator().manual_seed(11)
    n = 200_003
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float().to(dev)
    md = _lib.MgbErfBinModel()
    md.x, md.w, md.n = x.data_ptr(), None, n
    md.nbins, md.nparams, md.mu_idx, md.sig_idx = K, P, 0, 1
    md.sigma_link, md.loss_kind, md.loss_eps = 0, 0, 0.0
    md.x0 = 10.65625
    edges = np.linspace(9.0, 10.0, K + 1)
    for k in range(K + 1):
        md.edges[k] = float(edges[k])
    for k in range(K):
        md.norm[k] = 1.0 / (10.0 * n * 0.1)
        md.target[k] = 10.0 ** (-2.0 - 0.3 * k)
        md.inv_err[k] = 1.0
    return _lib, lib, dev, x, md


def _run(lib, _lib, md, params, dev, static, batch=1):
    """One MODE_LOSS_GRAD launch through either entry point."""
    p = torch.tensor(params, dtype=torch.float32, device=dev).reshape(batch, P)
    out = {"s": torch.zeros(batch, K, device=dev),
           "j": torch.zeros(batch, K, P, device=dev),
           "l": torch.zeros(batch, device=dev),
           "g": torch.zeros(batch, P, device=dev)}
    a = _lib.MgbStepArgs()
    a.mode, a.batch, a.reduce = _lib.MODE_LOSS_GRAD, batch, 0
    a.params = p.data_ptr()
    a.out_sumstats, a.out_jac = out["s"].data_ptr(), out["j"].data_ptr()
    a.out_loss, a.out_grad = out["l"].data_ptr(), out["g"].data_ptr()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if static:
        rc = lib.mgb_erfbin_step_static(C.byref(md), C.byref(a), None, stream)
    else:
        nbytes = int(lib.mgb_erfbin_scratch_bytes(K, batch))
        scratch = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        rc = lib.mgb_erfbin_step(C.byref(md), C.byref(a), None,
                                 scratch.data_ptr(), nbytes, stream)
    torch.cuda.synchronize()
    return rc, {k: v.cpu() for k, v in out.items()}


def test_static_scratch_entry_matches_caller_scratch_entry():
    _lib, lib, dev, x, md = _setup()
    # two calls in a row: the second one finds the array as the first left it
    for params in ([-1.9, 0.3], [-2.05, 0.22]):
        rc_s, got = _run(lib, _lib, md, params, dev, static=True)
        assert rc_s == 0, lib.mgb_last_error()
        rc_r, want = _run(lib, _lib, md, params, dev, static=False)
        assert rc_r == 0, lib.mgb_last_error()
        assert torch.isfinite(got["l"]).all() and float(got["s"].sum()) > 0
        for key in ("s", "j", "l", "g"):
            assert torch.allclose(got[key], want[key], rtol=3e-6, atol=1e-12), \
                (key, got[key], want[key])
    # a small batch still fits the built-in array
    rc_s, got = _run(lib, _lib, md, [[-1.9, 0.3], [-2.0, 0.2], [-2.1, 0.25]],
                     dev, static=True, batch=3)
    rc_r, want = _run(lib, _lib, md, [[-1.9, 0.3], [-2.0, 0.2], [-2.1, 0.25]],
                      dev, static=False, batch=3)
    assert rc_s == 0 and rc_r == 0
    assert torch.allclose(got["s"], want["s"], rtol=3e-6, atol=1e-12)
    assert torch.allclose(got["g"], want["g"], rtol=3e-6, atol=1e-12)


def test_static_scratch_entry_rejects_oversized_requests():
    _lib, lib, dev, x, md = _setup()
    batch = 64
    need = int(lib.mgb_erfbin_scratch_bytes(K, batch))
    assert need > (1 << 20)              # beyond the 1 MiB built-in array
    rc, _ = _run(lib, _lib, md, [[-1.9, 0.3]] * batch, dev, static=True,
                 batch=batch)
    assert rc == -2                      # MGB_ETOOBIG
    msg = lib.mgb_last_error().decode()
    assert "scratch" in msg and str(need) in msg
    # and the same request is fine with caller-provided scratch
    rc, out = _run(lib, _lib, md, [[-1.9, 0.3]] * batch, dev, static=False,
                   batch=batch)
    assert rc == 0 and torch.isfinite(out["l"]).all()
