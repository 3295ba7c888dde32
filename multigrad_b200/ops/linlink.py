"""Python face of the tensor-core linear-link kernel (``csrc/linlink.cu``)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from .reference import LOSS_KINDS, SIGMA_LINKS

__all__ = ["LinLinkPlan", "LinFitState", "launch_count", "MgbLinModel",
           "MgbLinArgs", "pack_features", "MAX_BINS"]

MAX_BINS = 11
_LAUNCHES = 0


def launch_count() -> int:
    return _LAUNCHES


class MgbLinModel(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("ow", C.c_void_p),
        ("n", C.c_int64),
        ("nbins", C.c_int32), ("nfeat", C.c_int32), ("pfeat", C.c_int32),
        ("sigma_link", C.c_int32), ("loss_kind", C.c_int32),
        ("loss_eps", C.c_float), ("x0", C.c_float),
        ("edges", C.c_float * (MAX_BINS + 1)),
        ("norm", C.c_float * MAX_BINS),
        ("target", C.c_float * MAX_BINS),
        ("inv_err", C.c_float * MAX_BINS),
    ]


class MgbLinArgs(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("reduce", C.c_int32),
        ("jac_at_params", C.c_int32), ("traj_rows", C.c_int32),
        ("step_base", C.c_int32),
        ("lr", C.c_float), ("b1", C.c_float), ("b2", C.c_float),
        ("eps", C.c_float),
        ("params", C.c_void_p),
        ("out_sumstats", C.c_void_p), ("out_jac", C.c_void_p),
        ("out_loss", C.c_void_p), ("out_grad", C.c_void_p),
        ("opt_p", C.c_void_p), ("opt_u", C.c_void_p),
        ("opt_m", C.c_void_p), ("opt_v", C.c_void_p),
        ("opt_step", C.c_void_p),
        ("traj", C.c_void_p), ("loss_hist", C.c_void_p),
        ("bound_kind", C.c_void_p), ("bound_lo", C.c_void_p),
        ("bound_hi", C.c_void_p),
    ]


def _declare(lib):
    if getattr(lib, "_linlink_declared", False):
        return
    lib.mgb_linlink_scratch_bytes.restype = C.c_uint64
    lib.mgb_linlink_scratch_bytes.argtypes = [C.c_int, C.c_int]
    lib.mgb_linlink_step.restype = C.c_int
    lib.mgb_linlink_step.argtypes = [
        C.POINTER(MgbLinModel), C.POINTER(MgbLinArgs),
        C.POINTER(_lib.MgbPeerTable), C.c_void_p, C.c_uint64, C.c_void_p]
    lib._linlink_declared = True


class LinFitState:
    """Device-resident optimiser state + trajectory / loss history."""

    def __init__(self, params, uparams, nsteps, device, step0=0, m=None,
                 v=None, rows=None):
        n = params.numel()
        f32 = dict(dtype=torch.float32, device=device)
        self.p = params.detach().to(**f32).clone().contiguous()
        self.u = uparams.detach().to(**f32).clone().contiguous()
        self.m = torch.zeros(n, **f32) if m is None else \
            m.detach().to(**f32).clone()
        self.v = torch.zeros(n, **f32) if v is None else \
            v.detach().to(**f32).clone()
        self.step = torch.full((1,), step0, dtype=torch.int32, device=device)
        self.step_base = int(step0)
        rows = (nsteps + 1) if rows is None else rows
        self.traj = torch.zeros(rows, n, **f32)
        self.loss_hist = torch.zeros(rows, **f32)


def pack_features(features: torch.Tensor, device) -> torch.Tensor:
    """The kernel's feature layout: bf16 ``[n][pfeat]``, ``pfeat`` = 64 or 128,
    zero padded.  Build it once and hand it to several plans (``packed=``)
    when one catalogue serves more than one launch plan."""
    n, nfeat = features.shape
    if nfeat > 128:
        raise ValueError("at most 128 features")
    pfeat = 64 if nfeat <= 64 else 128
    w = torch.zeros(max(n, 1), pfeat, dtype=torch.bfloat16, device=device)
    w[:n, :nfeat] = features.detach().to(device).to(torch.bfloat16)
    return w


def _aligned_f32(t: torch.Tensor, device) -> torch.Tensor:
    t = t.detach().to(device, torch.float32).contiguous()
    if t.data_ptr() % 16:
        t = t.clone()
    return t


class LinLinkPlan:
    """Launch plan: features are stored once as bf16 ``[n][pfeat]`` with
    ``pfeat`` = 64 or 128 (zero padded); optional per-object ``weights``."""

    def __init__(self, x: torch.Tensor, features: torch.Tensor, edges, norm,
                 target=None, inv_err=None, sigma_link="identity",
                 loss="log_mse", loss_eps=0.0, peer_table=None, weights=None,
                 packed: Optional[torch.Tensor] = None):
        if not x.is_cuda:
            raise ValueError("LinLinkPlan needs CUDA tensors")
        self.lib = _lib.load()
        _declare(self.lib)
        self.device = x.device
        n, nfeat = features.shape
        if x.numel() != n:
            raise ValueError("x and features disagree on the object count")
        w = pack_features(features, self.device) if packed is None else packed
        pfeat = int(w.shape[1])
        if (w.dtype != torch.bfloat16 or w.shape[0] != max(n, 1)
                or pfeat not in (64, 128) or nfeat > pfeat):
            raise ValueError("packed features do not match (n, nfeat)")
        x = _aligned_f32(x, self.device)
        self.x, self.w = x, w
        self.ow = None
        if weights is not None:
            if weights.numel() != n:
                raise ValueError("weights and x disagree on the object count")
            self.ow = _aligned_f32(weights, self.device)
        edges = np.asarray(edges, dtype=np.float64).ravel()
        K = edges.size - 1
        if K < 1 or K > MAX_BINS:
            raise ValueError(f"need 1..{MAX_BINS} bins, got {K}")
        self.K, self.nfeat, self.pfeat, self.NP = K, nfeat, pfeat, nfeat + 1
        md = MgbLinModel()
        md.x, md.w, md.n = x.data_ptr(), w.data_ptr(), n
        md.ow = None if self.ow is None else self.ow.data_ptr()
        md.nbins, md.nfeat, md.pfeat = K, nfeat, pfeat
        md.sigma_link = SIGMA_LINKS[sigma_link]
        md.loss_kind = LOSS_KINDS[loss]
        md.loss_eps = float(loss_eps)
        x0 = float(x.double().mean().item()) if n else 0.0
        md.x0 = round(x0 * 64.0) / 64.0
        norm = np.broadcast_to(np.asarray(norm, np.float64), (K,))
        tgt = np.ones(K) if target is None else np.asarray(target, np.float64)
        ierr = np.ones(K) if inv_err is None else \
            np.broadcast_to(np.asarray(inv_err, np.float64), (K,))
        for k in range(K + 1):
            md.edges[k] = float(edges[k])
        for k in range(K):
            md.norm[k], md.target[k] = float(norm[k]), float(tgt[k])
            md.inv_err[k] = float(ierr[k])
        self.model = md
        self.pt = peer_table
        self._pt_ref = C.byref(peer_table) if peer_table is not None else None
        self.world = peer_table.world if peer_table is not None else 1
        with torch.cuda.device(self.device):
            nbytes = int(self.lib.mgb_linlink_scratch_bytes(K, pfeat))
        if nbytes == 0:
            raise _lib.MgbError("mgb_linlink_scratch_bytes returned 0")
        self.scratch = torch.zeros(nbytes, dtype=torch.uint8,
                                   device=self.device)

    def set_target(self, target, inv_err=None):
        tgt = np.asarray(target, np.float64).ravel()
        for k in range(self.K):
            self.model.target[k] = float(tgt[k])
        if inv_err is not None:
            ie = np.broadcast_to(np.asarray(inv_err, np.float64), (self.K,))
            for k in range(self.K):
                self.model.inv_err[k] = float(ie[k])

    def launch(self, args: MgbLinArgs) -> None:
        global _LAUNCHES
        with torch.cuda.device(self.device):
            rc = self.lib.mgb_linlink_step(
                C.byref(self.model), C.byref(args), self._pt_ref,
                self.scratch.data_ptr(), self.scratch.numel(),
                C.c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "mgb_linlink_step")
        _LAUNCHES += 1

    def check_pipeline(self) -> None:
        """Raise if an in-kernel pipeline wait timed out (sticky flag)."""
        flag = int(self.scratch[4:8].view(torch.int32).item())
        if flag:
            raise RuntimeError("linlink kernel: an mbarrier wait timed out")

    def evaluate(self, params: torch.Tensor, mode: int, reduce: bool = True):
        p = params.detach().to(self.device, torch.float32).contiguous()
        assert p.numel() == self.NP
        f32 = dict(dtype=torch.float32, device=self.device)
        out = {"sumstats": torch.empty(self.K, **f32)}
        a = MgbLinArgs()
        a.mode = mode
        a.reduce = 1 if (reduce and self.world > 1) else 0
        a.params = p.data_ptr()
        a.out_sumstats = out["sumstats"].data_ptr()
        out["jac"] = torch.zeros(self.K, self.NP, **f32)
        a.out_jac = out["jac"].data_ptr()
        if mode >= _lib.MODE_LOSS_GRAD:
            out["loss"] = torch.empty(1, **f32)
            out["grad"] = torch.empty(self.NP, **f32)
            a.out_loss, a.out_grad = (out["loss"].data_ptr(),
                                      out["grad"].data_ptr())
        self.launch(a)
        out["_keepalive"] = p
        if "loss" in out:
            out["loss"] = out["loss"][0]
        return out

    def make_optimizer_args(self, mode, state: LinFitState, lr,
                            bounds: Optional[Sequence] = None,
                            jac_at_params=True, b1=0.9, b2=0.999, eps=1e-8):
        a = MgbLinArgs()
        a.mode = mode
        a.reduce = 1 if self.world > 1 else 0
        a.jac_at_params = 1 if jac_at_params else 0
        a.opt_p, a.opt_u = state.p.data_ptr(), state.u.data_ptr()
        a.opt_m, a.opt_v = state.m.data_ptr(), state.v.data_ptr()
        a.opt_step = state.step.data_ptr()
        a.traj, a.loss_hist = state.traj.data_ptr(), state.loss_hist.data_ptr()
        a.traj_rows, a.step_base = state.traj.shape[0], state.step_base
        a.lr, a.b1, a.b2, a.eps = float(lr), float(b1), float(b2), float(eps)
        if bounds is not None:
            kind = np.zeros(self.NP, np.int32)
            lo = np.zeros(self.NP, np.float32)
            hi = np.zeros(self.NP, np.float32)
            for i, bd in enumerate(bounds):
                if bd is None:
                    continue
                l, h = bd
                if l is not None and h is not None:
                    kind[i], lo[i], hi[i] = 1, l, h
                elif l is not None:
                    kind[i], lo[i] = 2, l
                elif h is not None:
                    kind[i], hi[i] = 3, h
            state._bk = torch.as_tensor(kind, device=self.device)
            state._bl = torch.as_tensor(lo, device=self.device)
            state._bh = torch.as_tensor(hi, device=self.device)
            a.bound_kind = state._bk.data_ptr()
            a.bound_lo, a.bound_hi = state._bl.data_ptr(), state._bh.data_ptr()
        return a
