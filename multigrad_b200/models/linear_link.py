"""Native model family with a per-object *linear predictor*:

    mu_h = x_h + sum_j features[h, j] * theta_j,      sigma = link(p_sigma)

with ``params = (theta_0 .. theta_{F-1}, p_sigma)``.  The features are basis
functions of object properties (a constant column reproduces the plain shift
of :class:`SMFModel`; more columns give a mass- or environment-dependent
stellar-to-halo-mass relation).  ``d sumstats / d theta`` is a tall-skinny
GEMM over the object axis, which the CUDA path runs on the tensor cores
(``csrc/linlink.cu``: TMA-streamed bf16 feature tiles, the forward product on
ldmatrix + mma.sync inside the compute warps, the Jacobian product on tcgen05
UMMA with TMEM accumulators), fused with the erfc evaluation, the cross-GPU
reduction and the optimiser.  Features are stored in bf16; that quantised matrix *is* the
model (forward and Jacobian use the same values), and theta is split into
three bf16 planes so the forward product is accurate to fp32.

Per-object ``weights`` multiply every object's contribution.  More than
11 bins are evaluated in slices of at most 11 (one launch per slice over the
same packed feature matrix; loss and chain rule then run as small torch ops),
so any bin count works at ``ceil(K / 11)`` passes over the features.

The reference has no such model class; it is what a user of the reference
would write as a JAX sumstats function with a ``features @ theta`` term and
differentiate with ``jax.vjp`` (``multigrad/multigrad.py:508-538``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any

import numpy as np
import scipy.optimize
import torch

from .. import adam as _adam
from .. import util
from ..core import OnePointModel
from ..ops import reference as _ref
from ..util import GradDescentResult

__all__ = ["LinearLinkBinnedModel"]


def _np(v):
    if isinstance(v, torch.Tensor):
        return v.detach().cpu().double().numpy()
    return np.asarray(v, dtype=np.float64)


@dataclass(eq=False)
class LinearLinkBinnedModel(OnePointModel):
    x: Any = None
    features: Any = None
    bin_edges: Any = None
    volume: float = 1.0
    target_sumstats: Any = None
    sigma_link: str = "identity"
    loss: str = "log_mse"
    loss_eps: float = 0.0
    inv_err: Any = None
    weights: Any = None

    def __post_init__(self):
        super().__post_init__()
        if self.x is None or self.features is None or self.bin_edges is None:
            raise ValueError("need x, features and bin_edges")
        self.x = torch.as_tensor(self.x).detach().to(
            self.device, torch.float32).contiguous()
        # the bf16-rounded feature matrix is the model
        f = torch.as_tensor(self.features).detach().to(self.device)
        self.features = f.to(torch.bfloat16).to(torch.float32)
        if self.weights is not None:
            self.weights = torch.as_tensor(self.weights).detach().to(
                self.device, torch.float32).contiguous()
            if self.weights.shape != self.x.shape:
                raise ValueError("weights must have the shape of x")
        self.nfeat = int(self.features.shape[1])
        self.nparams = self.nfeat + 1
        self._edges64 = _np(self.bin_edges).ravel()
        self.nbins = self._edges64.size - 1
        self._norm64 = 1.0 / (float(self.volume) * np.diff(self._edges64))
        self._edges_t = torch.as_tensor(self._edges64, dtype=torch.float32,
                                        device=self.device)
        self._norm_t = torch.as_tensor(self._norm64, dtype=torch.float32,
                                       device=self.device)
        self._target_t = None
        self._inv_err_t = None
        self._plans = None
        self._host = None
        from ..ops.linlink import MAX_BINS
        nsl = -(-self.nbins // MAX_BINS)
        per = -(-self.nbins // nsl)
        self._slices = [(k, min(k + per, self.nbins))
                        for k in range(0, self.nbins, per)]
        self._builtin_loss = (type(self).calc_loss_from_sumstats
                              is LinearLinkBinnedModel.calc_loss_from_sumstats)
        self._builtin_sumstats = (
            type(self).calc_partial_sumstats_from_params
            is LinearLinkBinnedModel.calc_partial_sumstats_from_params)
        if self.target_sumstats is not None:
            self.set_target(self.target_sumstats, self.inv_err)

    @property
    def fused(self) -> bool:
        return self.device.type == "cuda" and self._builtin_sumstats

    @property
    def _one_launch(self) -> bool:
        """Loss, gradient and optimiser inside the kernel: one slice of bins
        and the built-in loss."""
        return self.fused and self._builtin_loss and len(self._slices) == 1

    def set_target(self, target, inv_err=None):
        self.target_sumstats = target
        self._target_t = torch.as_tensor(_np(target), dtype=torch.float32,
                                         device=self.device)
        if inv_err is not None:
            self.inv_err = inv_err
            self._inv_err_t = torch.as_tensor(
                np.broadcast_to(_np(inv_err), (self.nbins,)).copy(),
                dtype=torch.float32, device=self.device)
        if self._plans is not None and len(self._slices) == 1:
            self._plans[0].set_target(_np(target), None if inv_err is None
                                      else _np(inv_err))

    def plans(self):
        """One launch plan per slice of at most 11 bins, all over the same
        packed feature matrix."""
        if self._plans is None:
            from ..ops.linlink import LinLinkPlan, pack_features
            table = None
            if self.comm is not None and self.comm.size > 1:
                table = self.comm.workspace().table
            single = len(self._slices) == 1
            packed = pack_features(self.features, self.device)
            out = []
            for k0, k1 in self._slices:
                out.append(LinLinkPlan(
                    self.x, self.features, self._edges64[k0:k1 + 1],
                    self._norm64[k0:k1],
                    target=_np(self.target_sumstats) if single and
                    self.target_sumstats is not None else None,
                    inv_err=_np(self.inv_err) if single and
                    self.inv_err is not None else None,
                    sigma_link=self.sigma_link, loss=self.loss,
                    loss_eps=self.loss_eps, peer_table=table,
                    weights=self.weights, packed=packed))
            self._plans = out
        return self._plans

    def plan(self):
        return self.plans()[0]

    def _check(self):
        for pl in self.plans():
            pl.check_pipeline()
        self._check_peers()

    def _check_peers(self):
        if self.comm is not None and self.comm.size > 1:
            self.comm.workspace().check_error()

    def _eval(self, params, reduce=True):
        """``(sumstats[K], jac[K, P])`` from the kernel, all slices."""
        from ..ops._lib import MODE_SUMSTATS_JAC
        p = self._params(params)
        outs = [pl.evaluate(p, MODE_SUMSTATS_JAC, reduce=reduce)
                for pl in self.plans()]
        if len(outs) == 1:
            return outs[0]["sumstats"], outs[0]["jac"]
        return (torch.cat([o["sumstats"] for o in outs]),
                torch.cat([o["jac"] for o in outs], dim=0))

    # ---- the two hooks as differentiable torch (CPU path and GPU oracle) ---
    def calc_partial_sumstats_from_params(self, params, randkey=None):
        p = params if isinstance(params, torch.Tensor) else \
            self._params(params)
        if self.fused and not p.requires_grad:
            return self._eval(p, reduce=False)[0]
        p = p.to(self.device)
        dt = p.dtype
        shift = self.features.to(dt) @ p[:self.nfeat]
        return _ref.erf_binned_partial(
            self.x.to(dt) + shift, self._edges_t.to(dt), self._norm_t.to(dt),
            torch.zeros((), dtype=dt, device=p.device), p[self.nfeat],
            self.sigma_link,
            None if self.weights is None else self.weights.to(dt))

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                randkey=None):
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        s = torch.as_tensor(sumstats)
        return _ref.binned_loss(
            s, self._target_t.to(s), self.loss, self.loss_eps,
            None if self._inv_err_t is None else self._inv_err_t.to(s))

    # ---- fused paths --------------------------------------------------------
    def calc_sumstats_from_params(self, params, total=True, randkey=None):
        if not self.fused:
            return super().calc_sumstats_from_params(params, total, randkey)
        return self._eval(params, reduce=total)[0]

    def calc_sumstats_and_jac_from_params(self, params, total=True):
        if not self.fused:
            p = self._params(params)
            jac = torch.autograd.functional.jacobian(
                lambda q: self.calc_partial_sumstats_from_params(q), p)
            s = self.calc_partial_sumstats_from_params(p)
            if total:
                from ..parallel.collectives import reduce_sum
                s, jac = reduce_sum(s, comm=self.comm), \
                    reduce_sum(jac, comm=self.comm)
            return s, jac
        return self._eval(params, reduce=total)

    def calc_loss_from_params(self, params, randkey=None):
        if not self.fused:
            return super().calc_loss_from_params(params, randkey)
        with torch.no_grad():
            return self.calc_loss_from_sumstats(
                self.calc_sumstats_from_params(params))

    def _vjp(self, params, randkey=None, include_loss=True):
        if not self.fused:
            return super()._vjp(params, randkey, include_loss)
        p = self._params(params)
        if self._one_launch:
            from ..ops._lib import MODE_LOSS_GRAD
            out = self.plan().evaluate(p, MODE_LOSS_GRAD)
            loss, grad = out["loss"], out["grad"]
        else:
            s, jac = self._eval(p)
            kw = {} if randkey is None else {"randkey": randkey}
            loss, dlds = self._loss_and_dloss(s, None, kw)
            grad = dlds.to(jac) @ jac
        return (loss, grad) if include_loss else grad

    # ---- host-driven evaluation --------------------------------------------
    def loss_and_grad_host(self, params_host) -> "tuple[float, np.ndarray]":
        """Loss and gradient for host parameters.  With one slice of bins and
        the built-in loss: ONE launch that reads the parameters straight from
        pinned host memory and writes ``(loss, grad)`` straight back into
        pinned host memory (zero copy over PCIe), then a stream synchronise --
        no copy-engine operation.  Otherwise the slices are evaluated on the
        device and the result is copied back.  Collective (every rank gets
        the same bits)."""
        if not self._one_launch:
            loss, grad = self._vjp(torch.as_tensor(
                np.asarray(params_host, dtype=np.float32)))
            return (float(loss),
                    grad.detach().cpu().double().numpy())
        from ..ops._lib import MODE_LOSS_GRAD
        from ..ops.linlink import MgbLinArgs
        plan = self.plan()
        if self._host is None:
            P = self.nparams
            h_in = torch.zeros(P, dtype=torch.float32).pin_memory()
            h_out = torch.zeros(1 + P, dtype=torch.float32).pin_memory()
            f32 = dict(dtype=torch.float32, device=self.device)
            d_s = torch.empty(self.nbins, **f32)
            d_j = torch.empty(self.nbins, P, **f32)
            a = MgbLinArgs()
            a.mode = MODE_LOSS_GRAD
            a.reduce = 1 if plan.world > 1 else 0
            a.params = h_in.data_ptr()
            a.out_sumstats, a.out_jac = d_s.data_ptr(), d_j.data_ptr()
            a.out_loss = h_out.data_ptr()
            a.out_grad = h_out.data_ptr() + 4
            self._host = (h_in, h_out, d_s, d_j, a, h_in.numpy(),
                          h_out.numpy())
        h = self._host
        h[5][:] = params_host
        plan.launch(h[4])
        torch.cuda.current_stream(self.device).synchronize()
        return float(h[6][0]), h[6][1:].astype(np.float64)

    def _host_session(self, warm_params):
        """``(evaluate, end)`` for the replicated L-BFGS-B driver
        (``models/bfgs_device.py``)."""
        self.loss_and_grad_host(np.asarray(warm_params, dtype=np.float64))
        return self.loss_and_grad_host, (lambda: None)

    def device_loss_grad_launcher(self, params: torch.Tensor,
                                  out: torch.Tensor):
        """Zero-argument callable that enqueues ONE fused loss+gradient launch
        reading the CUDA tensor ``params`` (at execution time) and writing
        ``(loss, grad)`` into the CUDA tensor ``out`` (``[1 + P]``); ``None``
        when the model cannot run that way.  This is what lets the model join
        a device-resident ``OnePointGroup`` fit (``ops/group.py``)."""
        if not self._one_launch:
            return None
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        from ..ops._lib import MODE_LOSS_GRAD
        from ..ops.linlink import MgbLinArgs
        plan = self.plan()
        f32 = dict(dtype=torch.float32, device=self.device)
        d_s = torch.empty(self.nbins, **f32)
        d_j = torch.empty(self.nbins, self.nparams, **f32)
        a = MgbLinArgs()
        a.mode = MODE_LOSS_GRAD
        a.reduce = 1 if plan.world > 1 else 0
        a.params = params.data_ptr()
        a.out_sumstats, a.out_jac = d_s.data_ptr(), d_j.data_ptr()
        a.out_loss = out.data_ptr()
        a.out_grad = out.data_ptr() + 4

        def launch(_keep=(params, out, d_s, d_j)):
            plan.launch(a)
        return launch

    # ---- optimisers ---------------------------------------------------------
    def run_bfgs(self, guess, maxsteps=100, param_bounds=None, randkey=None,
                 comm=None):
        """L-BFGS-B (scipy, float64, on the host) over :meth:`loss_and_grad_host`.
        With several GPUs every rank runs the same deterministic optimiser on
        the bit-identical reduced values (``models/bfgs_device.py``): no
        parameter hand-off and no result broadcast."""
        if not self.fused:
            return super().run_bfgs(guess, maxsteps, param_bounds, randkey,
                                    comm)
        comm = self.comm if comm is None else comm
        if comm is not None and comm.size > 1:
            from .bfgs_device import run_bfgs_replicated
            return run_bfgs_replicated(self, guess, maxsteps, param_bounds,
                                       comm)
        if param_bounds is not None and hasattr(param_bounds, "tolist"):
            param_bounds = param_bounds.tolist()
        x0 = _np(guess).ravel()
        pbar = util.trange(maxsteps, desc="BFGS Gradient Descent Progress",
                           leave=True)

        def callback(*_a, **_k):
            if hasattr(pbar, "update"):
                pbar.update()
        try:
            return scipy.optimize.minimize(
                self.loss_and_grad_host, x0=x0, method="L-BFGS-B", jac=True,
                options=dict(maxiter=maxsteps), callback=callback,
                bounds=param_bounds)
        finally:
            if hasattr(pbar, "close"):
                pbar.close()

    def run_adam_host(self, guess, nsteps=100, learning_rate=0.01,
                      init_state=None, return_state=False,
                      return_loss=False):
        """Adam with the optimiser state on the host, like the reference
        (``multigrad/adam.py:52-68``): per step one :meth:`loss_and_grad_host`
        and a float32 update on the CPU.  :meth:`run_adam` keeps the whole fit
        on the device and is the faster choice when nothing has to see the
        per-step values on the host.  Unbounded parameters only."""
        P = self.nparams
        p = _np(guess).ravel().astype(np.float32)
        step0 = 0
        m, v = np.zeros(P, np.float32), np.zeros(P, np.float32)
        if init_state is not None:
            step0 = int(init_state.step)
            m[:], v[:] = _np(init_state.m), _np(init_state.v)
        traj = np.zeros((nsteps + 1, P), np.float32)
        losses = np.zeros(max(nsteps, 1), np.float32)
        traj[0] = p
        b1, b2, eps = (np.float32(_adam.ADAM_B1), np.float32(_adam.ADAM_B2),
                       np.float32(_adam.ADAM_EPS))
        lr = np.float32(learning_rate)
        for i in range(nsteps):
            loss, g = self.loss_and_grad_host(p)
            g = g.astype(np.float32)
            t = step0 + i + 1
            m = b1 * m + (np.float32(1) - b1) * g
            v = b2 * v + (np.float32(1) - b2) * g * g
            mhat = m / np.float32(1.0 - float(b1) ** t)
            vhat = v / np.float32(1.0 - float(b2) ** t)
            p = (p - lr * mhat / (np.sqrt(vhat) + eps)).astype(np.float32)
            traj[i + 1], losses[i] = p, loss
        out = [torch.from_numpy(traj)]
        if return_loss:
            out.append(torch.from_numpy(losses[:nsteps]))
        if return_state:
            out.append(_adam.AdamState(step0 + nsteps, torch.from_numpy(m),
                                       torch.from_numpy(v)))
        return out[0] if len(out) == 1 else tuple(out)

    def run_adam(self, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, randkey=None, const_randkey=False,
                 comm=None, reference_compat=True, init_state=None,
                 return_state=False, return_loss=False):
        """Adam, one fused tensor-core launch per step."""
        if not self._one_launch:
            return super().run_adam(
                guess, nsteps, param_bounds, learning_rate, randkey,
                const_randkey, comm, reference_compat, init_state,
                return_state, return_loss)
        from ..ops._lib import MODE_ADAM
        from ..ops.linlink import LinFitState
        plan = self.plan()
        p0 = self._params(guess)
        bounds, u0 = None, p0
        if param_bounds is not None:
            bounds = _adam.normalize_bounds(param_bounds, p0.numel())
            u0 = _adam.apply_transforms(p0.cpu(), bounds).to(self.device)
        step0, m, v = 0, None, None
        if init_state is not None:
            step0, m, v = int(init_state.step), init_state.m, init_state.v
        st = LinFitState(p0, u0, nsteps, self.device, step0=step0, m=m, v=v)
        st.traj[0].copy_(p0)
        args = plan.make_optimizer_args(MODE_ADAM, st, learning_rate, bounds,
                                        jac_at_params=reference_compat)
        for _ in range(nsteps):
            plan.launch(args)
        self._check()
        out = [st.traj]
        if return_loss:
            out.append(st.loss_hist[:nsteps])
        if return_state:
            out.append(_adam.AdamState(step0 + nsteps, st.m, st.v))
        return out[0] if len(out) == 1 else tuple(out)

    def run_simple_grad_descent(self, guess, nsteps=100, learning_rate=0.01):
        if not self._one_launch:
            return super().run_simple_grad_descent(guess, nsteps,
                                                   learning_rate)
        from ..ops._lib import MODE_GD
        from ..ops.linlink import LinFitState
        plan = self.plan()
        p0 = self._params(guess)
        st = LinFitState(p0, p0, nsteps, self.device, rows=max(nsteps, 1))
        args = plan.make_optimizer_args(MODE_GD, st, learning_rate)
        for _ in range(nsteps):
            plan.launch(args)
        self._check()
        return GradDescentResult(loss=st.loss_hist[:nsteps],
                                 params=st.traj[:nsteps],
                                 aux=[None] * nsteps)
