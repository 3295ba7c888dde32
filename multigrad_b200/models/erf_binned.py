"""Native model family: erf-soft-binned one-point statistics.

``ErfBinnedModel`` is the B200-native counterpart of the stellar-mass-function
models the reference ships as examples (``tests/smf_example/
smf_grad_descent.py:17-82`` -- ``sigma = p1`` -- and ``docs/source/notebooks/
smf_gradient_descent.py:18-91`` -- ``sigma = 10**p1``, ``+1e-10`` in the
logs): every object ``h`` with coordinate ``x_h`` spreads a Gaussian of mean
``mu_h = x_h + params[mu_index]`` and width ``sigma = link(params[sigma_index])``
over the bins, and the summary statistic is the (volume- and width-normalised)
expected count per bin.

On CUDA every public method runs the fused kernel of ``csrc/erfbin.cu``: one
launch per loss+gradient evaluation or optimiser step, with the cross-GPU
reduction done inside the kernel over NVLink (no NCCL/MPI on the hot path).
On CPU the same mathematics runs as differentiable PyTorch through the generic
:class:`~multigrad_b200.core.OnePointModel` machinery, which is also the
oracle the GPU tests compare against.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any

import numpy as np
import scipy.optimize
import torch

from .. import adam as _adam
from .. import util
from ..core import OnePointModel
from ..ops import reference as _ref
from ..parallel.collectives import reduce_sum
from ..util import GradDescentResult

__all__ = ["ErfBinnedModel", "SMFModel", "AdamFit"]


class AdamFit:
    """Handle of a device-resident Adam fit (see
    :meth:`ErfBinnedModel.adam_fit`).  Parameters, moments, trajectory and
    loss history live on the GPU and are written by the fused kernel; nothing
    here synchronises with the device."""

    def __init__(self, model, plan, state, args, step0, nsteps):
        self.model, self._plan, self._st, self._args = model, plan, state, args
        self._step0, self.nsteps, self.done = step0, nsteps, 0

    def step(self, n: int = 1, use_graph=None) -> "AdamFit":
        """Enqueue the next ``n`` optimiser steps on the current stream."""
        n = int(n)
        if n < 0 or self.done + n > self.nsteps:
            raise ValueError(f"fit has room for {self.nsteps} steps, "
                             f"{self.done} done, {n} more asked for")
        if n == 1:
            self._args.nsteps = 1
            self._plan.launch(self._args)
        elif n > 1:
            self.model._run_steps(self._plan, self._args, n, use_graph)
        self.done += n
        return self

    @property
    def trajectory(self) -> torch.Tensor:
        """``(nsteps+1, ndim)``; rows ``0..done`` are valid."""
        return self._st.traj

    @property
    def losses(self) -> torch.Tensor:
        """Loss at the start of every completed step, ``(done,)``."""
        return self._st.loss_hist[:self.done]

    @property
    def params(self) -> torch.Tensor:
        """Current parameters (CUDA tensor, updated in place by the kernel)."""
        return self._st.p

    @property
    def state(self):
        return _adam.AdamState(self._step0 + self.done, self._st.m,
                               self._st.v)


def _np(v):
    if isinstance(v, torch.Tensor):
        return v.detach().cpu().double().numpy()
    return np.asarray(v, dtype=np.float64)


@dataclass(eq=False)
class ErfBinnedModel(OnePointModel):
    """Soft-binned statistic with a shift link for the mean and a scalar
    width.

    Parameters (besides those of :class:`OnePointModel`)
    ----------------------------------------------------
    x : tensor, this rank's shard of object coordinates (e.g. log10 M_halo)
    bin_edges : K+1 increasing edges
    volume : normalisation volume (``s_k = counts_k / volume / width_k``)
    target_sumstats : K targets for the built-in losses
    weights : optional per-object weights
    sigma_link : ``"identity"`` or ``"log10"``
    loss : ``"log_mse"`` | ``"mse"`` | ``"chi2"``;  loss_eps, inv_err
    nparams, mu_index, sigma_index : layout of the parameter vector
    deterministic : ``True`` = bit-reproducible results from run to run (fully
        static work split in the kernel, 2-3% slower; also
        ``MULTIGRAD_B200_DETERMINISTIC=1``).  The replicas on different GPUs
        agree bit for bit either way.

    ``aux_data`` may carry the same information under the reference example's
    keys (``log_halo_masses``, ``smf_bin_edges``, ``volume``,
    ``target_sumstats``).
    """
    x: Any = None
    bin_edges: Any = None
    volume: float = 1.0
    target_sumstats: Any = None
    weights: Any = None
    sigma_link: str = "identity"
    loss: str = "log_mse"
    loss_eps: float = 0.0
    inv_err: Any = None
    nparams: int = 2
    mu_index: int = 0
    sigma_index: int = 1
    deterministic: bool = False

    def __post_init__(self):
        super().__post_init__()
        aux = self.aux_data if isinstance(self.aux_data, dict) else {}
        if self.x is None:
            self.x = aux.get("log_halo_masses", aux.get("x"))
        if self.bin_edges is None:
            self.bin_edges = aux.get("smf_bin_edges", aux.get("bin_edges"))
        if self.target_sumstats is None:
            self.target_sumstats = aux.get("target_sumstats")
        if "volume" in aux and self.volume == 1.0:
            self.volume = float(aux["volume"])
        if self.weights is None:
            self.weights = aux.get("weights")
        if self.x is None or self.bin_edges is None:
            raise ValueError("ErfBinnedModel needs `x` and `bin_edges`")
        if isinstance(self.x, torch.Tensor) and self.x.is_cuda \
                and self.device.type != "cuda":
            self.device = self.x.device
        self.x = torch.as_tensor(self.x).detach().to(
            self.device, torch.float32).contiguous()
        if self.weights is not None:
            self.weights = torch.as_tensor(self.weights).detach().to(
                self.device, torch.float32).contiguous()
        self._edges64 = _np(self.bin_edges).ravel()
        self.nbins = self._edges64.size - 1
        self._norm64 = 1.0 / (float(self.volume) * np.diff(self._edges64))
        self._edges_t = torch.as_tensor(self._edges64, dtype=torch.float64,
                                        device=self.device)
        self._norm_t = torch.as_tensor(self._norm64, dtype=torch.float32,
                                       device=self.device)
        self._target_t = None
        self._inv_err_t = None
        self._builtin_loss = (type(self).calc_loss_from_sumstats
                              is ErfBinnedModel.calc_loss_from_sumstats)
        self._builtin_sumstats = (
            type(self).calc_partial_sumstats_from_params
            is ErfBinnedModel.calc_partial_sumstats_from_params)
        self._plan = None
        self._host = None
        if self.target_sumstats is not None:
            self.set_target(self.target_sumstats, self.inv_err)

    # ------------------------------------------------------------------ setup
    @property
    def fused(self) -> bool:
        """True when the CUDA kernels are the execution path: CUDA device and
        the built-in statistic (a subclass that overrides
        ``calc_partial_sumstats_from_params`` is evaluated by the generic
        autograd path of :class:`OnePointModel`, never silently by the
        built-in kernel)."""
        return self.device.type == "cuda" and self._builtin_sumstats

    def set_target(self, target, inv_err=None) -> None:
        """(Re)define the target summary statistics of the built-in loss."""
        self.target_sumstats = target
        self._target_t = torch.as_tensor(_np(target), dtype=torch.float32,
                                         device=self.device)
        if inv_err is not None:
            self.inv_err = inv_err
            self._inv_err_t = torch.as_tensor(
                np.broadcast_to(_np(inv_err), (self.nbins,)).copy(),
                dtype=torch.float32, device=self.device)
        if self._plan is not None:
            self._plan.set_target(_np(target), None if inv_err is None
                                  else _np(inv_err))

    def plan(self):
        """The launch plan (built on first use: allocating the communicator's
        peer workspace is a collective call)."""
        if self._plan is None:
            from ..ops.erfbin import ErfBinPlan
            table = None
            if self.comm is not None and self.comm.size > 1:
                table = self.comm.workspace().table
            self._plan = ErfBinPlan(
                self.x, self._edges64, self._norm64,
                target=None if self.target_sumstats is None
                else _np(self.target_sumstats),
                inv_err=None if self.inv_err is None else _np(self.inv_err),
                nparams=self.nparams, mu_idx=self.mu_index,
                sig_idx=self.sigma_index, sigma_link=self.sigma_link,
                loss=self.loss, loss_eps=self.loss_eps, weights=self.weights,
                peer_table=table)
            self._plan.deterministic = bool(self.deterministic)
            self.x = self._plan.x
        return self._plan

    def _check_peers(self):
        if self.comm is not None and self.comm.size > 1:
            self.comm.workspace().check_error()

    # --------------------------------------------------- user-hook equivalents
    def calc_partial_sumstats_from_params(self, params, randkey=None):
        """This rank's additive share of the binned statistic.  Differentiable
        PyTorch when ``params`` requires grad (oracle / generic path), the
        forward kernel otherwise."""
        p = params if isinstance(params, torch.Tensor) else \
            self._params(params)
        if self.fused and not p.requires_grad:
            from ..ops.erfbin import MODE_SUMSTATS
            return self.plan().evaluate(p, MODE_SUMSTATS,
                                        reduce=False)["sumstats"]
        p = p.to(self.device)
        return _ref.erf_binned_partial(
            self.x.to(p.dtype), self._edges_t.to(p.dtype),
            self._norm_t.to(p.dtype), p[self.mu_index], p[self.sigma_index],
            self.sigma_link,
            None if self.weights is None else self.weights.to(p.dtype))

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                randkey=None):
        """Built-in loss (override for a custom differentiable loss; the fused
        path then returns sumstats + Jacobian and applies autograd to it)."""
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        s = torch.as_tensor(sumstats)
        return _ref.binned_loss(
            s, self._target_t.to(s), self.loss, self.loss_eps,
            None if self._inv_err_t is None else self._inv_err_t.to(s))

    # ------------------------------------------------------------ fused paths
    def calc_sumstats_from_params(self, params, total=True, randkey=None):
        if not self.fused:
            return super().calc_sumstats_from_params(params, total, randkey)
        from ..ops.erfbin import MODE_SUMSTATS
        p = self._params(params)
        if p.dim() == 2 and total and self.comm.size > 1:
            part = self.plan().evaluate(p, MODE_SUMSTATS, reduce=False)
            return reduce_sum(part["sumstats"], comm=self.comm)
        return self.plan().evaluate(p, MODE_SUMSTATS, reduce=total)["sumstats"]

    def calc_sumstats_and_jac_from_params(self, params, total=True):
        """``(s[K], ds/dp[K, P])`` in one pass (no reference counterpart: the
        reference never materialises the Jacobian)."""
        p = self._params(params)
        if self.fused:
            from ..ops.erfbin import MODE_SUMSTATS_JAC
            out = self.plan().evaluate(p, MODE_SUMSTATS_JAC, reduce=total)
            return out["sumstats"], out["jac"]
        s, dmu, dsg = _ref.erf_binned_partial_and_jac(
            self.x, self._edges_t.to(self.x.dtype), self._norm_t,
            p[self.mu_index],
            p[self.sigma_index], self.sigma_link, self.weights)
        jac = torch.zeros(self.nbins, self.nparams, dtype=s.dtype,
                          device=s.device)
        jac[:, self.mu_index] += dmu
        jac[:, self.sigma_index] += dsg
        if total:
            s = reduce_sum(s, comm=self.comm)
            jac = reduce_sum(jac, comm=self.comm)
        return s, jac

    def calc_loss_from_params(self, params, randkey=None):
        if not self.fused:
            return super().calc_loss_from_params(params, randkey)
        s = self.calc_sumstats_from_params(params)
        with torch.no_grad():
            return self.calc_loss_from_sumstats(s)

    def _vjp(self, params, randkey=None, include_loss=True):
        if not self.fused:
            return super()._vjp(params, randkey, include_loss)
        from ..ops.erfbin import MODE_LOSS_GRAD, MODE_SUMSTATS_JAC
        p = self._params(params)
        if self._builtin_loss:
            if self._target_t is None:
                raise ValueError("no target_sumstats given")
            out = self.plan().evaluate(p, MODE_LOSS_GRAD)
            loss, grad = out["loss"], out["grad"]
        else:
            out = self.plan().evaluate(p, MODE_SUMSTATS_JAC)
            kw = {} if randkey is None else {"randkey": randkey}
            loss, dlds = self._loss_and_dloss(out["sumstats"], None, kw)
            grad = dlds.to(out["jac"]) @ out["jac"]
        return (loss, grad) if include_loss else grad

    # --------------------------------------------------------- host fast path
    def loss_and_grad_host(self, params_host) -> "tuple[float, np.ndarray]":
        """Loss+gradient for host parameters with minimal host<->device
        traffic, in ONE native call.  Default (``mgb_erfbin_eval_host_zc``):
        the fused kernel reads the ``P`` floats straight from a pinned host
        buffer and writes its ``1+P`` result floats straight into another one
        (zero copy over PCIe), the host polls for them: one launch, no
        copy-engine operation, no stream synchronise.  With
        ``MULTIGRAD_B200_ZEROCOPY=0`` (``mgb_erfbin_eval_host``): pinned H2D
        copy, launch, pinned D2H copy, stream sync.  Used by ``run_bfgs`` and
        by the end-to-end benchmark."""
        import ctypes as C
        plan = self.plan()
        if self._host is None:
            P = self.nparams
            h_in = torch.empty(P, dtype=torch.float32).pin_memory()
            # loss, gradient, + the hand-shake word of the pipelined host loop
            # (word 16, in its own cache line: MGB_HOST_FLAG_WORD)
            h_out = torch.zeros(32, dtype=torch.float32).pin_memory()
            d_in = torch.empty(P, dtype=torch.float32, device=self.device)
            d_out = torch.empty(1 + P, dtype=torch.float32,
                                device=self.device)
            scratch = plan._scratch_for(1)
            zc = os.environ.get("MULTIGRAD_B200_ZEROCOPY", "1") != "0"
            if zc:
                args = (C.byref(plan.model), plan._pt_ref,
                        C.c_void_p(h_in.data_ptr()),
                        C.c_void_p(h_out.data_ptr()),
                        C.c_void_p(scratch.data_ptr()), scratch.numel())
            else:
                args = (C.byref(plan.model), plan._pt_ref,
                        C.c_void_p(h_in.data_ptr()),
                        C.c_void_p(h_out.data_ptr()),
                        C.c_void_p(d_in.data_ptr()),
                        C.c_void_p(d_out.data_ptr()),
                        C.c_void_p(scratch.data_ptr()), scratch.numel())
            self._host = (h_in, h_out, d_in, d_out, scratch, h_in.numpy(),
                          h_out.numpy(), args, zc)
        h = self._host
        h[5][:] = params_host
        if torch.cuda.current_device() != self.device.index:
            torch.cuda.set_device(self.device)
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if h[8]:
            timeout = 60000
            if plan.pt is not None:
                timeout = max(timeout, 2 * int(plan.pt.spin_timeout_ms))
            rc = plan.lib.mgb_erfbin_eval_host_zc(*h[7], stream, timeout)
        else:
            rc = plan.lib.mgb_erfbin_eval_host(*h[7], stream)
        if rc:
            from ..ops import _lib
            _lib.check(rc, "mgb_erfbin_eval_host")
        from ..ops import erfbin as _eb
        _eb._LAUNCHES += 1
        return float(h[6][0]), h[6][1:1 + self.nparams].astype(np.float64)

    def run_adam_host(self, guess, nsteps=100, learning_rate=0.01,
                      init_state=None, return_state=False,
                      return_loss=False, _flush_l2=None):
        """Adam with the optimiser state on the **host**, like the reference
        (rank 0 owns it there, ``multigrad/adam.py:52-68``): every step writes
        the parameters to pinned host memory, runs one fused launch that reads
        them over PCIe and writes loss + gradient back into pinned host
        memory, and applies the update on the CPU -- all inside ONE native
        call (``mgb_erfbin_adam_host``), no Python per step.  Every rank runs
        the same loop on bit-identical gradients.  Use :meth:`run_adam` when
        nothing needs to see the per-step values on the host: it keeps the
        whole fit on the device.  Unbounded parameters only.

        Returns the trajectory ``(nsteps+1, ndim)`` as a CPU tensor (plus the
        loss history / final :class:`AdamState` when asked for)."""
        if not (self.fused and self._builtin_loss):
            return super().run_adam(
                guess, nsteps, None, learning_rate, init_state=init_state,
                return_state=return_state, return_loss=return_loss)
        import ctypes as C
        plan = self.plan()
        P = self.nparams
        if self._host is None:
            self.loss_and_grad_host(_np(guess).ravel())     # pinned buffers
        h = self._host
        h[5][:] = _np(guess).ravel()
        step0 = 0
        m = np.zeros(P, np.float32)
        v = np.zeros(P, np.float32)
        if init_state is not None:
            step0 = int(init_state.step)
            m[:] = _np(init_state.m)
            v[:] = _np(init_state.v)
        traj = np.zeros((nsteps + 1, P), np.float32)
        losses = np.zeros(max(nsteps, 1), np.float32)
        timeout = 60000
        if plan.pt is not None:
            timeout = max(timeout, 2 * int(plan.pt.spin_timeout_ms))
        if torch.cuda.current_device() != self.device.index:
            torch.cuda.set_device(self.device)
        scratch = h[4]
        # benchmarking aid: a CUDA uint8 tensor larger than L2 that is
        # overwritten in front of every step; self.host_timed_seconds then
        # excludes the wait for that fill (see include/multigrad_b200.h)
        timed = C.c_double(0.0)
        fl_ptr = C.c_void_p(_flush_l2.data_ptr()) if _flush_l2 is not None \
            else None
        fl_len = int(_flush_l2.numel()) if _flush_l2 is not None else 0
        rc = plan.lib.mgb_erfbin_adam_host(
            C.byref(plan.model), plan._pt_ref, C.c_void_p(h[0].data_ptr()),
            C.c_void_p(h[1].data_ptr()), m.ctypes.data_as(C.c_void_p),
            v.ctypes.data_as(C.c_void_p), step0, int(nsteps),
            float(learning_rate), _adam.ADAM_B1, _adam.ADAM_B2,
            _adam.ADAM_EPS, traj.ctypes.data_as(C.c_void_p),
            losses.ctypes.data_as(C.c_void_p),
            C.c_void_p(scratch.data_ptr()), scratch.numel(),
            C.c_void_p(torch.cuda.current_stream().cuda_stream), timeout,
            fl_ptr, fl_len, C.byref(timed))
        self.host_timed_seconds = float(timed.value)
        if rc:
            from ..ops import _lib
            _lib.check(rc, "mgb_erfbin_adam_host")
        from ..ops import erfbin as _eb
        _eb._LAUNCHES += int(nsteps)
        self._check_peers()
        out = [torch.from_numpy(traj)]
        if return_loss:
            out.append(torch.from_numpy(losses[:nsteps]))
        if return_state:
            out.append(_adam.AdamState(step0 + nsteps, torch.from_numpy(m),
                                       torch.from_numpy(v)))
        return out[0] if len(out) == 1 else tuple(out)

    # ------------------------------------------------- device-resident driver
    def device_loss_grad_launcher(self, params: torch.Tensor,
                                  out: torch.Tensor):
        """Zero-argument callable that enqueues ONE fused loss+gradient launch
        reading the parameters from the CUDA tensor ``params`` (``[P]``, read
        at execution time, not at call time) and writing ``(loss, grad)`` into
        the CUDA tensor ``out`` (``[1 + P]``).  ``None`` when the model cannot
        run that way (CPU, or a user-defined loss).  Used by
        ``OnePointGroup``'s device-resident optimiser loop."""
        if not (self.fused and self._builtin_loss):
            return None
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        from ..ops._lib import MgbStepArgs
        from ..ops.erfbin import MODE_LOSS_GRAD
        plan = self.plan()
        a = MgbStepArgs()
        a.mode, a.batch = MODE_LOSS_GRAD, 1
        a.reduce = 1 if plan.world > 1 else 0
        a.params = params.data_ptr()
        a.out_loss = out.data_ptr()
        a.out_grad = out.data_ptr() + 4

        def launch(_keep=(params, out)):
            plan.launch(a)
        return launch

    # -------------------------------------------------------------- optimisers
    def run_adam(self, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, randkey=None, const_randkey=False,
                 comm=None, reference_compat=True, init_state=None,
                 return_state=False, return_loss=False, use_graph=None):
        """Adam with one fused kernel launch per step (``csrc/erfbin.cu``
        mode 3): sumstats + Jacobian + NVLink all-reduce + loss + chain rule +
        bound transforms + update, trajectory and loss history written on the
        device.  Same return convention as :meth:`OnePointModel.run_adam`.
        ``randkey`` / ``const_randkey`` are validated like the reference does
        (``multigrad/multigrad.py:291-297``) but have no effect here: the
        built-in link draws no random numbers (a stochastic link is written
        with :class:`~multigrad_b200.models.LinkedBinnedModel`, whose kernel
        derives a ``(seed, step)`` key on the device)."""
        if const_randkey and randkey is None:
            raise ValueError("const_randkey=True requires a randkey")
        if randkey is not None:
            from ..utils import rng as _rng
            _rng.init_randkey(randkey)          # type check, as the reference
        if not (self.fused and self._builtin_loss):
            return super().run_adam(
                guess, nsteps, param_bounds, learning_rate, randkey,
                const_randkey, comm, reference_compat, init_state,
                return_state, return_loss)
        fit = self.adam_fit(guess, nsteps, param_bounds, learning_rate,
                            reference_compat=reference_compat,
                            init_state=init_state)
        fit.step(nsteps, use_graph=use_graph)
        self._check_peers()
        out = [fit.trajectory]
        if return_loss:
            out.append(fit.losses)
        if return_state:
            out.append(fit.state)
        return out[0] if len(out) == 1 else tuple(out)

    def adam_fit(self, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, reference_compat=True, init_state=None):
        """Device-resident Adam fit that the caller advances: returns an
        :class:`AdamFit` with room for ``nsteps`` steps; ``fit.step(n)``
        enqueues the next ``n`` of them (``n == 1``: one fused launch, ``n >
        1``: one persistent launch) and returns immediately, so the caller can
        interleave its own stream work -- logging, a cache flush for a
        benchmark, an early-stopping test on ``fit.losses`` -- between steps.
        ``run_adam(guess, nsteps)`` is ``fit.step(nsteps)`` + the results.
        CUDA + built-in loss only (``None`` otherwise)."""
        if not (self.fused and self._builtin_loss):
            return None
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        from ..ops.erfbin import FitState, MODE_ADAM
        plan = self.plan()
        p0 = self._params(guess)
        bounds = None
        u0 = p0
        if param_bounds is not None:
            bounds = _adam.normalize_bounds(param_bounds, p0.numel())
            u0 = _adam.apply_transforms(p0.cpu(), bounds).to(self.device)
        step0, m, v = 0, None, None
        if init_state is not None:
            step0, m, v = int(init_state.step), init_state.m, init_state.v
        st = FitState(p0, u0, nsteps, self.device, step0=step0, m=m, v=v)
        st.traj[0].copy_(p0)
        args = plan.make_optimizer_args(
            MODE_ADAM, st, learning_rate, bounds,
            jac_at_params=reference_compat)
        return AdamFit(self, plan, st, args, step0, int(nsteps))

    def run_simple_grad_descent(self, guess, nsteps=100, learning_rate=0.01,
                                use_graph=None):
        """Fixed-step gradient descent, one fused launch per step (mode 4).
        Records pre-update params like the reference (quirk Q6)."""
        if not (self.fused and self._builtin_loss):
            return super().run_simple_grad_descent(guess, nsteps,
                                                   learning_rate)
        from ..ops.erfbin import FitState, MODE_GD
        plan = self.plan()
        p0 = self._params(guess)
        st = FitState(p0, p0, nsteps, self.device, rows=max(nsteps, 1))
        args = plan.make_optimizer_args(MODE_GD, st, learning_rate)
        self._run_steps(plan, args, nsteps, use_graph)
        self._check_peers()
        return GradDescentResult(loss=st.loss_hist[:nsteps],
                                 params=st.traj[:nsteps],
                                 aux=[None] * nsteps)

    def _run_steps(self, plan, args, nsteps: int, use_graph) -> None:
        """Run ``nsteps`` optimiser steps.  Default: *persistent* launches --
        up to ``MULTIGRAD_B200_PERSISTENT_CHUNK`` (4096) steps inside ONE
        kernel, whose CTAs stay resident and hand the parameters on through
        device memory (the step index and the collective epoch live on the
        device, so consecutive launches continue each other).  With
        ``MULTIGRAD_B200_PERSISTENT=0`` or ``use_graph`` one launch per step,
        optionally as replays of a captured CUDA graph.  The trajectories of
        all three forms are bit-identical."""
        persistent = os.environ.get("MULTIGRAD_B200_PERSISTENT", "1") != "0"
        if use_graph is None:
            use_graph = os.environ.get("MULTIGRAD_B200_GRAPH", "0") == "1"
        if persistent and not use_graph and nsteps > 1:
            chunk = int(os.environ.get("MULTIGRAD_B200_PERSISTENT_CHUNK",
                                       "4096"))
            done = 0
            while done < nsteps:
                n = min(chunk, nsteps - done)
                args.nsteps = n
                plan.launch(args)
                done += n
            args.nsteps = 1
            return
        args.nsteps = 1
        chunk = int(os.environ.get("MULTIGRAD_B200_GRAPH_CHUNK", "50"))
        done = 0
        if use_graph and nsteps >= 2 * chunk:
            # warm-up launch outside capture, then capture `chunk` launches
            plan.launch(args)
            done = 1
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(chunk):
                    plan.launch(args)
            done += chunk      # capture does not execute: replay once now
            g.replay()
            while nsteps - done >= chunk:
                g.replay()
                done += chunk
            self._graph = g    # keep alive until the stream drains
        for _ in range(nsteps - done):
            plan.launch(args)

    def run_bfgs(self, guess, maxsteps=100, param_bounds=None, randkey=None,
                 comm=None):
        """L-BFGS-B (scipy, float64, on the host) with one fused
        loss+gradient launch per line-search evaluation, through an evaluation
        session (a grid always pre-launched and waiting on the GPU).  With
        several GPUs every rank runs the same deterministic optimiser on the
        bit-identical reduced values (``models/bfgs_device.py``): no
        parameter hand-off and no result broadcast."""
        if not (self.fused and self._builtin_loss):
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
        fn, end_session = self._host_session(x0)
        try:
            return scipy.optimize.minimize(
                fn, x0=x0, method="L-BFGS-B", jac=True,
                options=dict(maxiter=maxsteps), callback=callback,
                bounds=param_bounds)
        finally:
            end_session()
            if hasattr(pbar, "close"):
                pbar.close()

    def _host_session(self, warm_params):
        """``(evaluate, end)``: host-driven loss+gradient evaluations with
        one grid always pre-launched and waiting on the GPU for the next
        parameters (``mgb_erfbin_host_session``), so that a driver which
        returns to Python between evaluations -- scipy's L-BFGS-B -- does not
        pay the launch latency each time.  ``end()`` must be called (it
        cancels the waiting grid).  With ``MULTIGRAD_B200_HOST_PIPELINE=0``
        this is plain ``loss_and_grad_host``."""
        if os.environ.get("MULTIGRAD_B200_HOST_PIPELINE", "1") == "0":
            return self.loss_and_grad_host, (lambda: None)
        import ctypes as C
        plan = self.plan()
        P = self.nparams
        if self._host is None:
            self.loss_and_grad_host(warm_params)        # creates the buffers
        h = self._host
        seq = C.c_uint32(0)
        timeout = 60000
        if plan.pt is not None:
            timeout = max(timeout, 2 * int(plan.pt.spin_timeout_ms))
        if torch.cuda.current_device() != self.device.index:
            torch.cuda.set_device(self.device)
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        base = (C.byref(plan.model), plan._pt_ref,
                C.c_void_p(h[0].data_ptr()), C.c_void_p(h[1].data_ptr()),
                C.c_void_p(h[4].data_ptr()), h[4].numel(), stream, timeout)
        session = plan.lib.mgb_erfbin_host_session
        from ..ops import _lib, erfbin as _eb

        def call(op):
            rc = session(*base, op, C.byref(seq))
            if rc:
                _lib.check(rc, "mgb_erfbin_host_session")

        call(0)
        n_in, n_out = h[5], h[6]

        def evaluate(p):
            n_in[:] = p
            call(1)
            _eb._LAUNCHES += 1
            return float(n_out[0]), n_out[1:1 + P].astype(np.float64)

        def end():
            call(2)

        return evaluate, end

    def run_lhs_param_scan(self, xmins, xmaxs, n_dim, num_evaluations,
                           seed=None, randkey=None):
        """Latin-hypercube scan with **all draws evaluated in one launch**
        (batched forward kernel) instead of a host loop."""
        if not self.fused:
            return super().run_lhs_param_scan(xmins, xmaxs, n_dim,
                                              num_evaluations, seed, randkey)
        params = util.latin_hypercube_sampler(xmins, xmaxs, n_dim,
                                              num_evaluations, seed=seed)
        from ..ops._lib import MGB_MAX_BATCH
        p = torch.as_tensor(params, dtype=torch.float32, device=self.device)
        chunks = []
        for lo in range(0, p.shape[0], MGB_MAX_BATCH):
            chunks.append(self.calc_sumstats_from_params(
                p[lo:lo + MGB_MAX_BATCH].contiguous()))
        sumstats = torch.cat(chunks) if chunks else torch.zeros(
            0, self.nbins, device=self.device)
        with torch.no_grad():
            losses = torch.stack([self.calc_loss_from_sumstats(s)
                                  for s in sumstats]) \
                if len(sumstats) else torch.zeros(0)
        return params, sumstats.cpu().numpy(), losses.cpu().numpy()


@dataclass(eq=False)
class SMFModel(ErfBinnedModel):
    """Stellar-mass-function model of the reference examples.

    ``params = (log_shmrat, sigma_logsm)`` (``sigma_link="identity"``, the
    test-suite model) or ``(log_f, log_sigma)`` with ``sigma_link="log10"``
    and ``loss_eps=1e-10`` (the tutorial model)."""
    pass
