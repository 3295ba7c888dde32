"""The reference's stellar-mass-function example, written against the generic
user API (plain differentiable PyTorch in the two hooks).

This is the analogue of ``tests/smf_example/smf_grad_descent.py:17-82`` (test
model, ``sigma = p1``) and of ``docs/source/notebooks/
smf_gradient_descent.py:10-91`` (tutorial model, ``sigma = 10**p1``): it shows
what a user of the reference has to change -- ``jnp`` -> ``torch`` -- and it
is the oracle for the native :class:`~multigrad_b200.models.SMFModel`.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import NamedTuple

import numpy as np
import torch

from ..core import OnePointModel
from ..parallel import comm as _comm

__all__ = ["ParamTuple", "load_halo_masses", "load_halo_masses_tutorial",
           "calc_smf_cdf", "calc_smf_bin", "MySMFModel", "TutorialSMFModel",
           "TARGET_SUMSTATS_TEST", "make_smf_data", "synthetic_log_halo_masses",
           "synthetic_sumstats_and_jac", "synthetic_adam_oracle"]


class ParamTuple(NamedTuple):
    log_shmrat: float = -2.0
    sigma_logsm: float = 0.2


# SMF at truth params=(-2.0, 0.2) for 10 000 halos, V = 10 N
# (reference tests/test_mpi.py:44-47, tests/smf_example/smf_grad_descent.py:97-100)
TARGET_SUMSTATS_TEST = (
    2.30178721e-02, 1.69728529e-02, 1.16054425e-02, 7.10532581e-03,
    3.77187086e-03, 1.69136131e-03, 6.28149020e-04, 1.90466686e-04,
    4.66692982e-05, 9.17260695e-06)


def _shard(arr, comm):
    comm = _comm.get_world() if comm is None else comm
    if comm is None or comm.size == 1:
        return arr
    return torch.tensor_split(arr, comm.size)[comm.rank]


def load_halo_masses(num_halos=10_000, slope=-2, mmin=10.0 ** 10, qmax=0.95,
                     comm=None):
    """Fake power-law halo mass function, truncated so that the SMF has a
    knee; this rank's near-even shard (reference
    ``smf_grad_descent.py:23-28``).  float32 like the reference."""
    q = torch.linspace(0, qmax, num_halos, dtype=torch.float32)
    mhalo = mmin * (1 - q) ** (1 / (slope + 1))
    return _shard(mhalo, comm)


def load_halo_masses_tutorial(num_halos=10_000, comm=None):
    """Tutorial generator: ``1e10/(1-q)``, ``q`` in ``[0, 0.9]`` (reference
    ``docs/source/notebooks/smf_gradient_descent.py:10-16``)."""
    q = torch.linspace(0, 0.9, num_halos, dtype=torch.float32)
    return _shard(1e10 / (1 - q), comm)


def synthetic_log_halo_masses(num_halos, rank=0, world=1, device="cpu",
                              qmax=0.95, chunk=1 << 24):
    """``log10`` of :func:`load_halo_masses` for this rank's shard, generated
    *on the device* in chunks (1e8 halos never exist on the host).  For the
    power law ``m = 1e10 (1-q)^-1``: ``log10 m = 10 - log10(1-q)``."""
    bounds = np.linspace(0, num_halos, world + 1).astype(np.int64)
    # near-even split identical to np.array_split / torch.tensor_split
    base, extra = divmod(num_halos, world)
    lo = rank * base + min(rank, extra)
    hi = lo + base + (1 if rank < extra else 0)
    del bounds
    out = torch.empty(hi - lo, dtype=torch.float32, device=device)
    denom = max(num_halos - 1, 1)
    for s in range(lo, hi, chunk):
        e = min(s + chunk, hi)
        idx = torch.arange(s, e, device=device, dtype=torch.float64)
        q = idx * (qmax / denom)
        out[s - lo:e - lo] = (10.0 - torch.log10(1.0 - q)).to(torch.float32)
    return out


def calc_smf_cdf(logsm, mean_logsm, sigma_logsm):
    return 0.5 * (1 + torch.special.erf(
        (logsm - mean_logsm) / (math.sqrt(2) * sigma_logsm)))


def calc_smf_bin(params, logsm_low, logsm_high, volume, log_halo_masses):
    log_shmrat, sigma_logsm = params[0], params[1]
    mean_logsm = log_halo_masses + log_shmrat
    cdf_high = calc_smf_cdf(logsm_high, mean_logsm, sigma_logsm)
    cdf_low = calc_smf_cdf(logsm_low, mean_logsm, sigma_logsm)
    return torch.sum(cdf_high - cdf_low) / volume / (logsm_high - logsm_low)


@dataclass(eq=False)
class MySMFModel(OnePointModel):
    """User-style model: two differentiable torch hooks, nothing else."""
    aux_data: dict = None
    loss_func_has_aux: bool = False

    def calc_partial_sumstats_from_params(self, params):
        bin_edges = torch.as_tensor(self.aux_data["smf_bin_edges"])
        log_halo_masses = torch.as_tensor(self.aux_data["log_halo_masses"])
        volume = self.aux_data["volume"]
        smf = []
        logsm_low = bin_edges[0]
        for logsm_high in bin_edges[1:]:
            smf.append(calc_smf_bin(params, logsm_low, logsm_high, volume,
                                    log_halo_masses))
            logsm_low = logsm_high
        return torch.stack(smf)

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None):
        target = torch.log10(torch.as_tensor(
            self.aux_data["target_sumstats"]).to(sumstats))
        return torch.mean((torch.log10(sumstats) - target) ** 2)


@dataclass(eq=False)
class TutorialSMFModel(MySMFModel):
    """Tutorial variant: ``sigma = 10**log_sigma`` and ``+1e-10`` inside the
    logs (reference ``docs/.../smf_gradient_descent.py:21-35,86-91``)."""

    def calc_partial_sumstats_from_params(self, params):
        p = torch.stack([params[0], 10.0 ** params[1]])
        return super().calc_partial_sumstats_from_params(p)

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None):
        target = torch.log10(torch.as_tensor(
            self.aux_data["target_sumstats"]).to(sumstats) + 1e-10)
        return torch.mean((torch.log10(sumstats + 1e-10) - target) ** 2)


def make_smf_data(num_halos=10_000, comm=None, device="cpu"):
    """``aux_data`` dict of the reference test (``tests/test_mpi.py:39-48``)."""
    return dict(
        log_halo_masses=torch.log10(load_halo_masses(num_halos, comm=comm)
                                    ).to(device),
        smf_bin_edges=torch.linspace(9, 10, 11).to(device),
        volume=10.0 * num_halos,
        target_sumstats=torch.tensor(TARGET_SUMSTATS_TEST).to(device),
    )


# ---------------------------------------------------------------------------
# float64 oracle for the synthetic catalogue without touching its objects.
# ``synthetic_log_halo_masses`` places object i at x(q_i), q_i = i*qmax/(N-1):
# a sum over the objects of a smooth function is the trapezoid sum of
# f(x(q)), which Gauss-Legendre quadrature plus the Euler-Maclaurin end
# corrections reproduces to ~1e-12 relative -- the check ``bench.py`` runs at
# every GPU count against the fused kernels' trajectories.
# ---------------------------------------------------------------------------
def synthetic_sumstats_and_jac(num_halos, bin_edges, volume, params,
                               qmax=0.95, panels=4000, order=8):
    """``(s[K], ds/dmu[K], ds/dsigma[K])`` in float64 for the model
    ``mu = x + params[0]``, ``sigma = params[1]`` on the synthetic catalogue
    of ``num_halos`` objects (sum over ALL objects, i.e. all ranks)."""
    from scipy.special import erf
    edges = np.asarray(bin_edges, dtype=np.float64)
    p_mu, sig = float(params[0]), float(params[1])
    N = int(num_halos)
    h = qmax / max(N - 1, 1)
    xs, ws = np.polynomial.legendre.leggauss(order)
    bounds = np.linspace(0.0, qmax, panels + 1)
    half = 0.5 * np.diff(bounds)
    q = (0.5 * (bounds[1:] + bounds[:-1]))[:, None] + half[:, None] * xs
    wq = (half[:, None] * ws[None, :]).ravel()
    q = q.ravel()

    def terms(qv):
        x = 10.0 - np.log10(1.0 - qv)
        z = (edges[:, None] - (x + p_mu)[None, :]) / (math.sqrt(2.0) * sig)
        cdf = 0.5 * (1.0 + erf(z))
        g = np.exp(-z * z) / math.sqrt(math.pi)
        return cdf, -g / (math.sqrt(2.0) * sig), -g * z / sig

    def dterms(qv, eps=1e-6):
        a = terms(np.asarray([min(qv + eps, qmax)]))
        b = terms(np.asarray([max(qv - eps, 0.0)]))
        span = min(qv + eps, qmax) - max(qv - eps, 0.0)
        return [(u - v)[:, 0] / span for u, v in zip(a, b)]

    mid = terms(q)
    end0 = terms(np.asarray([0.0]))
    end1 = terms(np.asarray([qmax]))
    d0, d1 = dterms(0.0), dterms(qmax)
    out = []
    for k in range(3):
        tot = (mid[k] * wq[None, :]).sum(axis=1) / h
        tot += 0.5 * (end0[k][:, 0] + end1[k][:, 0])
        tot += (h / 12.0) * (d1[k] - d0[k])
        out.append(np.diff(tot) / (float(volume) * np.diff(edges)))
    return tuple(out)


def synthetic_adam_oracle(num_halos, bin_edges, volume, target, guess, nsteps,
                          learning_rate=0.01, b1=0.9, b2=0.999, eps=1e-8):
    """float64 Adam trajectory ``(nsteps+1, 2)`` and loss history of the
    log-MSE SMF fit on the synthetic catalogue (unbounded parameters)."""
    p = np.asarray(guess, dtype=np.float64).copy()
    tgt = np.asarray(target, dtype=np.float64)
    m = np.zeros(2)
    v = np.zeros(2)
    traj, losses = [p.copy()], []
    K = tgt.size
    for i in range(nsteps):
        s, dmu, dsg = synthetic_sumstats_and_jac(num_halos, bin_edges, volume,
                                                 p)
        r = np.log10(s) - np.log10(tgt)
        losses.append(float(np.mean(r * r)))
        dl = 2.0 * r / (K * s * math.log(10.0))
        g = np.array([dl @ dmu, dl @ dsg])
        m = (1 - b1) * g + b1 * m
        v = (1 - b2) * g * g + b2 * v
        mh, vh = m / (1 - b1 ** (i + 1)), v / (1 - b2 ** (i + 1))
        p = p - learning_rate * mh / (np.sqrt(vh) + eps)
        traj.append(p.copy())
    return np.asarray(traj), np.asarray(losses)
