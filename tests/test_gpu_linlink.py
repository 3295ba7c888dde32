"""Tensor-core linear-link kernel (csrc/linlink.cu) against a float64
PyTorch autograd oracle of the same model (same bf16-rounded features)."""
import ctypes as C

import numpy as np
import pytest
import torch

import multigrad_b200 as mg
from multigrad_b200.models import LinearLinkBinnedModel

pytestmark = pytest.mark.gpu
SELF = mg.Comm(None)


def dev():
    return torch.device("cuda", 0)


def make(n, F, k=10, seed=0, link="identity", **kw):
    g = torch.Generator().manual_seed(seed)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    feat = torch.cat([torch.ones(n, 1),
                      0.3 * torch.randn(n, F - 1, generator=g)], 1)
    edges = np.linspace(9.0, 10.0, k + 1)
    gpu = LinearLinkBinnedModel(x=x.to(dev()), features=feat.to(dev()),
                                bin_edges=edges, volume=10.0 * max(n, 1),
                                sigma_link=link, comm=SELF, device=dev(), **kw)
    cpu = LinearLinkBinnedModel(x=x.double(), features=feat, bin_edges=edges,
                                volume=10.0 * max(n, 1), sigma_link=link,
                                comm=SELF, device="cpu", **kw)
    theta = torch.cat([torch.tensor([-2.0]),
                       0.05 * torch.randn(F - 1, generator=g),
                       torch.tensor([0.2 if link == "identity" else -0.6])])
    return gpu, cpu, theta


def oracle(cpu, p):
    p64 = p.double().clone().requires_grad_(True)
    s = cpu.calc_partial_sumstats_from_params(p64)
    jac = torch.autograd.functional.jacobian(
        lambda q: cpu.calc_partial_sumstats_from_params(q), p.double())
    return s.detach(), jac


def test_selftest_building_blocks():
    from multigrad_b200.ops import _lib
    lib = _lib.load()
    for P in (64, 128):
        g = torch.Generator().manual_seed(P)
        W = torch.randn(128, P, generator=g).to(torch.bfloat16).to(dev())
        theta = torch.randn(P, generator=g).to(dev())
        G = torch.rand(32, 128, generator=g).to(dev())
        mu = torch.zeros(128, device=dev())
        J = torch.zeros(32, P, device=dev())
        status = torch.zeros(1, dtype=torch.int32, device=dev())
        rc = lib.mgb_linlink_selftest(W.data_ptr(), theta.data_ptr(),
                                      G.data_ptr(), mu.data_ptr(),
                                      J.data_ptr(), status.data_ptr(), P, None)
        torch.cuda.synchronize()
        assert rc == 0 and int(status) == 0
        mu_ref = W.double() @ theta.double()
        J_ref = G.to(torch.bfloat16).double() @ W.double()
        assert (mu.double() - mu_ref).abs().max() < 1e-5 * mu_ref.abs().max()
        assert (J.double() - J_ref).abs().max() < 1e-5 * J_ref.abs().max()


@pytest.mark.parametrize("n,F,k", [(128, 3, 10), (1000, 7, 10), (20_000, 33, 11),
                                   (50_001, 64, 10), (30_000, 100, 5),
                                   (0, 4, 3), (5, 2, 1)])
def test_sumstats_and_jacobian_vs_fp64(n, F, k):
    gpu, cpu, p = make(n, F, k)
    s, jac = gpu.calc_sumstats_and_jac_from_params(p)
    gpu.plan().check_pipeline()
    so, jo = oracle(cpu, p)
    assert s.shape == (k,) and jac.shape == (k, F + 1)
    assert torch.allclose(s.double().cpu(), so, rtol=5e-6, atol=1e-12)
    scale = float(jo.abs().max()) if n else 1.0
    assert torch.allclose(jac.double().cpu(), jo, rtol=2e-4,
                          atol=2e-5 * scale), (jac[:, :3], jo[:, :3])


def test_loss_grad_and_log10_link():
    gpu, cpu, p = make(40_000, 20, link="log10")
    truth = p.clone()
    tgt = gpu.calc_sumstats_from_params(truth)
    gpu.set_target(tgt)
    cpu.set_target(tgt.cpu())
    q = truth + 0.03
    loss, grad = gpu.calc_loss_and_grad_from_params(q)
    q64 = q.double().clone().requires_grad_(True)
    l64 = cpu.calc_loss_from_sumstats(cpu.calc_partial_sumstats_from_params(q64))
    (g64,) = torch.autograd.grad(l64, q64)
    assert abs(float(loss) - float(l64)) < 2e-4 * abs(float(l64))
    assert torch.allclose(grad.double().cpu(), g64, rtol=1e-3,
                          atol=1e-4 * float(g64.abs().max()))
    assert torch.allclose(gpu.calc_dloss_dparams(q), grad)


@pytest.mark.parametrize("bounds", [False, True])
def test_fused_adam_matches_generic(bounds):
    F = 6
    gpu, cpu, p = make(30_000, F)
    tgt = gpu.calc_sumstats_from_params(p)
    gpu.set_target(tgt)
    cpu32 = LinearLinkBinnedModel(
        x=gpu.x.cpu(), features=gpu.features.cpu(), bin_edges=gpu._edges64,
        volume=gpu.volume, target_sumstats=tgt.cpu(), comm=SELF, device="cpu")
    guess = p + 0.05
    pb = [(-5.0, 0.0)] + [None] * (F - 1) + [(0.01, None)] if bounds else None
    traj, losses = gpu.run_adam(guess, nsteps=20, learning_rate=0.01,
                                param_bounds=pb, return_loss=True)
    rt, rl = cpu32.run_adam(guess, nsteps=20, learning_rate=0.01,
                            param_bounds=pb, return_loss=True)
    assert traj.shape == (21, F + 1)
    assert torch.allclose(traj.cpu(), rt, atol=5e-4), (traj[-1], rt[-1])
    assert torch.allclose(losses.cpu(), rl, rtol=5e-3, atol=1e-7)
    gd = gpu.run_simple_grad_descent(guess, nsteps=5, learning_rate=1e-3)
    rgd = cpu32.run_simple_grad_descent(guess, nsteps=5, learning_rate=1e-3)
    assert torch.allclose(gd.params.cpu(), rgd.params, atol=2e-4)


def test_constant_feature_reproduces_smf_model():
    """One constant feature == the scalar-shift SMF model."""
    from multigrad_b200.models import SMFModel, smf
    data = smf.make_smf_data(10_000, comm=SELF, device=dev())
    n = data["log_halo_masses"].numel()
    lin = LinearLinkBinnedModel(
        x=data["log_halo_masses"], features=torch.ones(n, 1, device=dev()),
        bin_edges=np.linspace(9, 10, 11), volume=data["volume"],
        target_sumstats=data["target_sumstats"], comm=SELF, device=dev())
    ref = SMFModel(aux_data=data, comm=SELF, device=dev())
    p = torch.tensor([-1.9, 0.3])
    l1, g1 = lin.calc_loss_and_grad_from_params(p)
    l2, g2 = ref.calc_loss_and_grad_from_params(p)
    assert torch.allclose(l1, l2, rtol=1e-5)
    assert torch.allclose(g1, g2, rtol=1e-4, atol=1e-6)
    assert torch.allclose(data["target_sumstats"],
                          lin.calc_sumstats_from_params([-2.0, 0.2]))


def make_w(n, F, k=10, seed=3, **kw):
    g = torch.Generator().manual_seed(seed + 77)
    w = (0.2 + 1.6 * torch.rand(n, generator=g)).float()
    gpu, cpu, p = make(n, F, k, seed=seed, **kw)
    kwg = dict(x=gpu.x, features=gpu.features, bin_edges=gpu._edges64,
               volume=gpu.volume, sigma_link=gpu.sigma_link, comm=SELF)
    gw = LinearLinkBinnedModel(weights=w.to(dev()), device=dev(), **kwg)
    kwc = dict(kwg, x=cpu.x, features=cpu.features)
    cw = LinearLinkBinnedModel(weights=w.double(), device="cpu", **kwc)
    return gw, cw, p


@pytest.mark.parametrize("n,F,k", [(128, 3, 10), (10_000, 7, 10),
                                   (50_001, 64, 11), (30_003, 100, 5),
                                   (5, 2, 1)])
def test_weighted_objects_vs_fp64(n, F, k):
    """Per-object weights (whole tiles through the bulk copy, the ragged last
    tile through plain loads)."""
    gpu, cpu, p = make_w(n, F, k)
    s, jac = gpu.calc_sumstats_and_jac_from_params(p)
    gpu._check()
    so, jo = oracle(cpu, p)
    assert torch.allclose(s.double().cpu(), so, rtol=5e-6, atol=1e-12)
    scale = float(jo.abs().max())
    assert torch.allclose(jac.double().cpu(), jo, rtol=2e-4,
                          atol=2e-5 * scale)
    # unit weights are the unweighted model
    g1, _, _ = make(n, F, k, seed=3)
    gu = LinearLinkBinnedModel(
        x=g1.x, features=g1.features, bin_edges=g1._edges64, volume=g1.volume,
        weights=torch.ones(n, device=dev()), comm=SELF, device=dev())
    s1, j1 = g1.calc_sumstats_and_jac_from_params(p)
    su, ju = gu.calc_sumstats_and_jac_from_params(p)
    assert torch.allclose(s1, su, rtol=2e-6, atol=1e-12)
    assert torch.allclose(j1, ju, rtol=1e-5, atol=1e-6 * float(j1.abs().max()))


@pytest.mark.parametrize("k", [12, 25, 32])
def test_more_than_eleven_bins_in_slices(k):
    gpu, cpu, p = make(40_000, 9, k)
    assert len(gpu._slices) == -(-k // 11)
    s, jac = gpu.calc_sumstats_and_jac_from_params(p)
    gpu._check()
    so, jo = oracle(cpu, p)
    assert s.shape == (k,) and jac.shape == (k, 10)
    assert torch.allclose(s.double().cpu(), so, rtol=5e-6, atol=1e-12)
    assert torch.allclose(jac.double().cpu(), jo, rtol=2e-4,
                          atol=2e-5 * float(jo.abs().max()))
    tgt = gpu.calc_sumstats_from_params(p)
    gpu.set_target(tgt)
    cpu.set_target(tgt.cpu())
    q = p + 0.02
    loss, grad = gpu.calc_loss_and_grad_from_params(q)
    q64 = q.double().clone().requires_grad_(True)
    l64 = cpu.calc_loss_from_sumstats(cpu.calc_partial_sumstats_from_params(q64))
    (g64,) = torch.autograd.grad(l64, q64)
    assert abs(float(loss) - float(l64)) < 5e-4 * abs(float(l64))
    assert torch.allclose(grad.double().cpu(), g64, rtol=2e-3,
                          atol=2e-4 * float(g64.abs().max()))
    traj, losses = gpu.run_adam(q, nsteps=15, learning_rate=0.005,
                                return_loss=True)
    assert float(losses[-1]) < 0.5 * float(losses[0])
    l_h, g_h = gpu.loss_and_grad_host(q.numpy())
    assert abs(l_h - float(loss)) <= 1e-6 * abs(float(loss))


def test_host_evaluation_and_bfgs():
    F = 5
    gpu, cpu, p = make(60_000, F)
    tgt = gpu.calc_sumstats_from_params(p)
    gpu.set_target(tgt)
    q = p + 0.03
    from multigrad_b200.ops import linlink as ll
    n0 = ll.launch_count()
    l_h, g_h = gpu.loss_and_grad_host(q.numpy())
    assert ll.launch_count() == n0 + 1          # one launch, no copies
    loss, grad = gpu.calc_loss_and_grad_from_params(q)
    assert l_h == float(loss)
    assert np.array_equal(g_h.astype(np.float32), grad.cpu().numpy())
    res = gpu.run_bfgs(q, maxsteps=60,
                       param_bounds=[(None, None)] * F + [(0.02, 2.0)])
    assert res.fun < 1e-3 * l_h, res
    # host-state Adam walks the device-state trajectory
    th, lh = gpu.run_adam_host(q, nsteps=12, learning_rate=0.01,
                               return_loss=True)
    td, ld = gpu.run_adam(q, nsteps=12, learning_rate=0.01, return_loss=True)
    assert torch.allclose(th, td.cpu(), atol=2e-5)
    assert torch.allclose(lh, ld.cpu(), rtol=1e-4, atol=1e-9)


def test_joins_a_device_resident_group():
    """A linear-link model and a built-in SMF model sharing (shift, sigma) in
    one OnePointGroup: the device-resident group loop applies."""
    from multigrad_b200.models import SMFModel, smf
    from multigrad_b200.ops import group as grp
    data = smf.make_smf_data(30_000, comm=SELF, device=dev())
    n = data["log_halo_masses"].numel()
    lin = LinearLinkBinnedModel(
        x=data["log_halo_masses"] + 0.1,
        features=torch.ones(n, 1, device=dev()),
        bin_edges=np.linspace(9, 10, 8), volume=data["volume"], comm=SELF,
        device=dev())
    lin.set_target(lin.calc_sumstats_from_params([-2.0, 0.2]))
    ref = SMFModel(aux_data=data, comm=SELF, device=dev())
    group = mg.OnePointGroup(models=(ref, lin), main_comm=SELF)
    p = torch.tensor([-1.9, 0.3], device=dev())
    assert group._device_ok(p)
    lg, gg = group.calc_loss_and_grad_from_params(p.cpu())
    l1, g1 = ref.calc_loss_and_grad_from_params(p.cpu())
    l2, g2 = lin.calc_loss_and_grad_from_params(p.cpu())
    assert torch.allclose(torch.as_tensor(lg).cpu(), (l1 + l2).cpu(),
                          rtol=1e-5)
    assert torch.allclose(torch.as_tensor(gg).cpu(), (g1 + g2).cpu(),
                          rtol=1e-4, atol=1e-7)
    traj = group.run_adam(p.cpu(), nsteps=150, learning_rate=0.01)
    assert torch.allclose(traj[-1].cpu(), torch.tensor([-2.0, 0.2]),
                          atol=0.02), traj[-1]


def test_float64_default_dtype_does_not_change_the_buffers():
    """Buffers handed to the C ABI are float32 whatever torch's default is."""
    from multigrad_b200.models import SMFModel, smf
    gpu, _, p = make(20_000, 5)
    s0, j0 = gpu.calc_sumstats_and_jac_from_params(p)
    data = smf.make_smf_data(10_000, comm=SELF, device=dev())
    ref = SMFModel(aux_data=data, comm=SELF, device=dev(), deterministic=True)
    l0, g0 = ref.calc_loss_and_grad_from_params([-1.9, 0.3])
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.float64)
    try:
        s1, j1 = gpu.calc_sumstats_and_jac_from_params(p)
        l1, g1 = ref.calc_loss_and_grad_from_params([-1.9, 0.3])
    finally:
        torch.set_default_dtype(old)
    assert s1.dtype == torch.float32 and torch.equal(s0, s1)
    assert torch.equal(j0, j1)
    assert torch.equal(l0, l1) and torch.equal(g0, g1)


def test_fp8_gate_decision_is_reproducible():
    """The FP8 gate (tools/fp8_gate.py): MXFP8 features are emulated exactly
    on the bf16 kernel; the committed decision (profiles/linlink_fp8_gate.json)
    must be the one this measurement gives."""
    import json
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    sys.path.insert(0, str(root / "tools"))
    import fp8_gate
    # the emulation is exact: dequantised MXFP8 values survive bf16
    w = 0.3 * torch.randn(64, 64)
    q = fp8_gate.mxfp8_dequant(w)
    assert torch.equal(q, q.to(torch.bfloat16).float())
    assert float((q - w).abs().max()) < 0.07 * float(w.abs().max())
    res = fp8_gate.run(100_000, 64)
    assert res["bf16"]["jacobian_rel_err"] < 2e-4
    accepted = (res["mxfp8"]["jacobian_rel_err"] <= fp8_gate.JAC_TOL and
                res["mxfp8"]["sumstats_rel_err"] <= fp8_gate.SUM_TOL)
    assert res["decision"].startswith("accepted") == accepted
    rec = root / "profiles" / "linlink_fp8_gate.json"
    if rec.exists():
        lines = [json.loads(l) for l in rec.read_text().splitlines() if l]
        assert all(l["decision"].startswith("accepted") == accepted
                   for l in lines), (res, lines)
