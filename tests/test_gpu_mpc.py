"""GPU parity of the MPC-side rows: chance-constraint tightening and FITC weights.

The fixtures tests/golden/tighten_*.npz and fitc_*.npz are outputs of the reference's own methods
(tests/golden/make_golden_mpc.py); the CUDA path is compared with them and with the oracle restatement."""
import os

import numpy as np
import pytest
import torch

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
GP_IDX = [[0], [1, 2, 3], [4, 5, 6]]


def _gps(g):
    return [(torch.tensor(g[f"gp{k}_X"]), torch.tensor(g[f"gp{k}_y"]), torch.tensor(g[f"gp{k}_ell"]),
             float(g[f"gp{k}_sf2"]), float(g["noise"][k])) for k in range(3)]


def _tighten_raw(var, has_noise, z, noise, A, B, K, Bd, T, R, dt, icdf, idx=(6, 7)):
    """Straight through the C ABI with device buffers."""
    from gpmpc_b200 import _lib
    from gpmpc_b200._lib import call, ptr
    dev = torch.device("cuda", 0)
    nx, nu = A.shape[0], B.shape[1]
    hs = _lib.load().gpk_hyp_stride()
    hyp = torch.zeros(3, hs, dtype=torch.float64)
    hyp[:, 65] = torch.as_tensor(noise, dtype=torch.float64)
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64).to(dev)   # noqa: E731
    var_d, z_d, hyp_d, A_d, B_d, K_d, Bd_d = t(var), t(z), hyp.to(dev), t(A), t(B), t(K), t(Bd)
    sc = torch.full((R, 2 * nx, T + 1), float("nan"), dtype=torch.float64, device=dev)
    ic = torch.full((R, 2 * nu, T), float("nan"), dtype=torch.float64, device=dev)
    call("gpk_tighten_constraints", ptr(var_d), var_d.stride(0) if var_d.numel() else 0, has_noise, ptr(z_d),
         z_d.shape[1], idx[0], idx[1], ptr(hyp_d), ptr(A_d), ptr(B_d), ptr(K_d), ptr(Bd_d), nx, nu, T, R, dt, icdf,
         ptr(sc), ptr(ic), torch.cuda.current_stream().cuda_stream)
    return sc.cpu().numpy(), ic.cpu().numpy()


@pytest.mark.parametrize("name", ["tighten_T25", "tighten_T40"])
def test_tightening_kernel_matches_reference_output(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    R, nx, T1 = g["x_prev"].shape
    T = T1 - 1
    z = np.concatenate([np.hstack((g["x_prev"][r][:, :-1].T, g["u_prev"][r].T)) for r in range(R)])
    var_y = np.concatenate([g["var_y"][r] for r in range(R)], axis=1)             # (3, R*T) predictive variances
    sc, ic = _tighten_raw(var_y, 1, z, g["noise"], g["dfdx"], g["dfdu"], g["lqr_gain"], g["Bd"], T, R,
                          float(g["dt"]), float(g["inverse_cdf"]))
    assert o.rel_err(sc, g["state_constraint"]) < 1e-12
    assert o.rel_err(ic, g["input_constraint"]) < 1e-12
    # latent variances + the noise added twice is the same thing
    sc2, ic2 = _tighten_raw(var_y - g["noise"][:, None], 0, z, g["noise"], g["dfdx"], g["dfdu"], g["lqr_gain"], g["Bd"],
                            T, R, float(g["dt"]), float(g["inverse_cdf"]))
    assert o.rel_err(sc2, g["state_constraint"]) < 1e-12 and o.rel_err(ic2, g["input_constraint"]) < 1e-12


@pytest.mark.parametrize("name", ["tighten_T25", "tighten_T40"])
def test_propagate_constraint_limits_end_to_end(name):
    """GP variances from the tcgen05 prediction path feeding the tightening kernel, against the reference output."""
    from gpmpc_b200.mpc import ConstraintTightening
    g = np.load(os.path.join(GOLD, name + ".npz"))
    ct = ConstraintTightening(_gps(g), GP_IDX, g["dfdx"], g["dfdu"], g["lqr_gain"], g["Bd"], float(g["dt"]),
                              float(g["inverse_cdf"]), horizon=g["u_prev"].shape[2])
    sc0, ic0 = ct.propagate_constraint_limits(None, None)                      # no previous rollout: zeros
    assert sc0.shape == g["state_constraint"][0].shape and not sc0.any() and not ic0.any()
    sc, ic = ct.propagate_constraint_limits(g["x_prev"][0], g["u_prev"][0])      # the reference's call shape
    assert isinstance(sc, np.ndarray) and sc.shape == g["state_constraint"][0].shape
    assert o.rel_err(sc, g["state_constraint"][0]) < 1e-7 and o.rel_err(ic, g["input_constraint"][0]) < 1e-7
    scb, icb = ct.propagate_constraint_limits(torch.tensor(g["x_prev"]), torch.tensor(g["u_prev"]))   # batched
    assert scb.is_cuda and o.rel_err(scb.cpu(), g["state_constraint"]) < 1e-7
    assert o.rel_err(icb.cpu(), g["input_constraint"]) < 1e-7
    assert np.array_equal(scb[0].cpu().numpy(), sc)


@pytest.mark.parametrize("nx,nu,T,R", [(12, 4, 25, 257), (8, 2, 7, 5), (16, 8, 12, 33), (5, 1, 3, 2), (12, 4, 1, 1)])
def test_tightening_kernel_matches_oracle_on_random_systems(nx, nu, T, R):
    rng = np.random.default_rng(nx * 100 + T)
    A = np.eye(nx) + 0.05 * rng.standard_normal((nx, nx))
    B = 0.1 * rng.standard_normal((nx, nu))
    K = -0.3 * rng.standard_normal((nu, nx))
    Bd = np.zeros((nx, 5))
    Bd[rng.permutation(nx)[:5], np.arange(5)] = 1.0
    Bd += 0.01 * rng.standard_normal((nx, 5))                                 # not just a selection matrix
    z = rng.uniform(-1.0, 1.0, size=(R * T, nx + nu))
    idx = (nx - 2, nx - 1)
    var = rng.uniform(1e-4, 1e-2, size=(3, R * T))
    noise = np.array([1e-3, 2e-4, 5e-4])
    sc, ic = _tighten_raw(var, 1, z, noise, A, B, K, Bd, T, R, 0.02, 2.3, idx)
    for r in sorted({0, R // 2, R - 1}):
        s_ref, i_ref = o.propagate_constraint_limits(var[:, r * T:(r + 1) * T], z[r * T:(r + 1) * T], noise, A, B, K, Bd,
                                                     0.02, 2.3, idx_phi=idx[0], idx_theta=idx[1])
        assert o.rel_err(sc[r], s_ref) < 1e-11 and o.rel_err(ic[r], i_ref) < 1e-11
    assert not np.isnan(sc).any() and not np.isnan(ic).any()


def test_tightening_edge_cases():
    from gpmpc_b200._lib import GpkError
    A, B, K, Bd = np.eye(12), np.zeros((12, 4)), np.zeros((4, 12)), np.eye(12)[:, :5]
    sc, ic = _tighten_raw(np.zeros((3, 0)), 1, np.zeros((0, 16)), np.zeros(3), A, B, K, Bd, 0, 3, 0.02, 2.0)
    assert sc.shape == (3, 24, 1) and not sc.any() and ic.size == 0             # T = 0: only the initial zero covariance
    sc, ic = _tighten_raw(np.zeros((3, 0)), 1, np.zeros((0, 16)), np.zeros(3), A, B, K, Bd, 4, 0, 0.02, 2.0)
    assert sc.shape == (0, 24, 5)                                                # no rollouts: nothing launched
    # with A = I, K = 0 the covariance is the running sum of the stage disturbances
    T = 6
    var = np.full((3, T), 4.0)
    z = np.zeros((T, 16))
    sc, ic = _tighten_raw(var, 1, z, np.zeros(3), A, B, K, Bd, T, 1, 0.5, 1.0)
    # phi = theta = 0: factors (0, 0, 1, 1, 1); dt^2 var = 1 per stage on columns 2, 3, 4 of Bd
    for i in range(T + 1):
        assert np.allclose(sc[0, :12, i], -np.sqrt(i) * np.array([0, 0, 1, 1, 1] + [0] * 7), atol=1e-15)
    assert not ic.any()
    with pytest.raises(GpkError):
        _tighten_raw(var, 1, z, np.zeros(3), np.eye(17), np.zeros((17, 4)), np.zeros((4, 17)), np.eye(17)[:, :5], T, 1, 0.5, 1.0)


# tolerance follows cond(K_ss) (make_golden_mpc.py prints it: <= 5e3 and <= 2e6): Gamma = diag(K - K_xs K_ss^-1 K_sx)
# carries cond eps / noise relative error in either implementation; 1e-6 is BASELINE.json's fp64 bar
@pytest.mark.parametrize("name,tol", [("fitc_n45_m10", 1e-8), ("fitc_n200_m25", 1e-6)])
def test_fitc_weights_match_reference_output(name, tol):
    from gpmpc_b200 import GaussianProcess
    from gpmpc_b200 import engine as E
    from gpmpc_b200.mpc import _fitc_weights, sparse_gp
    g = np.load(os.path.join(GOLD, name + ".npz"))
    S = torch.tensor(g["sparse_inputs"])
    for k, (X, y, ell, sf2, noise) in enumerate(_gps(g)):
        fit, _ = E.factorize(X, y, ell, sf2, noise)
        w = _fitc_weights(fit, S[:, GP_IDX[k]].cuda().contiguous())
        assert o.rel_err(w.cpu(), g["posterior_mean"][k]) < tol
        # sparse prediction K_f(z, S) w through the mean + Jacobian kernel
        gp = GaussianProcess(X, y)
        gp._pred_fit = fit
        sf = sparse_gp(gp, S[:, GP_IDX[k]], w)
        Z = torch.rand(37, X.shape[1], dtype=torch.float64, generator=torch.Generator().manual_seed(k)) - 0.5
        mean, _, jac = E.predict(sf, Z.cuda(), want_var=False)
        ref = o.se_kernel(Z, S[:, GP_IDX[k]], ell, sf2) @ torch.tensor(g["posterior_mean"][k])
        assert o.rel_err(mean[0].cpu(), ref) < tol
        jref = torch.autograd.functional.jacobian(
            lambda zz: (o.se_kernel(zz, S[:, GP_IDX[k]], ell, sf2) @ torch.tensor(g["posterior_mean"][k])).sum(), Z)
        assert o.rel_err(jac[0].cpu(), jref) < tol


def test_precompute_sparse_posterior_mean_surface():
    """Same call shape as GPMPC.precompute_sparse_posterior_mean: rng-drawn inducing rows, (P, M) weights."""
    from gpmpc_b200 import GaussianProcess
    from gpmpc_b200.mpc import precompute_posterior_mean, precompute_sparse_posterior_mean
    g = np.load(os.path.join(GOLD, "fitc_n45_m10.npz"))
    gps = []
    for X, y, ell, sf2, noise in _gps(g):
        gp = GaussianProcess(X, y)
        gp.set_hyperparameters(ell, sf2, noise)
        gp.eval()
        gps.append(gp)
    w, S = precompute_sparse_posterior_mean(gps, GP_IDX, 10, np_random=np.random.default_rng(21))
    assert w.shape == (3, 10) and S.shape == (10, 7)
    assert np.array_equal(S, g["sparse_inputs"])                                 # same draw as the reference made
    for k, (X, y, ell, sf2, noise) in enumerate(_gps(g)):                      # hyper-parameters pass through float32
        ell32, sf32, n32 = gps[k].hyperparameters()
        assert o.rel_err(w[k], o.fitc_posterior_mean(X, y, torch.tensor(S[:, GP_IDX[k]]), ell32, sf32, n32)) < 1e-8
    assert o.rel_err(w, g["posterior_mean"]) < 1e-3
    alpha, train = precompute_posterior_mean(gps)
    assert alpha.shape == (3, 45) and train.shape == (45, 7)
    for k, (X, y, ell, sf2, noise) in enumerate(_gps(g)):
        ell32, sf32, n32 = gps[k].hyperparameters()
        assert o.rel_err(alpha[k], o.fit_state(X, y, ell32, sf32, n32)["alpha"]) < 1e-9


# ---- batched stage dynamics -------------------------------------------------------------
PRIOR = dict(a=12.1432, b=1.8118, c=-72.08, d=-7.5755, e=39.8653, f=-72.08, h=-7.5755, l=39.8653)   # config yaml:9-17


def _stage_reference(gps, x, u, prior_list, dt):
    means = []
    for X, y, ell, sf2, noise in gps:
        st = o.fit_state(X, y, ell, sf2, noise)
        means.append(o.gp_mean_fn(X, ell, sf2, st["alpha"]))
    out = [o.stage_linearisation(x[i], u[i], prior_list, means, dt) for i in range(x.shape[0])]
    return (torch.stack([t[k] for t in out]) for k in range(3))


@pytest.mark.parametrize("name,dt", [("tighten_T25", 1.0 / 50.0), ("tighten_T40", 1.0 / 60.0)])
def test_stage_dynamics_match_oracle(name, dt):
    from gpmpc_b200.mpc import StageDynamics
    g = np.load(os.path.join(GOLD, name + ".npz"))
    gps = _gps(g)
    sd = StageDynamics(gps, PRIOR, dt)
    gen = torch.Generator().manual_seed(3)
    Q = 19
    x = (torch.rand(Q, 12, generator=gen, dtype=torch.float64) - 0.5) * 0.8
    u = (torch.rand(Q, 4, generator=gen, dtype=torch.float64) - 0.3) * 0.6
    prior_list = [PRIOR[k] for k in "abcdefhl"] + [9.81]
    xr, Ar, Br = _stage_reference(gps, x, u, prior_list, dt)
    xn, A, B = sd.disc_dyn(x.cuda(), u.cuda())
    assert xn.is_cuda and A.shape == (Q, 12, 12) and B.shape == (Q, 12, 4)
    assert o.rel_err(xn.cpu(), xr) < 1e-12
    assert o.rel_err(A.cpu(), Ar) < 1e-11 and o.rel_err(B.cpu(), Br) < 1e-11
    # the GP residual matters in these numbers: without it the step differs visibly
    x0, *_ = _stage_reference([(X, torch.zeros_like(y), ell, sf2, noise) for X, y, ell, sf2, noise in gps], x, u, prior_list, dt)
    assert o.rel_err(x0, xr) > 1e-6
    # numpy in, numpy out, single pair; small scratch forces the chunk loop
    x1, A1, B1 = sd.disc_dyn(x[3].numpy(), u[3].numpy())
    assert isinstance(A1, np.ndarray) and np.array_equal(x1, xn[3].cpu().numpy())
    from gpmpc_b200 import _lib
    small = _lib.load().gpk_stage_rk4_work_bytes(4, sd.n, 3)
    xs, As, Bs = sd.disc_dyn(x.cuda(), u.cuda(), work_bytes=small)
    assert torch.equal(xs, xn) and torch.equal(As, A) and torch.equal(Bs, B)


def test_stage_dynamics_yaw_terms_and_larger_training_set():
    from gpmpc_b200.mpc import StageDynamics
    gen = torch.Generator().manual_seed(9)
    n = 700
    gps = []
    for k, d in enumerate((1, 3, 3)):
        X = torch.rand(n, d, generator=gen, dtype=torch.float64) - 0.5
        y = 0.3 * torch.sin(4.0 * X.sum(dim=1)) + 0.01 * torch.randn(n, generator=gen, dtype=torch.float64)
        gps.append((X, y, torch.full((d,), 0.3 + 0.1 * k, dtype=torch.float64), 0.5, 1e-3))
    prior_list = [PRIOR[k] for k in "abcdefhl"] + [9.81, -3.0, -1.5, 2.0]
    sd = StageDynamics(gps, prior_list, 0.02)
    Q = 7
    x = (torch.rand(Q, 12, generator=gen, dtype=torch.float64) - 0.5)
    u = (torch.rand(Q, 4, generator=gen, dtype=torch.float64) - 0.5)
    xr, Ar, Br = _stage_reference(gps, x, u, prior_list, 0.02)
    xn, A, B = sd.disc_dyn(x.numpy(), u.numpy())
    assert o.rel_err(xn, xr) < 1e-12 and o.rel_err(A, Ar) < 1e-10 and o.rel_err(B, Br) < 1e-10
    assert np.abs(A[:, 11, 8] - 0.02 * -3.0).max() < 1e-2                         # the yaw row is live


def test_stage_dynamics_fitc_model():
    """The sparse model of gpmpc.py:175-186: inducing inputs + FITC weights in place of (X, alpha)."""
    from gpmpc_b200 import GaussianProcess
    from gpmpc_b200 import engine as E
    from gpmpc_b200.mpc import StageDynamics, sparse_gp
    g = np.load(os.path.join(GOLD, "fitc_n45_m10.npz"))
    S = torch.tensor(g["sparse_inputs"])
    fits, means = [], []
    for k, (X, y, ell, sf2, noise) in enumerate(_gps(g)):
        gp = GaussianProcess(X, y)
        gp._pred_fit, _ = E.factorize(X, y, ell, sf2, noise)
        w = torch.tensor(g["posterior_mean"][k])
        fits.append(sparse_gp(gp, S[:, GP_IDX[k]], w))
        means.append(o.gp_mean_fn(S[:, GP_IDX[k]], ell, sf2, w))
    prior_list = [PRIOR[k] for k in "abcdefhl"] + [9.81]
    sd = StageDynamics(fits, PRIOR, 0.02)
    gen = torch.Generator().manual_seed(4)
    x = (torch.rand(5, 12, generator=gen, dtype=torch.float64) - 0.5) * 0.6
    u = (torch.rand(5, 4, generator=gen, dtype=torch.float64) - 0.3) * 0.5
    out = [o.stage_linearisation(x[i], u[i], prior_list, means, 0.02) for i in range(5)]
    xn, A, B = sd.disc_dyn(x.numpy(), u.numpy())
    assert o.rel_err(xn, torch.stack([t[0] for t in out])) < 1e-12
    assert o.rel_err(A, torch.stack([t[1] for t in out])) < 1e-11
    assert o.rel_err(B, torch.stack([t[2] for t in out])) < 1e-11


def test_mpc_kernels_write_only_their_outputs():
    """Outputs carved out of one sentinel-filled allocation: nothing before, between or after them changes."""
    from gpmpc_b200 import _lib
    from gpmpc_b200._lib import call, ptr
    from gpmpc_b200.mpc import StageDynamics
    dev = torch.device("cuda", 0)
    R, T, nx, nu, pad = 5, 7, 12, 4, 64
    sizes = [R * 2 * nx * (T + 1), R * 2 * nu * T]
    big = torch.full((sum(sizes) + 3 * pad,), -7.25, dtype=torch.float64, device=dev)
    sc = big[pad: pad + sizes[0]]
    ic = big[2 * pad + sizes[0]: 2 * pad + sizes[0] + sizes[1]]
    g = torch.Generator().manual_seed(1)
    rnd = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64).to(dev)   # noqa: E731
    A, B, K, Bd = torch.eye(nx, dtype=torch.float64, device=dev), rnd(nx, nu), -rnd(nu, nx), rnd(nx, 5)
    z, var = rnd(R * T, 16), rnd(3, R * T) * 1e-2
    hyp = torch.zeros(3, _lib.load().gpk_hyp_stride(), dtype=torch.float64, device=dev)
    call("gpk_tighten_constraints", ptr(var), var.stride(0), 1, ptr(z), 16, 6, 7, ptr(hyp), ptr(A), ptr(B), ptr(K),
         ptr(Bd), nx, nu, T, R, 0.02, 2.0, sc.data_ptr(), ic.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    guard = torch.ones_like(big, dtype=torch.bool)
    guard[pad: pad + sizes[0]] = False
    guard[2 * pad + sizes[0]: 2 * pad + sizes[0] + sizes[1]] = False
    assert bool((big[guard] == -7.25).all()) and not bool((sc == -7.25).any()) and not bool((ic == -7.25).any())

    q = 21
    gps = [(rnd(40, d).cpu() - 0.5, rnd(40).cpu(), torch.full((d,), 0.5, dtype=torch.float64), 0.5, 1e-2) for d in (1, 3, 3)]
    sd = StageDynamics(gps, [PRIOR[k] for k in "abcdefhl"], 0.02)
    sizes = [q * 12, q * 144, q * 48]
    big = torch.full((sum(sizes) + 4 * pad,), -7.25, dtype=torch.float64, device=dev)
    offs = [pad, 2 * pad + sizes[0], 3 * pad + sizes[0] + sizes[1]]
    outs = [big[o_: o_ + s_] for o_, s_ in zip(offs, sizes)]
    x, u = rnd(q, 12) - 0.5, rnd(q, 4) - 0.5
    lib = _lib.load()
    wb = lib.gpk_stage_rk4_work_bytes(8, sd.n, 3)                           # three chunks: 8 + 8 + 5
    work = torch.empty(wb, dtype=torch.uint8, device=dev)
    call("gpk_stage_rk4", ptr(x), 12, ptr(u), 4, q, sd.prior.data_ptr(), sd.prior.numel(), sd.dt, ptr(sd.hyp),
         ptr(sd.pack), sd.pack.stride(0), sd.n, outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(), ptr(work),
         wb, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    guard = torch.ones_like(big, dtype=torch.bool)
    for o_, s_ in zip(offs, sizes):
        guard[o_: o_ + s_] = False
    assert bool((big[guard] == -7.25).all())
    xr, Ar, Br = sd.disc_dyn(x, u)
    assert torch.equal(outs[0].view(q, 12), xr) and torch.equal(outs[1].view(q, 12, 12), Ar)
    assert torch.equal(outs[2].view(q, 12, 4), Br)
