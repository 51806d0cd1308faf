"""Pins the CPU oracle (oracle/gp_oracle.py) against independent re-derivations.

The reference has no tests or golden vectors (parity unpinned), so the oracle is
checked against: torch autograd of the same loss (what gpmpc/gp.py:61-63 does), the literal
explicit-inverse formulas of gpmpc/gp.py:46,85, an 80-bit long-double loop implementation, central
finite differences, analytic known-answer cases, and the committed fixtures in tests/golden/.
"""
import math
import os
import sys

import numpy as np
import pytest
import torch

from oracle import gp_oracle as o

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _problem(n, d, q, seed=0):
    p = o.synthetic_problem(n, d, q, p=1, seed=seed)
    return p["X"], p["Y"][:, 0], p["Z"], p["ell"], p["sf2"], p["noise"]


@pytest.mark.parametrize("n,d", [(15, 1), (45, 3), (200, 5), (64, 16)])
def test_kernel_formulations_agree(n, d):
    X, y, Z, ell, sf2, noise = _problem(n, d, 33)
    a = o.se_kernel(Z, X, ell, sf2)
    b = o.se_kernel_gpytorch(Z, X, ell, sf2)
    assert o.rel_err(a, b) < 1e-13
    # literal covSE_single formula, one pair at a time (gpmpc/gp.py:12-14)
    i, j = 3, 7
    ref = sf2 * math.exp(-0.5 * float((((Z[i] - X[j]) ** 2) / ell**2).sum()))
    assert abs(float(a[i, j]) - ref) < 1e-15


@pytest.mark.parametrize("n,d,ard", [(30, 1, False), (45, 3, False), (200, 5, True), (120, 16, True)])
def test_closed_form_gradient_matches_autograd(n, d, ard):
    X, y, *_ = _problem(n, d, 1, seed=3)
    g = torch.Generator().manual_seed(5)
    rn = torch.randn(1, generator=g, dtype=torch.float64) - 2.0
    rs = torch.randn((), generator=g, dtype=torch.float64)
    rl = torch.randn(1, d if ard else 1, generator=g, dtype=torch.float64) + 1.0
    l1, gn1, gs1, gl1 = o.loss_and_grad(X, y, rn, rs, rl)
    l2, gn2, gs2, gl2 = o.loss_autograd(X, y, rn, rs, rl)
    assert abs(float(l1 - l2)) < 1e-12 * max(1.0, abs(float(l2)))
    assert o.rel_err(gn1, gn2) < 1e-9
    assert o.rel_err(gs1, gs2) < 1e-9
    assert o.rel_err(gl1, gl2) < 1e-9


@pytest.mark.parametrize("n,d", [(15, 1), (45, 3), (200, 5)])
def test_predict_matches_explicit_inverse_and_autograd_jacobian(n, d):
    X, y, Z, ell, sf2, noise = _problem(n, d, 25, seed=11)
    st = o.fit_state(X, y, ell, sf2, noise)
    mean, var, jac = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    m2, v2 = o.predict_explicit_inverse(Z, X, y, ell, sf2, noise)
    assert o.rel_err(mean, m2) < 1e-9
    assert o.rel_err(var, v2) < 1e-9
    # Jacobian vs autograd of the mean
    Zg = Z.clone().requires_grad_(True)
    k = sf2 * torch.exp(-0.5 * (((Zg[:, None, :] - X[None, :, :]) / ell) ** 2).sum(-1))
    (k @ st["alpha"]).sum().backward()
    assert o.rel_err(jac, Zg.grad) < 1e-12
    # ... and vs central finite differences
    h = 1e-6
    for j in range(d):
        e = torch.zeros(d, dtype=torch.float64)
        e[j] = h
        mp, _, _ = o.predict(Z + e, X, ell, sf2, st["L"], st["alpha"], want_var=False, want_jac=False)
        mm, _, _ = o.predict(Z - e, X, ell, sf2, st["L"], st["alpha"], want_var=False, want_jac=False)
        assert o.rel_err((mp - mm) / (2 * h), jac[:, j]) < 1e-6


def _longdouble_gp(X, y, Z, ell, sf2, noise):
    """Naive loops in numpy long double (80-bit on x86): Cholesky, solves, mean, variance, MLL."""
    ld = np.longdouble
    X = X.numpy().astype(ld); y = y.numpy().astype(ld); Z = Z.numpy().astype(ld)
    ell = np.asarray(ell, dtype=ld); sf2 = ld(sf2); noise = ld(noise)
    n = X.shape[0]
    def kern(a, b):
        return sf2 * np.exp(ld(-0.5) * (((a - b) / ell) ** 2).sum())
    K = np.array([[kern(X[i], X[j]) for j in range(n)] for i in range(n)], dtype=ld) + noise * np.eye(n, dtype=ld)
    L = np.zeros((n, n), dtype=ld)
    for j in range(n):
        s = K[j, j] - (L[j, :j] ** 2).sum()
        L[j, j] = np.sqrt(s)
        for i in range(j + 1, n):
            L[i, j] = (K[i, j] - (L[i, :j] * L[j, :j]).sum()) / L[j, j]
    def fwd(b):
        x = np.zeros(n, dtype=ld)
        for i in range(n):
            x[i] = (b[i] - (L[i, :i] * x[:i]).sum()) / L[i, i]
        return x
    def bwd(b):
        x = np.zeros(n, dtype=ld)
        for i in reversed(range(n)):
            x[i] = (b[i] - (L[i + 1:, i] * x[i + 1:]).sum()) / L[i, i]
        return x
    t = fwd(y); alpha = bwd(t)
    mll = ld(-0.5) * ((t * t).sum() + 2 * np.log(np.diag(L)).sum() + n * np.log(2 * ld(np.pi)))
    means, vars_ = [], []
    for z in Z:
        k = np.array([kern(z, X[i]) for i in range(n)], dtype=ld)
        v = fwd(k)
        means.append((k * alpha).sum()); vars_.append(sf2 - (v * v).sum())
    return float(mll / n), np.array(means, dtype=np.float64), np.array(vars_, dtype=np.float64)


@pytest.mark.parametrize("n,d", [(15, 1), (30, 3), (45, 3)])
def test_oracle_matches_long_double_loops(n, d):
    X, y, Z, ell, sf2, noise = _problem(n, d, 8, seed=21)
    mll_ld, mean_ld, var_ld = _longdouble_gp(X, y, Z, ell.numpy(), sf2, noise)
    st = o.fit_state(X, y, ell, sf2, noise)
    mean, var, _ = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    assert abs(float(o.mll_per_n(X, y, ell, sf2, noise)) - mll_ld) < 1e-11 * abs(mll_ld)
    assert o.rel_err(mean, mean_ld) < 1e-10
    assert o.rel_err(var, var_ld) < 1e-10


def test_known_answers():
    sf2, noise, ell = 1.7, 0.3, torch.tensor([0.8, 1.1])
    # N = 1: mean = sf2 k/(sf2+noise) y ; var = sf2 - k^2/(sf2+noise)
    X = torch.tensor([[0.2, -0.4]], dtype=torch.float64); y = torch.tensor([1.3], dtype=torch.float64)
    Z = torch.tensor([[0.5, 0.1], [0.2, -0.4]], dtype=torch.float64)
    st = o.fit_state(X, y, ell, sf2, noise)
    mean, var, jac = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    k = o.se_kernel(Z, X, ell, sf2)[:, 0]
    assert torch.allclose(mean, k / (sf2 + noise) * y[0], rtol=1e-14)
    assert torch.allclose(var, sf2 - k * k / (sf2 + noise), rtol=1e-14)
    assert torch.allclose(jac[1], torch.zeros(2, dtype=torch.float64), atol=1e-15)  # at the data point
    # query == training point with tiny noise: mean -> y_i, var_f -> ~0
    X, yy, _, ell3, _, _ = _problem(20, 3, 1, seed=2)
    st = o.fit_state(X, yy, ell3, 1.0, 1e-6)
    mean, var, _ = o.predict(X[:5], X, ell3, 1.0, st["L"], st["alpha"])
    assert o.rel_err(mean, yy[:5]) < 1e-3 and float(var.abs().max()) < 1e-5
    # far-away query: mean -> 0, var -> sf2, jac -> 0
    mean, var, jac = o.predict(torch.full((1, 3), 50.0, dtype=torch.float64), X, ell3, 1.0, st["L"], st["alpha"])
    assert abs(float(mean)) < 1e-100 and abs(float(var) - 1.0) < 1e-14 and float(jac.abs().max()) < 1e-100
    # duplicated training rows survive with the noise floor
    Xd = torch.cat([X, X[:3]]); yd = torch.cat([yy, yy[:3]])
    st = o.fit_state(Xd, yd, ell3, 1.0, 1e-6)
    assert torch.isfinite(st["alpha"]).all()


def test_defaults_at_raw_zero():
    noise, sf2, ell = o.constrained(0.0, 0.0, 0.0)
    assert abs(float(noise) - (math.log(2) + 1e-6)) < 1e-15
    assert abs(float(sf2) - math.log(2)) < 1e-15 and abs(float(ell) - math.log(2)) < 1e-15
    y = torch.tensor(0.37, dtype=torch.float64)
    assert abs(float(o.softplus(o.inv_softplus(y))) - 0.37) < 1e-14


def test_fit_loop_decreases_loss_and_stops():
    X, y, *_ = _problem(30, 3, 1, seed=9)
    out = o.fit_hyperparameters(X, y, n_train=200, lr=0.05)
    assert out["losses"][-1] < out["losses"][0]
    assert len(out["losses"]) <= 200
    assert out["raw_ell"].shape == (1, 1) and out["raw_noise"].dtype == torch.float32


def test_fitc_reduces_to_exact_when_all_points_are_inducing():
    X, y, Z, ell, sf2, noise = _problem(25, 3, 10, seed=4)
    w = o.fitc_posterior_mean(X, y, X, ell, sf2, noise)
    st = o.fit_state(X, y, ell, sf2, noise)
    assert o.rel_err(w, st["alpha"]) < 1e-6


def test_golden_fixtures_reproduce():
    """tests/golden/*.npz were written by tests/golden/make_golden.py from this oracle; a change in the
    oracle's arithmetic shows up here."""
    files = sorted(f for f in os.listdir(GOLD) if f.endswith(".npz") and f.startswith(("ref_", "syn_")))
    assert files, "no golden fixtures committed"
    for f in files:
        g = np.load(os.path.join(GOLD, f))
        X, y, Z = (torch.from_numpy(g[k]) for k in ("X", "y", "Z"))
        ell, sf2, noise = torch.from_numpy(g["ell"]), float(g["sf2"]), float(g["noise"])
        st = o.fit_state(X, y, ell, sf2, noise)
        mean, var, jac = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
        assert o.rel_err(st["alpha"], g["alpha"]) < 1e-9, f
        assert o.rel_err(mean, g["mean"]) < 1e-10, f
        assert o.rel_err(var, g["var"]) < 1e-10, f
        assert o.rel_err(jac, g["jac"]) < 1e-10, f
        assert abs(float(o.mll_per_n(X, y, ell, sf2, noise)) - float(g["mll_per_n"])) < 1e-11 * abs(float(g["mll_per_n"])), f
        loss, gn, gs, gl = o.loss_and_grad(X, y, g["raw_noise"], g["raw_sf2"], g["raw_ell"])
        assert abs(float(loss) - float(g["loss_raw"])) < 1e-11 * abs(float(g["loss_raw"])), f
        assert o.rel_err(gl, g["g_raw_ell"]) < 1e-8 and o.rel_err(gn, g["g_raw_noise"]) < 1e-8, f


# ---- rows a1 / a2 / a6: fixtures written by the reference's own covSE_* and gpytorch_predict2casadi ---------
# (tests/golden/make_golden_covse.py runs the unmodified gpmpc/gp.py:12-21,72-85 over a NumPy casadi shim)
COVSE = [f"covse_d{d}_{k}" for d in (1, 3, 16) for k in ("shared", "ard")]


@pytest.mark.parametrize("name", COVSE)
def test_oracle_kernel_matches_reference_covse(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    x, Z = torch.from_numpy(g["x"])[None], torch.from_numpy(g["Z"])
    ell, sf2 = torch.from_numpy(g["ell"]).reshape(-1), float(g["sf2"])
    assert np.array_equal(g["k_single"], g["k_vec"]) or np.abs(g["k_single"] - g["k_vec"]).max() < 1e-15
    for fn in (o.se_kernel, o.se_kernel_gpytorch):
        k = fn(x, Z, ell, sf2).numpy()
        # element-wise relative error: the kernel is a product of exponentials, no cancellation
        assert np.abs(k - g["k_single"]).max() < 1e-14 * sf2, (name, fn.__name__)
        assert (np.abs(k - g["k_single"]) / g["k_single"]).max() < 2e-13, (name, fn.__name__)


@pytest.mark.parametrize("name", ["pred_d3_n45", "pred_d16_n300"])
def test_oracle_mean_matches_reference_casadi_mean(name):
    """mean = k(z,X) @ K_inv @ y exactly as the reference's exported CasADi function evaluates it (gp.py:85, with
    K_inv from torch.linalg.inv, gp.py:46) against the oracle's Cholesky form."""
    g = np.load(os.path.join(GOLD, name + ".npz"))
    X, y, Z = (torch.from_numpy(g[k]) for k in ("X", "y", "Z"))
    ell, sf2, noise = torch.from_numpy(g["ell"]).reshape(-1), float(g["sf2"]), float(g["noise"])
    st = o.fit_state(X, y, ell, sf2, noise)
    mean, _, _ = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    assert o.rel_err(mean, g["mean"]) < 1e-10, name
    if "K" in g:
        assert o.rel_err(st["K"], g["K"]) < 1e-14, name


def test_repo_casadi_twins_match_reference_bitwise(monkeypatch):
    """gpmpc_b200.gp.covSE_single / covSE_vectorized executed over the same NumPy shim reproduce the reference's
    outputs bit for bit (same operation order)."""
    import casadi_shim
    from gpmpc_b200 import gp as G

    monkeypatch.setattr(G, "cs", casadi_shim)
    for name in COVSE:
        g = np.load(os.path.join(GOLD, name + ".npz"))
        ks = np.asarray(G.covSE_single(g["x"][:, None], g["Z"].T, g["ell"].T, float(g["sf2"])))
        kv = np.asarray(G.covSE_vectorized(g["x"][:, None], g["Z"], g["ell"].T, float(g["sf2"])))
        assert np.array_equal(ks, g["k_single"]), name
        assert np.array_equal(kv, g["k_vec"]), name
    # symbolic use, as gpytorch_predict2casadi does: a Function of z
    g = np.load(os.path.join(GOLD, "covse_d3_ard.npz"))
    z = casadi_shim.SX.sym("z", 3)
    f = casadi_shim.Function("K_xz", [z], [G.covSE_single(z, g["Z"].T, g["ell"].T, float(g["sf2"]))], ["z"], ["K"])
    assert np.array_equal(np.asarray(f(z=g["x"])["K"]), g["k_single"])


# ---- MPC-side rows: fixtures written by the reference's own methods (make_golden_mpc.py) ------
@pytest.mark.parametrize("name", ["tighten_T25", "tighten_T40"])
def test_tightening_restatement_matches_reference_output(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    for r in range(g["x_prev"].shape[0]):
        z = np.hstack((g["x_prev"][r][:, :-1].T, g["u_prev"][r].T))          # gpmpc/gpmpc.py:437
        sc, ic = o.propagate_constraint_limits(g["var_y"][r], z, g["noise"], g["dfdx"], g["dfdu"], g["lqr_gain"],
                                               g["Bd"], float(g["dt"]), float(g["inverse_cdf"]))
        assert sc.shape == g["state_constraint"][r].shape and ic.shape == g["input_constraint"][r].shape
        assert o.rel_err(sc, g["state_constraint"][r]) < 1e-12
        assert o.rel_err(ic, g["input_constraint"][r]) < 1e-12
        assert np.all(sc[:, 0] == 0.0) and np.all(sc <= 0.0)                  # cov_x starts at 0; limits only tighten


@pytest.mark.parametrize("name", ["tighten_T25", "tighten_T40"])
def test_tightening_fixture_variances_come_from_the_gp_oracle(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    gp_idx = [[0], [1, 2, 3], [4, 5, 6]]
    z = torch.tensor(np.hstack((g["x_prev"][0][:, :-1].T, g["u_prev"][0].T)))
    for k, idx in enumerate(gp_idx):
        X, y, ell = (torch.tensor(g[f"gp{k}_{f}"]) for f in ("X", "y", "ell"))
        st = o.fit_state(X, y, ell, float(g[f"gp{k}_sf2"]), float(g["noise"][k]))
        _, var, _ = o.predict(z[:, idx], X, ell, float(g[f"gp{k}_sf2"]), st["L"], st["alpha"])
        assert o.rel_err(var + float(g["noise"][k]), g["var_y"][0][k]) < 1e-12


@pytest.mark.parametrize("name", ["fitc_n45_m10", "fitc_n200_m25"])
def test_fitc_restatement_matches_reference_output(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    gp_idx = [[0], [1, 2, 3], [4, 5, 6]]
    S = torch.tensor(g["sparse_inputs"])
    for k, idx in enumerate(gp_idx):
        X, y, ell = (torch.tensor(g[f"gp{k}_{f}"]) for f in ("X", "y", "ell"))
        w = o.fitc_posterior_mean(X, y, S[:, idx], ell, float(g[f"gp{k}_sf2"]), float(g["noise"][k]))
        assert o.rel_err(w, g["posterior_mean"][k]) < 1e-9


def test_stage_linearisation_matches_finite_differences_and_plain_rk4():
    """The stage-dynamics restatement (gpmpc/gpmpc.py:193-206): autograd Jacobians against central differences, and
    with zero GP residual the step is classical RK4 of the prior alone."""
    gen = torch.Generator().manual_seed(0)
    prior = [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653, 9.81]
    means = []
    for d in (1, 3, 3):
        X = torch.rand(20, d, generator=gen, dtype=torch.float64) - 0.5
        alpha = torch.randn(20, generator=gen, dtype=torch.float64)
        means.append(o.gp_mean_fn(X, torch.full((d,), 0.4, dtype=torch.float64), 0.7, alpha))
    x = (torch.rand(12, generator=gen, dtype=torch.float64) - 0.5) * 0.6
    u = (torch.rand(4, generator=gen, dtype=torch.float64) - 0.5) * 0.6
    dt = 0.02
    fx, A, B = o.stage_linearisation(x, u, prior, means, dt)
    h = 1e-6
    for j in range(12):
        e = torch.zeros(12, dtype=torch.float64)
        e[j] = h
        fd = (o.stage_f_disc(x + e, u, prior, means, dt) - o.stage_f_disc(x - e, u, prior, means, dt)) / (2 * h)
        assert float((fd - A[:, j]).abs().max()) < 1e-7
    for j in range(4):
        e = torch.zeros(4, dtype=torch.float64)
        e[j] = h
        fd = (o.stage_f_disc(x, u + e, prior, means, dt) - o.stage_f_disc(x, u - e, prior, means, dt)) / (2 * h)
        assert float((fd - B[:, j]).abs().max()) < 1e-7
    zero = [lambda z: torch.zeros((), dtype=torch.float64)] * 3
    f = lambda xx: o.attitude_prior_dynamics(xx, u, prior)   # noqa: E731
    k1 = f(x); k2 = f(x + dt / 2 * k1); k3 = f(x + dt / 2 * k2); k4 = f(x + dt * k3)   # noqa: E702
    assert torch.equal(o.stage_f_disc(x, u, prior, zero, dt), x + dt / 6 * (k1 + 2 * k2 + 2 * k3 + k4))
    # hover: thrust command (g - b) / a keeps the vertical acceleration at zero
    xh, uh = torch.zeros(12, dtype=torch.float64), torch.tensor([(9.81 - prior[1]) / prior[0], 0, 0, 0], dtype=torch.float64)
    assert float(o.stage_f_disc(xh, uh, prior, zero, dt).abs().max()) < 1e-15


# ---- rows a7-a9 against an INDEPENDENT third-party implementation (scikit-learn GPR; tests/golden/make_golden_sklearn.py) ---
SK = ["sk_d1_n15", "sk_d3_n45", "sk_d16_n300"]


def _oracle_log_gradient(X, y, ell, sf2, noise):
    """d log p(y|X) / d log(sf2, ell_1..d, noise) from the oracle's closed-form raw-parameter gradient."""
    n = X.shape[0]
    raw_noise = o.inv_softplus(torch.tensor([noise - o.NOISE_FLOOR], dtype=torch.float64))
    raw_sf2 = o.inv_softplus(torch.tensor(sf2, dtype=torch.float64))
    raw_ell = o.inv_softplus(ell.reshape(1, -1))
    loss, gn, gs, gl = o.loss_and_grad(X, y, raw_noise, raw_sf2, raw_ell)
    # loss = -lml / n ; d/d theta = (d/d raw) / sigmoid(raw) ; d/d log theta = theta d/d theta
    d_noise = -n * gn / torch.sigmoid(raw_noise) * noise
    d_sf2 = -n * gs / torch.sigmoid(raw_sf2) * sf2
    d_ell = -n * gl / torch.sigmoid(raw_ell) * ell.reshape(1, -1)
    return -n * float(loss), torch.cat([d_sf2.reshape(1), d_ell.reshape(-1), d_noise.reshape(1)])


@pytest.mark.parametrize("name", SK)
def test_oracle_matches_scikit_learn_gpr(name):
    """Exact log marginal likelihood, its gradient, posterior mean and predictive variance against scikit-learn's
    GaussianProcessRegressor (sf2 * RBF(ARD) + White) on the committed fixtures: an implementation that shares no code
    with the oracle.  gpytorch's ExactMarginalLogLikelihood (gp.py:57) is this lml divided by N."""
    g = np.load(os.path.join(GOLD, name + ".npz"))
    X, y, Z, ell = (torch.from_numpy(g[k]) for k in ("X", "y", "Z", "ell"))
    sf2, noise, n = float(g["sf2"]), float(g["noise"]), X.shape[0]
    lml, grad = _oracle_log_gradient(X, y, ell, sf2, noise)
    assert abs(lml - float(g["lml"])) < 1e-9 * abs(float(g["lml"]))
    assert abs(float(o.mll_per_n(X, y, ell, sf2, noise)) * n - float(g["lml"])) < 1e-9 * abs(float(g["lml"]))
    assert o.rel_err(grad, g["grad_log"]) < 1e-7
    st = o.fit_state(X, y, ell, sf2, noise)
    mean, var, _ = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    assert o.rel_err(mean, g["mean"]) < 1e-8
    assert float(((var + noise) - torch.from_numpy(g["var_y"])).abs().max() / torch.from_numpy(g["var_y"]).abs().min()) < 1e-6


def test_sklearn_fixtures_regenerate_when_sklearn_is_present():
    """The committed sk_*.npz are what scikit-learn computes here (skipped where scikit-learn is not installed)."""
    pytest.importorskip("sklearn")
    sys.path.insert(0, GOLD)
    import make_golden_sklearn as mk

    g = np.load(os.path.join(GOLD, "sk_d3_n45.npz"))
    lml, grad, mean, var_y = mk.sklearn_gp(g["X"], g["y"], g["Z"], g["ell"], float(g["sf2"]), float(g["noise"]))
    assert abs(lml - float(g["lml"])) < 1e-10 * abs(lml) and np.abs(grad - g["grad_log"]).max() < 1e-8 * np.abs(grad).max()
    assert np.abs(mean - g["mean"]).max() < 1e-10 and np.abs(var_y - g["var_y"]).max() < 1e-10
