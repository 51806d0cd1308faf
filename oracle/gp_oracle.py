"""CPU oracle for the gp-mpc Gaussian-process hot path  --  TEST INFRASTRUCTURE ONLY.

This file restates, in plain torch float64 on the CPU, the arithmetic that the reference's GP layer
(`gpmpc/gp.py`, called from `gpmpc/gpmpc.py`) runs inside gpytorch 1.12.0 / linear_operator / torch.
Nothing under `gpmpc_b200/` may import it: only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` use it, and only as the checker / the timed CPU
baseline, never as the product path.

PARITY STATUS.  The reference ships no tests, fixtures or golden vectors, and most of its
arithmetic lives in third-party packages that are not installed and not installable here (gpytorch==1.12.0 pinned in
reference `pyproject.toml:12`, `linear_operator`, casadi).
  PINNED to outputs of the reference's own code (fixtures under tests/golden/, generating scripts committed):
    * se_kernel / se_kernel_gpytorch  <- unmodified `covSE_single`, `covSE_vectorized` (gpmpc/gp.py:12-21) executed over a
      NumPy casadi shim: covse_*.npz (make_golden_covse.py), <= 2e-13 element-wise
    * assemble_K, the posterior mean  <- unmodified `gpytorch_predict2casadi` (gpmpc/gp.py:72-85) with K_inv from
      torch.linalg.inv as at gp.py:46: pred_*.npz, <= 1e-10
    * propagate_constraint_limits, fitc_posterior_mean  <- unmodified GPMPC methods (gpmpc/gpmpc.py:377-400, 425-507):
      tighten_*.npz, fitc_*.npz (make_golden_mpc.py)
  PARITY UNPINNED: what gpytorch does internally (softplus constraints and the 1e-6 noise floor, MLL / N, Cholesky,
    the variance of the prediction strategy, the Adam loop).  These follow the published gpytorch 1.12.0 semantics listed
    in SURVEY.md section 8c and are checked only against independent re-derivations (`tests/test_oracle.py`: torch
    autograd, explicit-inverse formulas, an 80-bit long-double loop implementation, finite differences and analytic
    known-answer cases).

Conventions (reference file:line in brackets):
  K_f(a, b) = sf2 * exp(-1/2 * sum_j (a_j - b_j)^2 / ell_j^2)              [gpmpc/gp.py:12-14, 17-21]
  K         = K_f(X, X) + noise * I                                        [gpmpc/gp.py:45]
  noise     = softplus(raw_noise) + 1e-6 ; sf2 = softplus(raw_sf2) ; ell = softplus(raw_ell)
                                                                           [gpmpc/gp.py:31,34; gpytorch constraints]
  loss      = -MLL / N,  MLL = -1/2 (y^T K^-1 y + log det K + N log 2 pi)  [gpmpc/gp.py:57,62]
  mean(z)   = K_f(z, X) K^-1 y                                             [gpmpc/gp.py:85]
  var_f(z)  = sf2 - K_f(z, X) K^-1 K_f(X, z) ; var_y = var_f + noise       [gpmpc/gpmpc.py:443-445]
  jac(z)    = d mean / d z = ((k * alpha) X - z * mean) / ell^2            [CasADi AD of gpmpc/gp.py:12-14,85]
Exact Cholesky at every N (the shipped config never leaves gpytorch's exact branch, N <= 45 < 800).
"""

from __future__ import annotations

import math

import numpy as np
import torch

NOISE_FLOOR = 1e-6  # GreaterThan(1e-6), gpmpc/gp.py:31
LOG_2PI = math.log(2.0 * math.pi)


# ----------------------------------------------------------------------------------------------
# constraints (gpytorch Positive / GreaterThan use softplus)
# ----------------------------------------------------------------------------------------------
def softplus(x: torch.Tensor) -> torch.Tensor:
    return torch.nn.functional.softplus(x)


def inv_softplus(y: torch.Tensor) -> torch.Tensor:
    return y + torch.log(-torch.expm1(-y))


def constrained(raw_noise, raw_sf2, raw_ell):
    """raw (unconstrained) parameters -> (noise, sf2, ell) as float64 tensors."""
    raw_noise = torch.as_tensor(raw_noise, dtype=torch.float64)
    raw_sf2 = torch.as_tensor(raw_sf2, dtype=torch.float64)
    raw_ell = torch.as_tensor(raw_ell, dtype=torch.float64)
    return softplus(raw_noise) + NOISE_FLOOR, softplus(raw_sf2), softplus(raw_ell)


def _ell_vec(ell, d: int) -> torch.Tensor:
    ell = torch.as_tensor(ell, dtype=torch.float64).reshape(-1)
    if ell.numel() == 1:
        ell = ell.expand(d)
    assert ell.numel() == d, f"lengthscale has {ell.numel()} entries, expected 1 or {d}"
    return ell


# ----------------------------------------------------------------------------------------------
# kernel
# ----------------------------------------------------------------------------------------------
def se_kernel(A: torch.Tensor, B: torch.Tensor, ell, sf2) -> torch.Tensor:
    """ARD squared-exponential kernel by direct differences (gpmpc/gp.py:12-14). (n,d),(m,d)->(n,m)."""
    A = A.to(torch.float64)
    B = B.to(torch.float64)
    ell = _ell_vec(ell, A.shape[1])
    As, Bs = A / ell, B / ell
    r2 = torch.zeros(A.shape[0], B.shape[0], dtype=torch.float64)
    for j in range(A.shape[1]):  # column loop: O(nm) memory, no (n,m,d) temporary
        diff = As[:, j : j + 1] - Bs[:, j : j + 1].T
        r2 += diff * diff
    return float(sf2) * torch.exp(-0.5 * r2)


def se_kernel_gpytorch(A: torch.Tensor, B: torch.Tensor, ell, sf2) -> torch.Tensor:
    """Same kernel the way gpytorch's RBFKernel forms it: inputs / ell,
    mean-centred, ||a||^2 + ||b||^2 - 2 a.b by GEMM, clamped at 0.  Used only to check that the two
    formulations agree to rounding."""
    A = A.to(torch.float64)
    B = B.to(torch.float64)
    ell = _ell_vec(ell, A.shape[1])
    As, Bs = A / ell, B / ell
    adj = As.mean(dim=0, keepdim=True)
    As, Bs = As - adj, Bs - adj
    r2 = (As * As).sum(1, keepdim=True) + (Bs * Bs).sum(1, keepdim=True).T - 2.0 * As @ Bs.T
    return float(sf2) * torch.exp(-0.5 * r2.clamp_min(0.0))


def assemble_K(X: torch.Tensor, ell, sf2, noise) -> torch.Tensor:
    """K = K_f(X,X) + noise I   (gpmpc/gp.py:45)."""
    K = se_kernel(X, X, ell, sf2)
    K.diagonal().add_(float(noise))
    return K


# ----------------------------------------------------------------------------------------------
# fit state, MLL, gradient
# ----------------------------------------------------------------------------------------------
def fit_state(X: torch.Tensor, y: torch.Tensor, ell, sf2, noise) -> dict:
    """Cholesky state used by every prediction: L L^T = K, alpha = K^-1 y (gpytorch mean_cache)."""
    y = y.to(torch.float64)
    K = assemble_K(X, ell, sf2, noise)
    L, info = torch.linalg.cholesky_ex(K)
    if int(info) != 0:
        raise np.linalg.LinAlgError(f"K is not positive definite (potrf info={int(info)})")
    alpha = torch.cholesky_solve(y[:, None], L)[:, 0]
    logdet = 2.0 * torch.log(L.diagonal()).sum()
    quad = torch.dot(y, alpha)
    return {"K": K, "L": L, "alpha": alpha, "logdet": logdet, "quad": quad}


def mll_per_n(X, y, ell, sf2, noise) -> torch.Tensor:
    """Exact marginal log likelihood divided by N (what ExactMarginalLogLikelihood returns)."""
    st = fit_state(X, y, ell, sf2, noise)
    n = X.shape[0]
    return -0.5 * (st["quad"] + st["logdet"] + n * LOG_2PI) / n


def loss_and_grad(X: torch.Tensor, y: torch.Tensor, raw_noise, raw_sf2, raw_ell):
    """loss = -MLL/N and its gradient w.r.t. the RAW parameters, closed form:
        G = alpha alpha^T - K^-1
        dMLL/dnoise = 1/2 tr G ;  dMLL/dsf2 = 1/2 sum(G * K_f) / sf2
        dMLL/dell_j = 1/2 sum_ab G_ab K_f,ab (x_aj - x_bj)^2 / ell_j^3
        chain through softplus: d softplus(raw) / d raw = sigmoid(raw).
    raw_ell may have 1 entry (the reference's shared lengthscale, gpmpc/gp.py:34) or d entries (ARD).
    Returns (loss, g_raw_noise, g_raw_sf2, g_raw_ell) as float64 tensors; g_raw_ell has raw_ell's shape."""
    X = X.to(torch.float64)
    y = y.to(torch.float64)
    n, d = X.shape
    raw_noise = torch.as_tensor(raw_noise, dtype=torch.float64)
    raw_sf2 = torch.as_tensor(raw_sf2, dtype=torch.float64)
    raw_ell_t = torch.as_tensor(raw_ell, dtype=torch.float64)
    noise, sf2, ell = constrained(raw_noise, raw_sf2, raw_ell_t)
    ellv = _ell_vec(ell, d)
    st = fit_state(X, y, ellv, sf2, noise)
    Kf = st["K"].clone()
    Kf.diagonal().sub_(float(noise))
    Kinv = torch.cholesky_inverse(st["L"])
    a = st["alpha"]
    G = torch.outer(a, a) - Kinv
    mll = -0.5 * (st["quad"] + st["logdet"] + n * LOG_2PI)
    g_noise = 0.5 * G.diagonal().sum()
    GK = G * Kf
    g_sf2 = 0.5 * GK.sum() / sf2
    g_ell = torch.empty(d, dtype=torch.float64)
    for j in range(d):
        diff = X[:, j : j + 1] - X[:, j : j + 1].T
        g_ell[j] = 0.5 * (GK * diff * diff).sum() / ellv[j] ** 3
    if raw_ell_t.numel() == 1:
        g_ell = g_ell.sum().reshape(raw_ell_t.shape)
    else:
        g_ell = g_ell.reshape(raw_ell_t.shape)
    # loss = -MLL/N ; chain rule through softplus
    s = -1.0 / n
    return (
        s * mll,
        s * g_noise * torch.sigmoid(raw_noise),
        s * g_sf2 * torch.sigmoid(raw_sf2),
        s * g_ell * torch.sigmoid(raw_ell_t),
    )


def loss_autograd(X, y, raw_noise, raw_sf2, raw_ell):
    """Same loss and gradient through torch autograd (what the reference does at gpmpc/gp.py:61-63)."""
    X = X.to(torch.float64)
    y = y.to(torch.float64)
    n, d = X.shape
    rn = torch.as_tensor(raw_noise, dtype=torch.float64).clone().requires_grad_(True)
    rs = torch.as_tensor(raw_sf2, dtype=torch.float64).clone().requires_grad_(True)
    rl = torch.as_tensor(raw_ell, dtype=torch.float64).clone().requires_grad_(True)
    noise = softplus(rn) + NOISE_FLOOR
    sf2 = softplus(rs)
    ell = softplus(rl).reshape(-1)
    ell = ell.expand(d) if ell.numel() == 1 else ell
    Xs = X / ell
    r2 = ((Xs[:, None, :] - Xs[None, :, :]) ** 2).sum(-1)
    K = sf2 * torch.exp(-0.5 * r2) + noise * torch.eye(n, dtype=torch.float64)
    L = torch.linalg.cholesky(K)
    alpha = torch.cholesky_solve(y[:, None], L)[:, 0]
    mll = -0.5 * (torch.dot(y, alpha) + 2.0 * torch.log(L.diagonal()).sum() + n * LOG_2PI)
    loss = -mll / n
    loss.backward()
    return loss.detach(), rn.grad, rs.grad, rl.grad


# ----------------------------------------------------------------------------------------------
# prediction
# ----------------------------------------------------------------------------------------------
def predict(Z: torch.Tensor, X: torch.Tensor, ell, sf2, L: torch.Tensor, alpha: torch.Tensor,
            want_var: bool = True, want_jac: bool = True, chunk: int = 4096):
    """Posterior mean (Q,), latent variance var_f (Q,) and input-Jacobian (Q,d) at queries Z (Q,d).
    mean = k alpha [gp.py:85]; var_f = sf2 - ||L^-1 k||^2 [gpmpc.py:443-445 keeps only this diagonal,
    plus the likelihood noise]; jac = ((k*alpha) X - z mean)/ell^2."""
    Z = Z.to(torch.float64)
    X = X.to(torch.float64)
    q, d = Z.shape
    ellv = _ell_vec(ell, d)
    mean = torch.empty(q, dtype=torch.float64)
    var = torch.empty(q, dtype=torch.float64) if want_var else None
    jac = torch.empty(q, d, dtype=torch.float64) if want_jac else None
    for s in range(0, q, chunk):
        z = Z[s : s + chunk]
        k = se_kernel(z, X, ellv, sf2)  # (c, N)
        m = k @ alpha
        mean[s : s + chunk] = m
        if want_jac:
            jac[s : s + chunk] = ((k * alpha) @ X - z * m[:, None]) / (ellv * ellv)
        if want_var:
            v = torch.linalg.solve_triangular(L, k.T, upper=False)  # (N, c)
            var[s : s + chunk] = float(sf2) - (v * v).sum(0)
    return mean, var, jac


def predict_explicit_inverse(Z, X, y, ell, sf2, noise):
    """The literal reference formulas with a dense inverse (gpmpc/gp.py:46,85): independent check."""
    K = assemble_K(X, ell, sf2, noise)
    Kinv = torch.linalg.inv(K)
    k = se_kernel(Z, X, ell, sf2)
    mean = k @ Kinv @ y.to(torch.float64)
    var = float(sf2) - ((k @ Kinv) * k).sum(1)
    return mean, var


# ----------------------------------------------------------------------------------------------
# hyper-parameter fit (restates gpmpc/gp.py:49-69)
# ----------------------------------------------------------------------------------------------
def fit_hyperparameters(X, y, n_train=500, lr=0.01, ard=False, param_dtype=torch.float32):
    """Adam on loss=-MLL/N from raw=0, early stop when |last_loss - loss| < 1e-3 (gpmpc/gp.py:56-67).
    gpytorch parameters are float32 while the data is float64: the raw
    parameters and Adam moments are kept in `param_dtype`, the linear algebra runs in float64.
    Returns dict(raw_noise, raw_sf2, raw_ell, losses)."""
    d = X.shape[1]
    rn = torch.zeros(1, dtype=param_dtype, requires_grad=True)
    rs = torch.zeros((), dtype=param_dtype, requires_grad=True)
    rl = torch.zeros(1, d if ard else 1, dtype=param_dtype, requires_grad=True)
    optim = torch.optim.Adam([rn, rs, rl], lr=lr)
    last = torch.tensor(torch.inf, dtype=torch.float64)
    losses = []
    for _ in range(n_train):
        optim.zero_grad()
        loss, gn, gs, gl = loss_and_grad(X, y, rn.detach().double(), rs.detach().double(), rl.detach().double())
        rn.grad = gn.to(param_dtype).reshape(rn.shape)
        rs.grad = gs.to(param_dtype).reshape(rs.shape)
        rl.grad = gl.to(param_dtype).reshape(rl.shape)
        optim.step()
        losses.append(float(loss))
        if torch.abs(last - loss) < 1e-3:
            break
        last = loss
    return {"raw_noise": rn.detach(), "raw_sf2": rs.detach(), "raw_ell": rl.detach(), "losses": losses}


# ----------------------------------------------------------------------------------------------
# FITC sparse posterior weights (restates gpmpc/gpmpc.py:391-398)
# ----------------------------------------------------------------------------------------------
def fitc_posterior_mean(X, y, S, ell, sf2, noise):
    """w (M,) such that mean_sparse(z) = K_f(z, S) w   (gpmpc/gpmpc.py:391-398)."""
    X = X.to(torch.float64)
    y = y.to(torch.float64)
    S = S.to(torch.float64)
    K = assemble_K(X, ell, sf2, noise)
    K_ss = se_kernel(S, S, ell, sf2)
    K_xs = se_kernel(X, S, ell, sf2)
    Gamma = torch.diagonal(K - K_xs @ torch.linalg.solve(K_ss, K_xs.T))
    Gi = 1.0 / Gamma
    Sigma_inv = K_ss + (K_xs.T * Gi) @ K_xs
    return torch.linalg.solve(Sigma_inv, K_xs.T) @ (Gi * y)


# ----------------------------------------------------------------------------------------------
# chance-constraint tightening (restates gpmpc/gpmpc.py:425-498)
# PINNED: tests/golden/tighten_*.npz hold outputs of the reference's own GPMPC.propagate_constraint_limits
# executed here (tests/golden/make_golden_mpc.py loads gpmpc/gpmpc.py with its absent third-party imports
# stubbed and feeds it GP objects that return prescribed variances); tests/test_oracle.py checks this
# restatement against them.
# ----------------------------------------------------------------------------------------------
def propagate_constraint_limits(var_y, z, noise, dfdx, dfdu, lqr_gain, Bd, dt, inverse_cdf, idx_phi=6, idx_theta=7):
    """One rollout.  var_y (3, T): predictive variances diag(gp.likelihood(gp(z)).covariance_matrix) of the
    thrust / roll / pitch GPs [gpmpc.py:441-445]; z (T, >= 8) stage inputs; noise (3,) likelihood noises.
    Returns state_constraint (2 nx, T+1), input_constraint (2 nu, T)."""
    var_y = np.asarray(var_y, dtype=np.float64)
    z = np.asarray(z, dtype=np.float64)
    T = z.shape[0]
    nx, nu = dfdx.shape[0], dfdu.shape[1]
    state_constraint = np.zeros((2 * nx, T + 1))
    input_constraint = np.zeros((2 * nu, T))
    cps2 = np.cos(z[:, idx_phi]) * np.sin(z[:, idx_theta]) ** 2                 # [gpmpc.py:448-450]
    sp2 = (-np.sin(z[:, idx_phi])) ** 2
    cpct2 = (np.cos(z[:, idx_phi]) * np.cos(z[:, idx_theta])) ** 2
    fac = np.stack([cps2, sp2, cpct2, np.ones(T), np.ones(T)], axis=1)          # (T, 5)
    gp_of = [0, 0, 0, 1, 2]
    cov_d = np.stack([var_y[g] for g in gp_of], axis=1) * fac * dt**2           # [gpmpc.py:452-457, 469]
    cov_n = np.stack([np.full(T, float(noise[g])) for g in gp_of], axis=1) * fac * dt**2   # [gpmpc.py:459-468]
    absA_x = np.abs(np.concatenate([-np.eye(nx), np.eye(nx)], axis=0))          # [gpmpc.py:472-476]
    absA_u = np.abs(np.concatenate([-np.eye(nu), np.eye(nu)], axis=0))
    cov_x = np.zeros((nx, nx))                                                  # [gpmpc.py:436]
    for i in range(T):                                                          # [gpmpc.py:478-495]
        cov_xu = cov_x @ lqr_gain.T
        cov_u = lqr_gain @ cov_x @ lqr_gain.T
        cd = np.diag(cov_d[i] + cov_n[i])
        state_constraint[:, i] = -inverse_cdf * absA_x @ np.sqrt(np.diag(cov_x))
        input_constraint[:, i] = -inverse_cdf * absA_u @ np.sqrt(np.diag(cov_u))
        cov_x = (dfdx @ cov_x @ dfdx.T + dfdx @ cov_xu @ dfdu.T + dfdu @ cov_xu.T @ dfdx.T
                 + dfdu @ cov_u @ dfdu.T + Bd @ cd @ Bd.T)
    state_constraint[:, -1] = -inverse_cdf * absA_x @ np.sqrt(np.diag(cov_x))   # [gpmpc.py:497]
    return state_constraint, input_constraint


# ----------------------------------------------------------------------------------------------
# discrete stage dynamics of the acados model (restates gpmpc/gpmpc.py:166-209)
# PARITY UNPINNED for the prior: `prior_dynamics` is crazyflow.sim.symbolic.symbolic_attitude(...).fc_func
# (scripts/run_gp_mpc.py:146; crazyflow ^0.0.1, pyproject.toml:11), a third-party package that is absent here.
# It is restated from how the reference uses it -- acc_prior(thrust) = a T + b along the body z axis with the
# rotation column of gpmpc.py:194-196, attitude rates c phi + d dphi + e cmd etc. (gpmpc.py:128-146,
# scripts/gp_mpc_config.yaml:9-17).  The GP residual, its placement and the RK4 are the reference's own lines.
# Jacobians come from torch autograd here (CasADi AD in the reference).
# ----------------------------------------------------------------------------------------------
def attitude_prior_dynamics(x, u, p):
    """p = (a, b, c, d, e, f, h, l, g[, yaw c, d, e]); x (12,) = (x, dx, y, dy, z, dz, phi, theta, psi, dphi, dtheta,
    dpsi); u (4,) = (thrust, roll, pitch, yaw commands)."""
    p = list(p) + [0.0] * (12 - len(p))
    a, b, c, d, e, f, h, l, g, yc, yd, ye = p
    acc = a * u[0] + b
    cp, sp, ct, st = torch.cos(x[6]), torch.sin(x[6]), torch.cos(x[7]), torch.sin(x[7])
    return torch.stack([
        x[1], acc * cp * st, x[3], acc * (-sp), x[5], acc * cp * ct - g, x[9], x[10], x[11],
        c * x[6] + d * x[9] + e * u[1], f * x[7] + h * x[10] + l * u[2], yc * x[8] + yd * x[11] + ye * u[3]])


def gp_mean_fn(X, ell, sf2, alpha):
    """z -> k(z, X) alpha as a differentiable torch expression (gpmpc/gp.py:85)."""
    return lambda z: (se_kernel(z.reshape(1, -1), X, ell, sf2) @ alpha).reshape(())


def stage_f_disc(x, u, prior, gp_means, dt):
    """One RK4 step of f_cont = prior + res_dyn [gpmpc.py:172-206]; gp_means = (T, R, P) mean functions."""
    nx = 12
    zero = torch.zeros((), dtype=x.dtype)

    def f_cont(xx):
        z = torch.cat([xx, u])
        T_pred = gp_means[0](z[[0 + nx]])                                  # idx_T  [gpmpc.py:172]
        R_pred = gp_means[1](z[[6, 9, 1 + nx]])                            # idx_R
        P_pred = gp_means[2](z[[7, 10, 2 + nx]])                           # idx_P
        ax = T_pred * (torch.cos(xx[6]) * torch.sin(xx[7]))               # [gpmpc.py:194-196]
        ay = T_pred * (-torch.sin(xx[6]))
        az = T_pred * (torch.cos(xx[6]) * torch.cos(xx[7]))
        res = torch.stack([zero, ax, zero, ay, zero, az, zero, zero, zero, R_pred, P_pred, zero])   # [197]
        return attitude_prior_dynamics(xx, u, prior) + res                # [199]

    k1 = f_cont(x)                                                         # [gpmpc.py:202-206]
    k2 = f_cont(x + dt / 2 * k1)
    k3 = f_cont(x + dt / 2 * k2)
    k4 = f_cont(x + dt * k3)
    return x + dt / 6 * (k1 + 2 * k2 + 2 * k3 + k4)


def stage_linearisation(x, u, prior, gp_means, dt):
    """(f_disc, A = d f_disc/dx, B = d f_disc/du) for one (x, u)."""
    fx = stage_f_disc(x, u, prior, gp_means, dt)
    A, B = torch.autograd.functional.jacobian(lambda xx, uu: stage_f_disc(xx, uu, prior, gp_means, dt), (x, u))
    return fx, A, B


# ----------------------------------------------------------------------------------------------
# synthetic data recipe shared by tests and bench
# ----------------------------------------------------------------------------------------------
def synthetic_problem(n: int, d: int, q: int, p: int = 1, seed: int = 1337):
    """X ~ U(-1,1)^(n,d); y_p = sin(X w_p) + 0.05 eps; Z ~ U(-1.2,1.2)^(q,d);
    ell_j = 0.5 sqrt(d), sf2 = 1, noise = 1e-2.  float64."""
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, generator=g, dtype=torch.float64) * 2.0 - 1.0
    W = torch.randn(d, p, generator=g, dtype=torch.float64)
    eps = torch.randn(n, p, generator=g, dtype=torch.float64)
    Y = torch.sin(X @ W) + 0.05 * eps
    Z = (torch.rand(q, d, generator=g, dtype=torch.float64) * 2.0 - 1.0) * 1.2
    ell = torch.full((d,), 0.5 * math.sqrt(d), dtype=torch.float64)
    return {"X": X, "Y": Y, "Z": Z, "ell": ell, "sf2": 1.0, "noise": 1e-2}


def rel_err(a, b) -> float:
    """Normwise relative error max|a-b| / max|b| used by every parity test."""
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    denom = float(b.abs().max())
    return float((a - b).abs().max()) / max(denom, 1e-300)
