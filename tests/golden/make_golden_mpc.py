"""Golden vectors for the MPC-side rows of SURVEY.md 8(f) -- produced by the REFERENCE'S OWN CODE.

`gpmpc/gpmpc.py` cannot be imported as is (casadi, acados_template, gpytorch are absent), but the two methods
these fixtures pin are plain numpy / torch:

  GPMPC.propagate_constraint_limits        gpmpc/gpmpc.py:425-498   -> tighten_*.npz
  GPMPC.precompute_sparse_posterior_mean   gpmpc/gpmpc.py:377-400   -> fitc_*.npz
  GPMPC.setup_prior_dynamics               gpmpc/gpmpc.py:500-507   (A, B, LQR gain of the tighten cases)

So this script loads that file from /root/reference with the absent third-party modules replaced by
`unittest.mock.MagicMock` stubs (they are only touched at import time and as the `fast_pred_var` context
managers) and calls the unmodified methods on a stand-in `self` whose GP objects return prescribed kernel
matrices / variances (computed by oracle/gp_oracle.py).  What is pinned is therefore the reference's
recursion and FITC algebra, given the GP quantities; the GP quantities themselves stay pinned as described
in oracle/gp_oracle.py.  /root/reference exists only in the build container: the fixtures are committed.

Run from the repo root:  python tests/golden/make_golden_mpc.py
"""
import importlib.util
import os
import sys
import types
from unittest.mock import MagicMock

import numpy as np
import scipy.stats
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import gp_oracle as o  # noqa: E402

REF = "/root/reference/gpmpc/gpmpc.py"


def load_reference():
    for m in ["casadi", "acados_template", "gpytorch", "gpytorch.settings", "gpmpc", "gpmpc.gp", "gpmpc.mpc"]:
        sys.modules.setdefault(m, MagicMock())
    spec = importlib.util.spec_from_file_location("_ref_gpmpc", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def hover_linearisation():
    """A plausible continuous-time linearisation of the 12-state quadrotor model about hover; state order
    (x, dx, y, dy, z, dz, phi, theta, psi, dphi, dtheta, dpsi), inputs (thrust, roll, pitch, yaw commands)."""
    nx, nu, g = 12, 4, 9.81
    A = np.zeros((nx, nx))
    B = np.zeros((nx, nu))
    for p in (0, 2, 4, 6, 7, 8):
        A[p, p + 1 if p < 6 else p + 3] = 1.0
    A[1, 7], A[3, 6] = g, -g
    B[5, 0] = 20.9
    for k, (a0, a1, b0) in enumerate([(-130.3, -16.33, 119.3), (-99.94, -13.3, 84.73), (0.0, -8.0, 6.0)]):
        A[9 + k, 6 + k], A[9 + k, 9 + k], B[9 + k, 1 + k] = a0, a1, b0
    return A, B


class FakeLikelihood:
    def __init__(self, noise):
        self.noise = torch.tensor([noise], dtype=torch.float64)

    def __call__(self, dist):
        return types.SimpleNamespace(covariance_matrix=dist.cov + self.noise * torch.eye(dist.cov.shape[0], dtype=torch.float64))


class FakeGP:
    """Stands in for gpmpc.gp.GaussianProcess: exact GP quantities from the oracle."""

    def __init__(self, X, y, ell, sf2, noise):
        self.X, self.y, self.ell, self.sf2, self.noise_v = X, y, ell, sf2, noise
        self.train_inputs, self.train_targets = (X,), y
        self.likelihood = FakeLikelihood(noise)
        self.st = o.fit_state(X, y, ell, sf2, noise)
        self.K = o.assemble_K(X, ell, sf2, noise)

    def eval(self):
        return self

    def covar_module(self, a, b=None):
        return types.SimpleNamespace(to_dense=lambda: o.se_kernel(a, a if b is None else b, self.ell, self.sf2))

    def __call__(self, z):
        _, var, _ = o.predict(z, self.X, self.ell, self.sf2, self.st["L"], self.st["alpha"])
        return types.SimpleNamespace(cov=torch.diag(var))


def make_gps(seed, n, gp_idx, ell0=0.35):
    g = torch.Generator().manual_seed(seed)
    X16 = torch.rand(n, 16, generator=g, dtype=torch.float64) * 0.8 - 0.4
    gps = []
    for k, idx in enumerate(gp_idx):
        X = X16[:, idx].contiguous()
        w = torch.randn(len(idx), generator=g, dtype=torch.float64)
        y = torch.sin(3.0 * X @ w) * 0.3 + 0.02 * torch.randn(n, generator=g, dtype=torch.float64)
        ell = torch.full((len(idx),), (ell0 if k == 0 else 0.35 + 0.1 * k), dtype=torch.float64)
        gps.append(FakeGP(X, y, ell, 0.4 + 0.2 * k, [2e-3, 5e-4, 1e-3][k]))
    return gps, X16


def tighten_case(mod, name, seed, T, n_train, rollouts, prob=0.955, dt=1.0 / 50.0):
    nx, nu = 12, 4
    gp_idx = [[0], [1, 2, 3], [4, 5, 6]]                                  # gpmpc.py:59
    gps, _ = make_gps(seed, n_train, gp_idx)
    A_c, B_c = hover_linearisation()
    Q = np.diag([8.0, 1.5, 8.0, 1.5, 8.0, 1.5, 0.5, 0.5, 0.5, 0.001, 0.001, 0.001])
    R = np.diag([3.0, 3.0, 3.0, 0.1])
    dfdx, dfdu, lqr_gain = mod.GPMPC.setup_prior_dynamics(A_c, B_c, Q, R, dt)
    inverse_cdf = scipy.stats.norm.ppf(1 - (1 / nx - (prob + 1) / (2 * nx)))   # gpmpc.py:63-65
    Bd = np.eye(nx)[:, [1, 3, 5, 9, 10]]                                   # gpmpc.py:68-69
    rng = np.random.default_rng(seed)
    xs, us, scs, ics, vars_ = [], [], [], [], []
    for r in range(rollouts):
        x_prev = rng.uniform(-0.45, 0.45, size=(nx, T + 1))
        u_prev = rng.uniform(-0.3, 0.5, size=(nu, T))
        fake = types.SimpleNamespace(
            model=types.SimpleNamespace(nx=nx, nu=nu), T=T, x_prev=x_prev, u_prev=u_prev, device="cpu",
            gaussian_process=gps, gp_idx=gp_idx, dt=dt, inverse_cdf=inverse_cdf, lqr_gain=lqr_gain,
            discrete_dfdx=dfdx, discrete_dfdu=dfdu, Bd=Bd)
        sc, ic = mod.GPMPC.propagate_constraint_limits(fake)
        z = torch.tensor(np.hstack((x_prev[:, :-1].T, u_prev.T)))
        vy = np.stack([(torch.diag(gp(z[:, idx]).cov) + gp.noise_v).numpy() for gp, idx in zip(gps, gp_idx)])
        xs.append(x_prev), us.append(u_prev), scs.append(sc), ics.append(ic), vars_.append(vy)
    np.savez(
        os.path.join(HERE, name + ".npz"), x_prev=np.stack(xs), u_prev=np.stack(us), state_constraint=np.stack(scs),
        input_constraint=np.stack(ics), var_y=np.stack(vars_), noise=np.array([gp.noise_v for gp in gps]),
        dfdx=dfdx, dfdu=dfdu, lqr_gain=lqr_gain, Bd=Bd, dt=np.float64(dt), inverse_cdf=np.float64(inverse_cdf),
        **{f"gp{k}_{f}": np.asarray(getattr(gp, a)) for k, gp in enumerate(gps)
           for f, a in (("X", "X"), ("y", "y"), ("ell", "ell"), ("sf2", "sf2"))})
    print(name, "T", T, "rollouts", rollouts, "max |state limit|", float(np.abs(np.stack(scs)).max()))


def fitc_case(mod, name, seed, n_train, n_samples, ell0):
    """ell0: lengthscale of the 1-D thrust GP, chosen so that K_ss of the drawn inducing points is well conditioned.
    With the tighten cases' 0.35, cond(K_ss) is 1e13 .. 1e18 and the reference's own output (an LU solve of K_ss
    without jitter, gpmpc.py:394) is dominated by rounding -- nothing another implementation could reproduce."""
    gp_idx = [[0], [1, 2, 3], [4, 5, 6]]
    gps, _ = make_gps(seed, n_train, gp_idx, ell0=ell0)
    # the reference concatenates the GPs' own inputs column-wise and re-splits with gp_idx (gpmpc.py:383, 392)
    fake = types.SimpleNamespace(gaussian_process=gps, gp_idx=gp_idx, np_random=np.random.default_rng(seed))
    w, sparse_inputs = mod.GPMPC.precompute_sparse_posterior_mean(fake, n_samples)
    for gp, idx in zip(gps, gp_idx):
        Kss = o.se_kernel(torch.tensor(sparse_inputs[:, idx]), torch.tensor(sparse_inputs[:, idx]), gp.ell, gp.sf2)
        print("   cond(K_ss) = %.3g" % np.linalg.cond(Kss.numpy()))
    np.savez(
        os.path.join(HERE, name + ".npz"), posterior_mean=w, sparse_inputs=sparse_inputs,
        noise=np.array([gp.noise_v for gp in gps]),
        **{f"gp{k}_{f}": np.asarray(getattr(gp, a)) for k, gp in enumerate(gps)
           for f, a in (("X", "X"), ("y", "y"), ("ell", "ell"), ("sf2", "sf2"))})
    print(name, "N", n_train, "M", n_samples, "max |w|", float(np.abs(w).max()))


def main():
    mod = load_reference()
    tighten_case(mod, "tighten_T25", seed=11, T=25, n_train=30, rollouts=4)     # the shipped horizon, config yaml
    tighten_case(mod, "tighten_T40", seed=12, T=40, n_train=45, rollouts=3, prob=0.9, dt=1.0 / 60.0)
    fitc_case(mod, "fitc_n45_m10", seed=21, n_train=45, n_samples=10, ell0=0.05)
    fitc_case(mod, "fitc_n200_m25", seed=22, n_train=200, n_samples=25, ell0=0.012)


if __name__ == "__main__":
    main()
