"""MPC-side consumers of the GP hot path, same names as the reference's methods.

  ConstraintTightening.propagate_constraint_limits   gpmpc/gpmpc.py:425-498   (GPMPC.propagate_constraint_limits)
  precompute_sparse_posterior_mean                    gpmpc/gpmpc.py:377-400   (GPMPC.precompute_sparse_posterior_mean)
  precompute_posterior_mean                           gpmpc/gpmpc.py:370-375

Both stay on the device end to end: the tightening kernel reads the variance written by the prediction
kernels, the FITC weights are built from the same kernel-assembly / factorisation / GEMM entries as the exact
fit.  torch is device memory and a few elementwise scalings only; there is no CPU path.
"""

from __future__ import annotations

import numpy as np
import torch

from . import _lib
from . import engine as E
from ._lib import call, ptr

F64 = torch.float64


def stack_padded(gps: list, digits: int | None = None, dpad: int | None = None) -> tuple[E.GPBatch, int]:
    """One GPBatch from GPs with the same N but different input dimension (the reference's thrust / roll / pitch
    GPs have d = 1, 3, 3: gpmpc/gpmpc.py:59): inputs are padded with zero columns, which leaves every kernel
    value unchanged.  Elements are GaussianProcess objects or (X, y, ell, sf2, noise) tuples.  Returns
    (batch, padded d)."""
    raw = []
    for gp in gps:
        if isinstance(gp, tuple):
            X, y, ell, sf2, noise = gp
            ell = torch.as_tensor(ell, dtype=F64).reshape(-1)
        else:
            ell, sf2, noise = gp.hyperparameters()
            X, y = gp.train_inputs[0], gp.train_targets
        raw.append((X if X.dim() == 2 else X[:, None], y, ell, float(sf2), float(noise)))
    dmax = max(max(X.shape[1] for X, *_ in raw), dpad or 0)
    fits = []
    for X, y, ell, sf2, noise in raw:
        dev = _lib.require_cuda(X.device if X.is_cuda else None)
        Xp = torch.zeros(X.shape[0], dmax, dtype=F64, device=dev)
        Xp[:, : X.shape[1]] = X.to(dev, F64)
        ellp = torch.ones(dmax, dtype=F64)
        ellp[: X.shape[1]] = ell if ell.numel() == X.shape[1] else ell.expand(X.shape[1])
        fit, _ = E.factorize(Xp, y.to(dev), ellp, sf2, noise, digits=digits)
        fits.append(fit)
    return E.GPBatch.from_fits(fits), dmax


class ConstraintTightening:
    """Device-resident state of GPMPC.propagate_constraint_limits (gpmpc/gpmpc.py:425-498), batched over rollouts.

    gps: the trained GaussianProcess objects (thrust, roll, pitch); gp_idx: their input columns in z = (x, u)
    (gpmpc.py:59); discrete_dfdx, discrete_dfdu, lqr_gain: GPMPC.setup_prior_dynamics' outputs (gpmpc.py:84, 500-507);
    Bd (nx, 5) (gpmpc.py:68-69); inverse_cdf (gpmpc.py:63-65); dt the control period."""

    def __init__(self, gps: list, gp_idx: list, discrete_dfdx, discrete_dfdu, lqr_gain, Bd, dt: float,
                 inverse_cdf: float, idx_phi: int = 6, idx_theta: int = 7, horizon: int | None = None):
        assert len(gps) == 3 and len(gp_idx) == 3, "thrust, roll and pitch GPs"
        self.batch, self.dpad = stack_padded(gps)
        dev = self.batch.device
        self.gp_idx = [list(i) for i in gp_idx]
        self.A = torch.as_tensor(np.asarray(discrete_dfdx), dtype=F64).contiguous().to(dev)
        self.B = torch.as_tensor(np.asarray(discrete_dfdu), dtype=F64).contiguous().to(dev)
        self.K = torch.as_tensor(np.asarray(lqr_gain), dtype=F64).contiguous().to(dev)
        self.Bd = torch.as_tensor(np.asarray(Bd), dtype=F64).contiguous().to(dev)
        self.nx, self.nu = self.A.shape[0], self.B.shape[1]
        assert self.A.shape == (self.nx, self.nx) and self.K.shape == (self.nu, self.nx) and self.Bd.shape == (self.nx, 5)
        self.dt, self.inverse_cdf = float(dt), float(inverse_cdf)
        self.idx_phi, self.idx_theta = idx_phi, idx_theta
        self.T = horizon
        # gather map: column j of GP g's padded query comes from column cols[g][j] of z, or is zero
        cols = torch.zeros(3, self.dpad, dtype=torch.long)
        mask = torch.zeros(3, self.dpad, dtype=F64)
        for g, idx in enumerate(self.gp_idx):
            cols[g, : len(idx)] = torch.tensor(idx)
            mask[g, : len(idx)] = 1.0
        self._cols, self._mask = cols.to(dev), mask.to(dev)

    @torch.no_grad()
    def stage_variances(self, z: torch.Tensor) -> torch.Tensor:
        """Latent posterior variances (3, len(z)) of the three GPs at the stage inputs z (.., nx+nu) [gpmpc.py:441-445]."""
        Zg = z[:, self._cols] * self._mask          # (Q, 3, dpad)
        Zg = Zg.permute(1, 0, 2).contiguous()
        return E.predict(self.batch, Zg, want_mean_jac=False)[1]      # variance only: no Jacobian GEMM in the front kernel

    @torch.no_grad()
    def propagate_constraint_limits(self, x_prev, u_prev):
        """x_prev (nx, T+1) and u_prev (nu, T) as in the reference, or (R, nx, T+1) / (R, nu, T) for R rollouts.
        Returns (state_constraint (.., 2 nx, T+1), input_constraint (.., 2 nu, T)): numpy arrays for numpy input,
        device tensors for torch input.  No previous rollout (None) gives zeros [gpmpc.py:432-433]; the horizon is
        then the constructor's `horizon` or that of the last call."""
        dev = self.batch.device
        if x_prev is None or u_prev is None:
            assert self.T is not None, "no horizon known yet: pass horizon= to the constructor"
            return np.zeros((2 * self.nx, self.T + 1)), np.zeros((2 * self.nu, self.T))
        as_numpy = isinstance(x_prev, np.ndarray)
        x = torch.as_tensor(x_prev, dtype=F64).to(dev)
        u = torch.as_tensor(u_prev, dtype=F64).to(dev)
        single = x.dim() == 2
        if single:
            x, u = x[None], u[None]
        R, nx, T1 = x.shape
        T = self.T = T1 - 1
        assert nx == self.nx and u.shape == (R, self.nu, T), "x_prev (.., nx, T+1), u_prev (.., nu, T)"
        # z = hstack(x_prev[:, :-1].T, u_prev.T) per rollout [gpmpc.py:437]
        z = torch.cat([x[:, :, :-1].transpose(1, 2), u.transpose(1, 2)], dim=2).reshape(R * T, nx + self.nu).contiguous()
        sc = torch.empty(R, 2 * nx, T + 1, dtype=F64, device=dev)
        ic = torch.empty(R, 2 * self.nu, T, dtype=F64, device=dev)
        var = self.stage_variances(z) if T > 0 else torch.empty(3, 0, dtype=F64, device=dev)
        call("gpk_tighten_constraints", ptr(var), var.stride(0), 0, ptr(z), z.stride(0), self.idx_phi, self.idx_theta,
             ptr(self.batch.hyp), ptr(self.A), ptr(self.B), ptr(self.K), ptr(self.Bd), nx, self.nu, T, R, self.dt,
             self.inverse_cdf, ptr(sc), ptr(ic), _lib.stream_ptr(dev))
        if single:
            sc, ic = sc[0], ic[0]
        if as_numpy:
            return sc.cpu().numpy(), ic.cpu().numpy()
        return sc, ic


# --------------------------------------------------------------------------------------------------
# posterior-mean weights handed to the acados model (gpmpc/gpmpc.py:370-400)
# --------------------------------------------------------------------------------------------------
@torch.no_grad()
def precompute_posterior_mean(gps: list) -> tuple[np.ndarray, np.ndarray]:
    """alpha = K^-1 y per GP stacked (P, N) and the column-wise concatenated training inputs [gpmpc.py:370-375]."""
    alphas = [gp._prediction_fit().alpha[: gp.n_ind_points] for gp in gps]
    train = torch.cat([gp.train_inputs[0] for gp in gps], dim=-1)
    return torch.stack(alphas).cpu().numpy(), train.cpu().numpy()


def _fitc_weights(fit_x: E.GPFit, S: torch.Tensor) -> torch.Tensor:
    """FITC weights w (M,) of one GP for inducing inputs S (M, d) [gpmpc.py:391-398]:
         Gamma = diag(K - K_xs K_ss^-1 K_sx),  Sigma^-1 = K_ss + K_sx Gamma^-1 K_xs,  w = Sigma K_sx Gamma^-1 y.
    Gamma_i is sf2 + noise minus the posterior-variance reduction of a noise-free GP conditioned on S, evaluated
    at x_i -- the variance path of the exact GP with (S, noise 0) in place of (X, noise).
    K_ss is factorised by Cholesky with the exact fit's jitter policy; the reference LU-solves it without jitter
    (gpmpc.py:394), so the two agree to ~cond(K_ss) eps and the reference's result is rounding noise once that
    reaches the likelihood noise (e.g. many inducing points of a 1-D GP inside one lengthscale)."""
    dev, n, d = fit_x.device, fit_x.n, fit_x.d
    m = S.shape[0]
    mp = E.round_up(m, E.TILE)
    npad = fit_x.npad
    st = _lib.stream_ptr(dev)
    X, y = fit_x.X, fit_x.y
    # noise-free GP on the inducing points: W_s = chol(K_ss)^-1
    fit_s, _ = E.factorize(S, torch.zeros(m, dtype=F64, device=dev), fit_x.ell, fit_x.sf2, 0.0, center=fit_x.center,
                           want_image=False)
    _, var_s, _ = E.predict(fit_s, X, var_path="f64")           # sf2 - |W_s k_s(x_i)|^2
    gamma = var_s[0] + fit_x.noise                              # diag(K) = sf2 + noise
    g_isqrt = torch.rsqrt(gamma)
    # A = K_sx Gamma^-1/2 (M x N, zero padded), Sigma^-1 = K_ss + A A^T
    A = torch.empty(mp, npad, dtype=F64, device=dev)
    call("gpk_cross_kernel", 1, ptr(S), S.stride(0), m, 0, ptr(X), X.stride(0), n, 0, d, ptr(fit_s.hyp), ptr(A),
         npad, 0, mp, npad, 0, 1, st)
    A[:, :n] *= g_isqrt
    Sig = torch.empty(mp, mp, dtype=F64, device=dev)
    call("gpk_cross_kernel", 1, ptr(S), S.stride(0), m, 0, ptr(S), S.stride(0), m, 0, d, ptr(fit_s.hyp), ptr(Sig),
         mp, 0, mp, mp, 1, 1, st)                                # K_ss (noise 0), identity on the padding
    call("gpk_dgemm_nt", ptr(A), npad, ptr(A), npad, ptr(Sig), mp, mp, mp, npad, 1.0, 1.0, 0, 0, st)
    rhs = torch.zeros(mp, dtype=F64, device=dev)
    rhs[:m] = A[:m, :n] @ (g_isqrt * y)
    # solve Sigma^-1 w = rhs through the same factorisation entries as the exact fit
    ws = E.get_fit_scratch(mp, dev)
    ws.L.copy_(Sig)
    call("gpk_potrf", ptr(ws.L), mp, mp, ptr(ws.dinv), ptr(ws.info), st)
    if int(ws.info.item()) != 0:
        raise E.NotPSDError("FITC: K_ss + K_sx Gamma^-1 K_xs is not positive definite")
    Wt = torch.empty(mp, mp, dtype=F64, device=dev)
    call("gpk_trtri", ptr(ws.L), mp, mp, ptr(ws.dinv), ptr(Wt), mp, ptr(ws.U), mp, ptr(ws.T), ws.T.numel(), st)
    w = torch.empty(mp, dtype=F64, device=dev)
    call("gpk_solve_alpha", ptr(Wt), mp, ptr(ws.U), mp, ptr(ws.L), mp, mp, ptr(rhs), ptr(ws.t_work), ptr(w),
         ptr(ws.scal), st)
    return w[:m].clone()


@torch.no_grad()
def precompute_sparse_posterior_mean(gps: list, gp_idx: list, n_samples: int, np_random=None,
                                     rand_idx=None) -> tuple[np.ndarray, np.ndarray]:
    """FITC posterior-mean weights (P, n_samples) and the inducing inputs (n_samples, sum d) [gpmpc.py:377-400].
    The inducing points are `n_samples` training rows drawn without replacement by `np_random` exactly as the
    reference draws them (gpmpc.py:387), or the rows `rand_idx` when given."""
    inputs = torch.cat([gp.train_inputs[0] for gp in gps], dim=-1)
    if rand_idx is None:
        rand_idx = np_random.choice(range(inputs.shape[0]), size=n_samples, replace=False)
    sparse_inputs = inputs[torch.as_tensor(np.asarray(rand_idx), dtype=torch.long)]
    weights = []
    for gp, idx in zip(gps, gp_idx):
        fit = gp._prediction_fit()
        S = sparse_inputs[:, idx].to(fit.device, F64).contiguous()
        weights.append(_fitc_weights(fit, S))
    return torch.stack(weights).cpu().numpy(), sparse_inputs.cpu().numpy()


def sparse_gp(gp, S: torch.Tensor, w: torch.Tensor) -> E.GPFit:
    """Prediction state of the FITC mean K_f(z, S) w (the `sparse_gp_kernels_cs` expression of gpmpc.py:402-423
    times the precomputed weights): the exact GP's mean + Jacobian kernel with (S, w) in place of (X, alpha).
    Use with engine.predict(fit, Z, want_var=False)."""
    fx = gp._prediction_fit()
    dev = fx.device
    S = S.to(dev, F64).contiguous()
    m = S.shape[0]
    mp = E.round_up(m, E.TILE)
    alpha = torch.zeros(mp, dtype=F64, device=dev)
    alpha[:m] = torch.as_tensor(w, dtype=F64).to(dev)
    return E.GPFit(X=S, y=None, n=m, d=fx.d, npad=mp, center=fx.center, ell=fx.ell, sf2=fx.sf2, noise=fx.noise,
                   hyp=fx.hyp, W=None, alpha=alpha)


# --------------------------------------------------------------------------------------------------
# batched discrete stage dynamics of the acados model (gpmpc/gpmpc.py:166-209)
# --------------------------------------------------------------------------------------------------
PRIOR_KEYS = ("a", "b", "c", "d", "e", "f", "h", "l")     # scripts/gp_mpc_config.yaml:9-17


class StageDynamics:
    """f_disc and its Jacobians for batches of (x, u): the `disc_dyn_expr` GPMPC.setup_acados_model hands to acados
    (gpmpc/gpmpc.py:166-209), evaluated on the GPU for every stage of every rollout at once.

    gps: the thrust / roll / pitch GPs as GaussianProcess objects, (X, y, ell, sf2, noise) tuples, or engine.GPFit
    objects with d <= 3 (e.g. mpc.sparse_gp(...) for the FITC model, gpmpc.py:175-186); prior_params: dict with the
    keys a..l of the config (or a sequence a, b, c, d, e, f, h, l[, g, yaw c, d, e]); dt: the control period."""

    def __init__(self, gps: list, prior_params, dt: float, g: float = 9.81):
        assert len(gps) == 3, "thrust, roll and pitch GPs"
        if all(isinstance(f, E.GPFit) for f in gps):
            fits = [self._pad_fit(f) for f in gps]
        else:
            fits = stack_padded(gps, dpad=3)[0].fits
        assert all(f.n == fits[0].n for f in fits), "the three GPs share the number of (inducing) points"
        self.fits = fits
        dev = fits[0].device
        self.device = dev
        self.n = fits[0].n
        self.pack = torch.stack([f.pack(F64) for f in fits])
        self.hyp = torch.stack([f.hyp for f in fits])
        if isinstance(prior_params, dict):
            p = [float(prior_params[k]) for k in PRIOR_KEYS] + [float(g)]
            p += [float(prior_params.get(k, 0.0)) for k in ("yaw_c", "yaw_d", "yaw_e")]
        else:
            p = [float(v) for v in prior_params]
            if len(p) == 8:
                p.append(float(g))
        self.prior = torch.tensor(p, dtype=F64)
        self.dt = float(dt)

    @staticmethod
    def _pad_fit(f: E.GPFit) -> E.GPFit:
        if f.d == 3:
            return f
        assert f.d < 3
        X = torch.zeros(f.n, 3, dtype=F64, device=f.device)
        X[:, : f.d] = f.X
        ell = torch.ones(3, dtype=F64)
        ell[: f.d] = f.ell
        center = torch.zeros(3, dtype=F64)
        center[: f.d] = f.center
        return E.GPFit(X=X, y=f.y, n=f.n, d=3, npad=f.npad, center=center, ell=ell, sf2=f.sf2, noise=f.noise,
                       hyp=E.make_hyp(center, ell, f.sf2, f.noise, 3, f.device), W=None, alpha=f.alpha)

    @torch.no_grad()
    def disc_dyn(self, x, u, work_bytes: int = 2 << 30):
        """x (Q, 12), u (Q, 4) -> x_next (Q, 12), A (Q, 12, 12), B (Q, 12, 4); numpy in, numpy out; torch in, device
        tensors out."""
        as_numpy = isinstance(x, np.ndarray)
        dev = self.device
        x = torch.as_tensor(x, dtype=F64).to(dev).contiguous()
        u = torch.as_tensor(u, dtype=F64).to(dev).contiguous()
        single = x.dim() == 1
        if single:
            x, u = x[None], u[None]
        q = x.shape[0]
        assert x.shape == (q, 12) and u.shape == (q, 4), "x (Q, 12), u (Q, 4)"
        xn = torch.empty(q, 12, dtype=F64, device=dev)
        A = torch.empty(q, 12, 12, dtype=F64, device=dev)
        B = torch.empty(q, 12, 4, dtype=F64, device=dev)
        lib = _lib.load()
        need = lib.gpk_stage_rk4_work_bytes(max(q, 1), self.n, 3)
        wbytes = max(min(need, work_bytes), lib.gpk_stage_rk4_work_bytes(1, self.n, 3))
        work = E.get_work("stage_rk4", wbytes, dev)
        call("gpk_stage_rk4", ptr(x), x.stride(0), ptr(u), u.stride(0), q, self.prior.data_ptr(),
             self.prior.numel(), self.dt, ptr(self.hyp), ptr(self.pack), self.pack.stride(0),
             self.n, ptr(xn), ptr(A), ptr(B), ptr(work), wbytes, _lib.stream_ptr(dev))
        if single:
            xn, A, B = xn[0], A[0], B[0]
        if as_numpy:
            return xn.cpu().numpy(), A.cpu().numpy(), B.cpu().numpy()
        return xn, A, B
