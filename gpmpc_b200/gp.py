"""Drop-in for the reference GP layer `gpmpc/gp.py`, backed by the sm_100a CUDA path (no gpytorch).

Same public names, signatures and attribute surface as the reference (file:line under /root/reference):
  GaussianProcess(x, y)                 gpmpc/gp.py:24-46   (gpytorch ExactGP subclass there)
  fit_gp(gp, n_train, lr, device)       gpmpc/gp.py:49-69
  gpytorch_predict2casadi(gp)           gpmpc/gp.py:72-85   (needs casadi; import-guarded)
  covSE_single / covSE_vectorized       gpmpc/gp.py:12-21   (CasADi symbolic; import-guarded)
and everything `gpmpc/gpmpc.py` touches on a GP object: train_inputs, train_targets, K,
K_inv, n_ind_points, input_dimension, train()/eval()/to()/parameters(), gp(x) -> distribution,
gp.likelihood(dist), .covariance_matrix, gp.likelihood.noise, gp.covar_module(a[,b]).to_dense(),
.add_diag(), covar_module.base_kernel.lengthscale, covar_module.outputscale, compute_covariances().

New (no reference counterpart): `GaussianProcess.predict(Z)` -> mean, variance, Jacobian in one batched
call, and `predict_batch` for several same-shape GPs.  All arithmetic runs in libgpmpc_b200.so on the GPU;
without a CUDA device every compute entry raises (there is no CPU fallback).
"""

from __future__ import annotations

import math

import torch
from torch import nn

from . import engine as E

try:  # casadi is an optional dependency of the symbolic exports only
    import casadi as cs
except ImportError:  # pragma: no cover - casadi is absent in the build image
    cs = None

NOISE_LOWER_BOUND = 1e-6  # GreaterThan(1e-6), gpmpc/gp.py:31
USE_FIT_GRAPH = True      # fit_gp iterations at N >= engine.GRAPH_FIT_MIN_N replay a CUDA graph (engine.GraphedMLL)
SMALL_FIT_MAX_N = 128     # fit_gp iterations of GPs up to this size run as ONE fused kernel (engine.small_loss_and_grad)


def _need_casadi():
    if cs is None:
        raise ImportError("casadi is required for the symbolic kernel / prediction exports (pip install casadi)")


# --------------------------------------------------------------------------------------------------
# CasADi symbolic kernels (gpmpc/gp.py:12-21)
# --------------------------------------------------------------------------------------------------
def covSE_single(x, z, ell, sf2):
    """sf2 * exp(-1/2 sum_j (x_j - z_j)^2 / ell_j^2), CasADi symbolic; x (d,), z (d,N)."""
    _need_casadi()
    r2 = cs.sum1((x - z) ** 2 / ell**2)      # same operation order as the reference: bit-equal outputs
    return sf2 * cs.exp(-0.5 * r2)


def covSE_vectorized(x, Z, ell, sf2):
    """Same kernel with the second argument given row-major, Z (M,d)."""
    _need_casadi()
    m = Z.shape[0]
    r2 = cs.sum1((cs.repmat(x, 1, m) - Z.T) ** 2 / ell**2)
    return sf2 * cs.exp(-0.5 * r2)


# --------------------------------------------------------------------------------------------------
# minimal stand-ins for the gpytorch objects the controller touches
# --------------------------------------------------------------------------------------------------
def _softplus(x):
    return torch.nn.functional.softplus(x)


class _NoiseCovar(nn.Module):
    def __init__(self):
        super().__init__()
        self.raw_noise = nn.Parameter(torch.zeros(1))

    @property
    def noise(self) -> torch.Tensor:
        return _softplus(self.raw_noise) + NOISE_LOWER_BOUND


class GaussianLikelihood(nn.Module):
    """Homoskedastic Gaussian noise, noise = softplus(raw) + 1e-6 (gpmpc/gp.py:31)."""

    def __init__(self):
        super().__init__()
        self.noise_covar = _NoiseCovar()

    @property
    def noise(self) -> torch.Tensor:
        return self.noise_covar.noise

    def forward(self, dist: "MultivariateNormal") -> "MultivariateNormal":
        return dist._with_added_diag(float(self.noise.detach().double()))


class RBFKernel(nn.Module):
    """lengthscale = softplus(raw); shape (1,1) as in the reference (no ard_num_dims, gpmpc/gp.py:34) or
    (1,d) with ard=True."""

    def __init__(self, ard_num_dims: int | None = None):
        super().__init__()
        self.ard_num_dims = ard_num_dims
        self.raw_lengthscale = nn.Parameter(torch.zeros(1, ard_num_dims or 1))

    @property
    def lengthscale(self) -> torch.Tensor:
        return _softplus(self.raw_lengthscale)


class LazyKernelMatrix:
    """What gp.covar_module(a[, b]) returns: densified on the GPU by .to_dense()."""

    def __init__(self, gp: "GaussianProcess", a: torch.Tensor, b: torch.Tensor | None, diag: float = 0.0):
        self._gp, self._a, self._b, self._diag = gp, a, b, diag

    def add_diag(self, value) -> "LazyKernelMatrix":
        v = float(torch.as_tensor(value).detach().double().reshape(-1)[0])
        return LazyKernelMatrix(self._gp, self._a, self._b, self._diag + v)

    def to_dense(self) -> torch.Tensor:
        K = E.cross_K(self._gp._hyper_fit(), self._a, self._b)
        if self._diag:
            K = K.clone()
            K.diagonal().add_(self._diag)
        return K.to(self._a.device) if not self._a.is_cuda else K

    evaluate = to_dense

    @property
    def shape(self):
        b = self._a if self._b is None else self._b
        return torch.Size([self._a.shape[0], b.shape[0]])


class ScaleKernel(nn.Module):
    def __init__(self, base_kernel: RBFKernel):
        super().__init__()
        self.base_kernel = base_kernel
        self.raw_outputscale = nn.Parameter(torch.zeros(()))
        self._gp = None  # set by GaussianProcess (not a submodule: avoid a reference cycle in state_dict)

    @property
    def outputscale(self) -> torch.Tensor:
        return _softplus(self.raw_outputscale)

    def forward(self, a: torch.Tensor, b: torch.Tensor | None = None) -> LazyKernelMatrix:
        return LazyKernelMatrix(self._gp[0], a, b)


class MultivariateNormal:
    """Distribution returned by gp(x).  Only what the controller reads is provided: mean, variance,
    covariance_matrix (dense, for modest Q as at gpmpc/gpmpc.py:444), and the batched extras."""

    MAX_DENSE_Q = 8192

    def __init__(self, gp: "GaussianProcess", x: torch.Tensor, posterior: bool, diag: float = 0.0):
        self._gp, self._x, self._posterior, self._diag = gp, x, posterior, diag
        self._cache = None

    def _with_added_diag(self, v: float) -> "MultivariateNormal":
        out = MultivariateNormal(self._gp, self._x, self._posterior, self._diag + v)
        out._cache = self._cache
        return out

    def _pred(self):
        if self._cache is None:
            self._cache = E.predict(self._gp._prediction_fit(), self._x)
        return self._cache

    def _out(self, t: torch.Tensor) -> torch.Tensor:
        return t if self._x.is_cuda else t.to(self._x.device)

    @property
    def mean(self) -> torch.Tensor:
        if not self._posterior:
            return torch.zeros(self._x.shape[0], dtype=self._x.dtype, device=self._x.device)
        return self._out(self._pred()[0][0])

    loc = mean

    @property
    def variance(self) -> torch.Tensor:
        if not self._posterior:
            sf2 = float(self._gp.covar_module.outputscale.detach().double())
            return torch.full((self._x.shape[0],), sf2 + self._diag, dtype=self._x.dtype, device=self._x.device)
        return self._out(self._pred()[1][0] + self._diag)

    @property
    def jacobian(self) -> torch.Tensor:
        """d mean / d x, (Q, d) -- what CasADi AD gives acados for the exported mean (gpmpc/gp.py:85)."""
        return self._out(self._pred()[2][0])

    @property
    def covariance_matrix(self) -> torch.Tensor:
        q = self._x.shape[0]
        if not self._posterior:
            return LazyKernelMatrix(self._gp, self._x, None, self._diag).to_dense()
        if q > self.MAX_DENSE_Q:
            raise ValueError(f"dense {q}x{q} posterior covariance refused; use .variance (the diagonal the "
                             "controller keeps at gpmpc/gpmpc.py:445)")
        cov = E.posterior_covariance(self._gp._prediction_fit(), self._x)
        if self._diag:
            cov.diagonal().add_(self._diag)
        return self._out(cov)


# --------------------------------------------------------------------------------------------------
# the GP itself
# --------------------------------------------------------------------------------------------------
class GaussianProcess(nn.Module):
    """Exact GP, zero mean, ScaleKernel(RBFKernel()), Gaussian likelihood with noise >= 1e-6.

    Mirrors gpmpc/gp.py:24-46.  Parameters are float32 like gpytorch's defaults while the data (and all
    linear algebra) is float64.  `ard=True` is an extension (one lengthscale per input)."""

    def __init__(self, x: torch.Tensor, y: torch.Tensor, ard: bool = False):
        assert isinstance(x, torch.Tensor), "x must be a torch.Tensor"
        assert isinstance(y, torch.Tensor), "y must be a torch.Tensor"
        super().__init__()
        if x.dim() == 1:
            x = x[:, None]
        self.train_inputs = (x,)
        self.train_targets = y
        self.likelihood = GaussianLikelihood()
        self.covar_module = ScaleKernel(RBFKernel(x.shape[1] if ard else None))
        self.covar_module._gp = (self,)
        # Save dimensions for later use (gpmpc/gp.py:36-37)
        self.n_ind_points = x.shape[0]
        self.input_dimension = x.shape[1]
        self.K, self.K_inv = None, None  # only computed once the GP is trained (gpmpc/gp.py:38)
        self._pred_fit = None            # cached prediction state, dropped by train() like gpytorch's
        self._hyp_fit = None

    # ---- hyper-parameter access -----------------------------------------------------------------
    def hyperparameters(self) -> tuple[torch.Tensor, float, float]:
        """(ell (d,) float64, sf2, noise): constrained values, computed in the parameter dtype then promoted."""
        with torch.no_grad():
            ell = self.covar_module.base_kernel.lengthscale.double().reshape(-1).cpu()
            if ell.numel() == 1:
                ell = ell.expand(self.input_dimension).clone()
            sf2 = float(self.covar_module.outputscale.double())
            noise = float(self.likelihood.noise.double())
        return ell, sf2, noise

    def set_hyperparameters(self, ell, sf2: float, noise: float):
        """Set constrained values directly (used by tests/benchmarks that compare at fixed hyper-parameters)."""
        def inv_softplus(v):
            v = torch.as_tensor(v, dtype=torch.float64)
            return v + torch.log(-torch.expm1(-v))

        with torch.no_grad():
            rl = self.covar_module.base_kernel.raw_lengthscale
            ell_t = torch.as_tensor(ell, dtype=torch.float64).reshape(1, -1)
            if ell_t.shape[1] != rl.shape[1]:
                assert rl.shape[1] == 1 and bool((ell_t == ell_t[0, 0]).all()), "non-ARD GP takes one lengthscale"
                ell_t = ell_t[:, :1]
            rl.copy_(inv_softplus(ell_t).to(rl.dtype))
            self.covar_module.raw_outputscale.copy_(inv_softplus(sf2).to(rl.dtype))
            self.likelihood.noise_covar.raw_noise.copy_(inv_softplus(noise - NOISE_LOWER_BOUND).to(rl.dtype).reshape(1))
        self._invalidate()

    def _invalidate(self):
        self._pred_fit = None
        self._hyp_fit = None

    def train(self, mode: bool = True):
        if mode:
            self._invalidate()   # gpytorch drops its prediction strategy on train()
        return super().train(mode)

    # ---- device-side state ----------------------------------------------------------------------
    def _compute_device(self) -> torch.device:
        x = self.train_inputs[0]
        return E._lib.require_cuda(x.device if x.is_cuda else None)

    def _prediction_fit(self) -> E.GPFit:
        """Cholesky state at the current hyper-parameters (gpytorch: DefaultPredictionStrategy caches)."""
        if self._pred_fit is None:
            ell, sf2, noise = self.hyperparameters()
            dev = self._compute_device()
            self._pred_fit, _ = E.factorize(self.train_inputs[0].to(dev), self.train_targets.to(dev), ell, sf2, noise)
        return self._pred_fit

    def _hyper_fit(self) -> E.GPFit:
        """Hyper-parameter block only (kernel evaluation without factorising)."""
        if self._pred_fit is not None:
            return self._pred_fit
        if self._hyp_fit is None:
            ell, sf2, noise = self.hyperparameters()
            dev = self._compute_device()
            X = self.train_inputs[0].detach().to(dev, torch.float64).contiguous()
            center = X.mean(dim=0).cpu()
            self._hyp_fit = E.GPFit(X=X, y=None, n=X.shape[0], d=X.shape[1], npad=E.round_up(X.shape[0], E.TILE),
                                    center=center, ell=ell, sf2=sf2, noise=noise,
                                    hyp=E.make_hyp(center, ell, sf2, noise, X.shape[1], dev), W=None, alpha=None)
        return self._hyp_fit

    # ---- reference surface ----------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> MultivariateNormal:
        """Prior over x: zero mean, K_f(x, x) (gpmpc/gp.py:40-41)."""
        return MultivariateNormal(self, x, posterior=False)

    def __call__(self, x: torch.Tensor) -> MultivariateNormal:
        if self.training:
            return self.forward(x)
        return MultivariateNormal(self, x, posterior=True)

    def compute_covariances(self) -> tuple[torch.Tensor, torch.Tensor]:
        """K(X,X) + sigma^2 I and its inverse (gpmpc/gp.py:43-46), both dense (N,N) float64."""
        ell, sf2, noise = self.hyperparameters()
        dev = self._compute_device()
        x = self.train_inputs[0]
        if self.n_ind_points <= SMALL_FIT_MAX_N:
            # the reference's sizes: both matrices from the one-launch small-GP kernel (no jitter was needed if it returns)
            cov = {}
            *_, jitter = E.small_loss_and_grad(x.to(dev), self.train_targets.to(dev), ell, sf2, noise, covariances=cov)
            if jitter == 0.0:
                K, K_inv = cov["K"], cov["K_inv"]
                return (K, K_inv) if x.is_cuda else (K.to(x.device), K_inv.to(x.device))
        fit, ws = E.factorize(self.train_inputs[0].to(dev), self.train_targets.to(dev), ell, sf2, noise, want_kinv=True)
        self._pred_fit = fit
        n = fit.n
        K = E.dense_K(fit).clone()
        K_inv = ws.Kinv[:n, :n].clone()
        x = self.train_inputs[0]
        if not x.is_cuda:
            K, K_inv = K.to(x.device), K_inv.to(x.device)
        return K, K_inv

    # ---- batched prediction (new) ---------------------------------------------------------------
    @torch.no_grad()
    def predict(self, Z: torch.Tensor, include_noise: bool = False):
        """mean (Q,), variance (Q,), Jacobian (Q,d) at Z (Q,d) in Z's dtype (float64 or float32).
        variance is the latent var_f; include_noise=True adds sigma^2 like gp.likelihood(gp(z))."""
        mean, var, jac = E.predict(self._prediction_fit(), Z)
        var = var[0]
        if include_noise:
            var = var + self.hyperparameters()[2]
        if not Z.is_cuda:
            return mean[0].to(Z.device), var.to(Z.device), jac[0].to(Z.device)
        return mean[0], var, jac[0]


@torch.no_grad()
def predict_batch(gps: list, Z: torch.Tensor, want_var: bool = True):
    """One-launch prediction for several GPs of identical (N, d): mean (P,Q), var_f (P,Q), jac (P,Q,d)."""
    batch = E.GPBatch.from_fits([gp._prediction_fit() for gp in gps])
    return E.predict(batch, Z, want_var=want_var)


# --------------------------------------------------------------------------------------------------
# hyper-parameter fit (gpmpc/gp.py:49-69)
# --------------------------------------------------------------------------------------------------
def _loss_and_set_grads(gp: GaussianProcess, dev: torch.device, W_buf: torch.Tensor | None, grad_reduce=None):
    """One MLL evaluation on the GPU: loss = -MLL/N, closed-form gradient written to the raw parameters'
    .grad (replaces output = gp(train_x); loss = -mll(output, y); loss.backward(), gpmpc/gp.py:61-63)."""
    ell, sf2, noise = gp.hyperparameters()
    X = gp.train_inputs[0]
    cache = getattr(gp, "_fit_cache", None)
    if cache is None or cache[0] is not X or cache[4] is not gp.train_targets:
        # training data on the compute device and its centre (read back once per fit, not once per iteration)
        Xd = X.detach().to(dev, torch.float64).contiguous()
        cache = gp._fit_cache = (X, Xd, gp.train_targets.detach().to(dev, torch.float64).contiguous(), Xd.mean(dim=0).cpu(),
                                 gp.train_targets)
    fit = None
    if grad_reduce is None and gp.n_ind_points <= SMALL_FIT_MAX_N:
        # the reference's own sizes (N <= 45): the whole iteration is one launch and one host round trip
        loss, g_noise, g_sf2, g_ell, _ = E.small_loss_and_grad(cache[1], cache[2], ell, sf2, noise, center=cache[3])
    elif grad_reduce is None and gp.n_ind_points >= E.GRAPH_FIT_MIN_N and USE_FIT_GRAPH:
        # mid / large N: the iteration's ~170+ launches on two streams replayed as one CUDA graph
        graphed = getattr(gp, "_fit_graph", None)
        if graphed is None or graphed.key is not cache[1]:
            graphed = gp._fit_graph = E.GraphedMLL(cache[1], cache[2], center=cache[3])
        loss, g_noise, g_sf2, g_ell, _ = graphed(ell, sf2, noise)
    else:
        # a row-split reducer (sharding.RowSplitGradReduce) names the rows of K^-1 this rank needs; everyone else gets all
        want_kinv = grad_reduce.rows(gp.n_ind_points) if hasattr(grad_reduce, "rows") else True
        fit, ws = E.factorize(cache[1], cache[2], ell, sf2, noise, center=cache[3], want_kinv=want_kinv, W_out=W_buf,
                              want_image=False)
        if grad_reduce is None:
            partials = E.grad_partials(fit, ws)
        else:
            partials = grad_reduce(fit, ws)    # sharding layer: row-split + all-reduce
        loss, g_noise, g_sf2, g_ell = E.loss_and_grad_from_partials(fit, partials)
    rn = gp.likelihood.noise_covar.raw_noise
    rs = gp.covar_module.raw_outputscale
    rl = gp.covar_module.base_kernel.raw_lengthscale
    with torch.no_grad():
        g_ell_t = torch.tensor(g_ell, dtype=torch.float64)
        if rl.shape[1] == 1:
            g_ell_t = g_ell_t.sum().reshape(1, 1)
        else:
            g_ell_t = g_ell_t.reshape(1, -1)
        rn.grad = (g_noise * torch.sigmoid(rn.detach().double())).to(rn.dtype).reshape(rn.shape).to(rn.device)
        rs.grad = (g_sf2 * torch.sigmoid(rs.detach().double())).to(rs.dtype).reshape(rs.shape).to(rs.device)
        rl.grad = (g_ell_t.to(rl.device) * torch.sigmoid(rl.detach().double())).to(rl.dtype)
    return loss, fit


def fit_gp(gp: GaussianProcess, n_train=500, lr=0.01, device: str = "cpu", *, compute_covariances: bool = True,
           grad_reduce=None):
    """Fit a GP to its training data: Adam on -MLL/N, early stop at |delta loss| < 1e-3, then cache
    gp.K and gp.K_inv.  Same signature and stopping rule as gpmpc/gp.py:49-69.

    `device` keeps its reference meaning for where the GP's tensors live afterwards; the arithmetic always
    runs on a CUDA device (the one named, else the current one) -- a missing GPU raises, there is no CPU path.
    """
    assert isinstance(gp, GaussianProcess), f"gp must be a GaussianProcess, got {type(gp)}"
    dev = E._lib.require_cuda(device if "cuda" in str(device) else None)
    gp.train_inputs = (gp.train_inputs[0].to(device),)
    gp.train_targets = gp.train_targets.to(device)
    gp.train().to(device)
    optim = torch.optim.Adam(gp.parameters(), lr=lr)
    npad = E.round_up(gp.n_ind_points, E.TILE)
    small = grad_reduce is None and (gp.n_ind_points <= SMALL_FIT_MAX_N or
                                     (USE_FIT_GRAPH and gp.n_ind_points >= E.GRAPH_FIT_MIN_N))
    W_buf = None if small else torch.empty(npad, npad, dtype=torch.float64, device=dev)
    last_loss = math.inf
    gp.fit_losses = []
    for _ in range(n_train):
        optim.zero_grad()
        loss, _ = _loss_and_set_grads(gp, dev, W_buf, grad_reduce)
        optim.step()
        gp._invalidate()
        gp.fit_losses.append(loss)
        if abs(last_loss - loss) < 1e-3:  # early stopping if converged (gpmpc/gp.py:65)
            break
        last_loss = loss
    del W_buf
    gp._fit_cache = None
    gp._fit_graph = None
    if compute_covariances:
        gp.K, gp.K_inv = gp.compute_covariances()


def _set_raw_grads(gp: GaussianProcess, g_noise: float, g_sf2: float, g_ell: list):
    """Chain rule through softplus onto the raw parameters' .grad (closed form instead of loss.backward(), gp.py:63)."""
    rn = gp.likelihood.noise_covar.raw_noise
    rs = gp.covar_module.raw_outputscale
    rl = gp.covar_module.base_kernel.raw_lengthscale
    with torch.no_grad():
        g_ell_t = torch.tensor(g_ell, dtype=torch.float64)
        g_ell_t = g_ell_t.sum().reshape(1, 1) if rl.shape[1] == 1 else g_ell_t.reshape(1, -1)
        rn.grad = (g_noise * torch.sigmoid(rn.detach().double())).to(rn.dtype).reshape(rn.shape).to(rn.device)
        rs.grad = (g_sf2 * torch.sigmoid(rs.detach().double())).to(rs.dtype).reshape(rs.shape).to(rs.device)
        rl.grad = (g_ell_t.to(rl.device) * torch.sigmoid(rl.detach().double())).to(rl.dtype)


def fit_gps(gps: list, n_train=500, lr=0.01, device: str = "cpu", *, compute_covariances: bool = True):
    """fit_gp for several small GPs AT ONCE (no reference counterpart: gpmpc/gpmpc.py:159-162 calls fit_gp for its three
    GPs one after the other).  Every GP keeps its own Adam state and its own early stop (|delta loss| < 1e-3, gp.py:65), so
    the result equals three fit_gp calls; per iteration ALL GPs are evaluated by one launch and one host round trip
    (engine.small_batch_loss_and_grad).  Needs the same number of training points N <= 128 in every GP (the reference
    draws the same samples for all three); anything else falls back to fit_gp per GP."""
    assert all(isinstance(gp, GaussianProcess) for gp in gps), "gps must be GaussianProcess objects"
    n = gps[0].n_ind_points
    if len(gps) < 2 or n > SMALL_FIT_MAX_N or any(gp.n_ind_points != n for gp in gps):
        for gp in gps:
            fit_gp(gp, n_train=n_train, lr=lr, device=device, compute_covariances=compute_covariances)
        return
    dev = E._lib.require_cuda(device if "cuda" in str(device) else None)
    B, dp = len(gps), max(gp.input_dimension for gp in gps)
    Xpad = torch.zeros(B, n, dp, dtype=torch.float64, device=dev)
    Y = torch.empty(B, n, dtype=torch.float64, device=dev)
    for b, gp in enumerate(gps):
        gp.train_inputs = (gp.train_inputs[0].to(device),)
        gp.train_targets = gp.train_targets.to(device)
        gp.train().to(device)
        Xpad[b, :, : gp.input_dimension] = gp.train_inputs[0].to(dev, torch.float64)
        Y[b] = gp.train_targets.to(dev, torch.float64)
        gp.fit_losses = []
    centers = Xpad.mean(dim=1).cpu()
    optims = [torch.optim.Adam(gp.parameters(), lr=lr) for gp in gps]
    last = [math.inf] * B
    active = [True] * B

    def evaluate(cov=None):
        hp = [gp.hyperparameters() for gp in gps]
        res = E.small_batch_loss_and_grad(Xpad, Y, centers, [h[0] for h in hp], [h[1] for h in hp], [h[2] for h in hp],
                                          covariances=cov)
        for b, r in enumerate(res):
            if r[4] != 0:       # not positive definite: this GP alone, with the jitter retries
                ell, sf2, noise = hp[b]
                res[b] = (*E.small_loss_and_grad(Xpad[b, :, : gps[b].input_dimension].contiguous(), Y[b], ell, sf2, noise)[:4], -1)
        return res

    for _ in range(n_train):
        if not any(active):
            break
        res = evaluate()
        for b, gp in enumerate(gps):
            if not active[b]:
                continue
            loss, g_noise, g_sf2, g_ell, _ = res[b]
            optims[b].zero_grad()
            _set_raw_grads(gp, g_noise, g_sf2, g_ell)
            optims[b].step()
            gp._invalidate()
            gp.fit_losses.append(loss)
            if abs(last[b] - loss) < 1e-3:          # early stopping if converged (gpmpc/gp.py:65)
                active[b] = False
            last[b] = loss
    if compute_covariances:
        cov = {}
        res = evaluate(cov)
        for b, gp in enumerate(gps):
            if res[b][4] == 0:
                x = gp.train_inputs[0]
                K, K_inv = cov["K"][b].clone(), cov["K_inv"][b].clone()
                gp.K, gp.K_inv = (K, K_inv) if x.is_cuda else (K.to(x.device), K_inv.to(x.device))
            else:
                gp.K, gp.K_inv = gp.compute_covariances()


# --------------------------------------------------------------------------------------------------
# exports of the fitted mean (gpmpc/gp.py:72-85)
# --------------------------------------------------------------------------------------------------
def gpytorch_predict2casadi(gp: GaussianProcess):
    """CasADi Function "pred": z -> mean = k(z, X) K^-1 y with X, y, ell, sf2, K^-1 baked in as constants
    (gpmpc/gp.py:72-85).  alpha = K^-1 y is folded on the GPU so the expression is O(N) per evaluation."""
    assert isinstance(gp, GaussianProcess), f"Expected a GaussianProcess, got {type(gp)}"
    _need_casadi()
    X = gp.train_inputs[0].numpy(force=True)
    assert X.ndim == 2, "train_inputs must be a 2D array"
    ell = gp.covar_module.base_kernel.lengthscale.to_dense().numpy(force=True)
    sf2 = gp.covar_module.outputscale.to_dense().numpy(force=True)
    fit = gp._prediction_fit()
    alpha = fit.alpha[: fit.n].cpu().numpy()
    z = cs.SX.sym("z", X.shape[1])
    k = covSE_single(z, X.T, ell.T, sf2)
    return cs.Function("pred", [z], [k @ alpha], ["z"], ["mean"])


def predict_callable(gp: GaussianProcess):
    """CasADi-free stand-in for the exported "pred" function: z (d,) or (Q,d) array-like ->
    dict(mean=..., jac=...) evaluated by the fused CUDA kernel (batched over Q)."""
    def pred(z):
        zt = torch.as_tensor(z, dtype=torch.float64)
        single = zt.dim() == 1
        zt = zt.reshape(1, -1) if single else zt
        mean, _, jac = E.predict(gp._prediction_fit(), zt.to(gp._compute_device()), want_var=False)
        mean, jac = mean[0].cpu().numpy(), jac[0].cpu().numpy()
        return {"mean": mean[0] if single else mean, "jac": jac[0] if single else jac}

    return pred


def gp_casadi_callback(gp: GaussianProcess, name: str = "pred_gpu"):
    """CasADi bridge: a `casadi.Callback` z (d,) -> mean (1,) whose value AND Jacobian come from
    the fused CUDA mean + Jacobian kernel, so that a CasADi graph (an NLP solved with ipopt / sqpmethod, or
    `cs.Function`s evaluated in Python) can use the GPU posterior in place of the symbolic `pred` of
    gpytorch_predict2casadi (gpmpc/gp.py:72-85, called at gpmpc/gpmpc.py:189-191).  Needs casadi.

    Scope note: acados generates C from the symbolic expression graph of `AcadosModel.disc_dyn_expr`; a Callback has
    no C source, so the acados OCP of `setup_acados_model` keeps using the symbolic export (or gets its stage
    linearisations from mpc.StageDynamics directly).  The Callback serves CasADi-side consumers."""
    _need_casadi()
    d = gp.input_dimension
    pred = predict_callable(gp)

    class _GPJacobian(cs.Callback):
        """Jacobian function in CasADi's convention: inputs (z, mean), output d mean / d z (1, d)."""

        def __init__(self, cname):
            cs.Callback.__init__(self)
            self.construct(cname, {})

        def get_n_in(self): return 2
        def get_n_out(self): return 1
        def get_sparsity_in(self, i): return cs.Sparsity.dense(d, 1) if i == 0 else cs.Sparsity.dense(1, 1)
        def get_sparsity_out(self, i): return cs.Sparsity.dense(1, d)

        def eval(self, arg):
            import numpy as np
            z = np.asarray(arg[0], dtype=np.float64).reshape(-1)
            return [pred(z)["jac"].reshape(1, d)]

    class _GPMean(cs.Callback):
        def __init__(self, cname):
            cs.Callback.__init__(self)
            self._jac = None
            self.construct(cname, {"enable_fd": False})

        def get_n_in(self): return 1
        def get_n_out(self): return 1
        def get_name_in(self, i): return "z"
        def get_name_out(self, i): return "mean"
        def get_sparsity_in(self, i): return cs.Sparsity.dense(d, 1)
        def get_sparsity_out(self, i): return cs.Sparsity.dense(1, 1)

        def eval(self, arg):
            import numpy as np
            z = np.asarray(arg[0], dtype=np.float64).reshape(-1)
            return [np.asarray(pred(z)["mean"]).reshape(1, 1)]

        def has_jacobian(self): return True

        def get_jacobian(self, jname, inames, onames, opts):
            self._jac = _GPJacobian(jname)       # kept alive by the owning callback, as CasADi requires
            return self._jac

    return _GPMean(name)


_copy_streams: dict = {}


@torch.no_grad()
def predict_host(batch, Z_host: torch.Tensor, out: tuple | None = None):
    """End-to-end prediction from HOST buffers: copies the (pinned) queries to the GPU, runs the batched
    mean + variance + Jacobian path and returns (mean, var, jac) in pinned host memory; blocks until the
    results are on the host.  `batch` is an engine.GPBatch / GPFit or a list of fitted GaussianProcess.
    The mean and Jacobian (94 % of the result bytes) are copied out on a second stream while the variance
    contraction still runs."""
    if isinstance(batch, (list, tuple)):
        batch = E.GPBatch.from_fits([gp._prediction_fit() for gp in batch])
    dev = batch.device
    main = torch.cuda.current_stream(dev)
    side = _copy_streams.get(dev.index)
    if side is None:
        side = _copy_streams[dev.index] = torch.cuda.Stream(dev)
    Zd = Z_host.to(dev, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(main)                      # creates the CUDA event; re-recorded by the library when mean/jac are final
    mean, var, jac = E.predict(batch, Zd, mj_event=ev)
    if out is None:
        out = tuple(torch.empty(r.shape, dtype=r.dtype, pin_memory=True) for r in (mean, var, jac))
    with torch.cuda.stream(side):
        side.wait_event(ev)
        out[0].copy_(mean, non_blocking=True)
        out[2].copy_(jac, non_blocking=True)
    out[1].copy_(var, non_blocking=True)
    mean.record_stream(side)
    jac.record_stream(side)
    main.synchronize()
    side.synchronize()
    return out
