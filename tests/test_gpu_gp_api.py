"""GPU parity tests of the reference-facing Python surface (GaussianProcess / fit_gp / eval-mode
prediction) against the CPU oracle and the committed golden fixtures.  Run with -m gpu."""
import os

import numpy as np
import pytest
import torch

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL64, TOL32 = 1e-6, 1e-4


def _gp_at(X, y, ell, sf2, noise, ard=True):
    from gpmpc_b200.gp import GaussianProcess

    gp = GaussianProcess(X, y, ard=ard)
    gp.set_hyperparameters(ell, sf2, noise)
    # compare at the float32-rounded values the GP actually holds (parameters are float32)
    return gp, gp.hyperparameters()


@pytest.mark.parametrize("name", sorted(f[:-4] for f in os.listdir(GOLD) if f.endswith(".npz") and f.startswith(("ref_", "syn_"))))
def test_golden_fixtures(name):
    from gpmpc_b200 import engine as E

    g = np.load(os.path.join(GOLD, name + ".npz"))
    X, y, Z = (torch.from_numpy(g[k]) for k in ("X", "y", "Z"))
    ell, sf2, noise = torch.from_numpy(g["ell"]), float(g["sf2"]), float(g["noise"])
    dev = torch.device("cuda", 0)
    fit, ws = E.factorize(X.to(dev), y.to(dev), ell, sf2, noise, want_kinv=True)
    assert o.rel_err(fit.alpha[: fit.n].cpu(), g["alpha"]) < TOL64
    assert abs(fit.mll_per_n() - float(g["mll_per_n"])) < TOL64 * abs(float(g["mll_per_n"]))
    for dtype, tol in ((torch.float64, TOL64), (torch.float32, TOL32)):
        mean, var, jac = E.predict(fit, Z.to(dev).to(dtype))
        assert o.rel_err(mean[0].cpu(), g["mean"]) < tol, (name, dtype)
        assert o.rel_err(var[0].cpu(), g["var"]) < tol, (name, dtype)
        assert o.rel_err(jac[0].cpu(), g["jac"]) < tol, (name, dtype)
    loss, g_noise, g_sf2, g_ell = E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
    assert abs(loss - float(g["loss_raw"])) < TOL64 * abs(float(g["loss_raw"]))
    rl = torch.from_numpy(g["raw_ell"])
    gl = torch.tensor(g_ell, dtype=torch.float64)
    gl = (gl.reshape(1, -1) if rl.shape[1] > 1 else gl.sum().reshape(1, 1)) * torch.sigmoid(rl)
    assert o.rel_err(gl, g["g_raw_ell"]) < TOL64
    gn = g_noise * float(torch.sigmoid(torch.from_numpy(g["raw_noise"])))
    assert abs(gn - float(g["g_raw_noise"].reshape(-1)[0])) < TOL64 * max(abs(float(g["g_raw_noise"].reshape(-1)[0])), 1e-3)


@pytest.mark.parametrize("n,d", [(15, 1), (30, 3), (45, 3)])
def test_fit_gp_follows_reference_adam_loop(n, d):
    """fit_gp (gpmpc/gp.py:49-69): same losses and final raw parameters as the oracle's restatement of the
    loop; then gp.K / gp.K_inv as compute_covariances (gp.py:43-46)."""
    from gpmpc_b200.gp import GaussianProcess, fit_gp

    p = o.synthetic_problem(n, d, 1, seed=50 + n)
    X, y = p["X"], p["Y"][:, 0]
    gp = GaussianProcess(X, y)
    fit_gp(gp, n_train=60, lr=0.05, device="cpu")
    ref = o.fit_hyperparameters(X, y, n_train=60, lr=0.05)
    assert len(gp.fit_losses) == len(ref["losses"])
    assert o.rel_err(torch.tensor(gp.fit_losses), torch.tensor(ref["losses"])) < 1e-5
    assert torch.allclose(gp.likelihood.noise_covar.raw_noise.detach().cpu(), ref["raw_noise"], atol=2e-4)
    assert torch.allclose(gp.covar_module.raw_outputscale.detach().cpu(), ref["raw_sf2"], atol=2e-4)
    assert torch.allclose(gp.covar_module.base_kernel.raw_lengthscale.detach().cpu(), ref["raw_ell"], atol=2e-4)
    ell, sf2, noise = gp.hyperparameters()
    K_ref = o.assemble_K(X, ell, sf2, noise)
    assert gp.K.shape == (n, n) and gp.K.device.type == "cpu" and gp.K.dtype == torch.float64
    assert o.rel_err(gp.K, K_ref) < 1e-12
    assert o.rel_err(gp.K_inv, torch.linalg.inv(K_ref)) < TOL64
    assert len(gp.fit_losses) < 60 or gp.fit_losses[-1] < gp.fit_losses[0]


def test_fit_gp_early_stop_and_ard():
    from gpmpc_b200.gp import GaussianProcess, fit_gp

    p = o.synthetic_problem(40, 3, 1, seed=77)
    X, y = p["X"], p["Y"][:, 0]
    gp = GaussianProcess(X, y, ard=True)
    fit_gp(gp, n_train=400, lr=0.05, device="cuda")
    ref = o.fit_hyperparameters(X, y, n_train=400, lr=0.05, ard=True)
    assert len(gp.fit_losses) == len(ref["losses"]) < 400       # |delta loss| < 1e-3 stops both at the same step
    assert gp.K.is_cuda and next(gp.parameters()).is_cuda


@pytest.mark.parametrize("n,d,q", [(15, 1, 25), (45, 3, 25), (300, 16, 25)])
def test_eval_mode_distribution_like_gpmpc_call_site(n, d, q):
    """gp.eval(); cov = gp.likelihood(gp(z)).covariance_matrix; diag(cov)   (gpmpc/gpmpc.py:441-445)."""
    p = o.synthetic_problem(n, d, q, seed=9 + n)
    X, y, Z = p["X"], p["Y"][:, 0], p["Z"]
    gp, (ell, sf2, noise) = _gp_at(X, y, p["ell"] * 0.9, 1.4, 0.03)
    gp.eval()
    dist = gp(Z)
    cov = gp.likelihood(dist).covariance_matrix
    st = o.fit_state(X, y, ell, sf2, noise)
    m_ref, v_ref, j_ref = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    kzx = o.se_kernel(Z, X, ell, sf2)
    v = torch.linalg.solve_triangular(st["L"], kzx.T, upper=False)
    cov_ref = o.se_kernel(Z, Z, ell, sf2) - v.T @ v + noise * torch.eye(q, dtype=torch.float64)
    assert cov.shape == (q, q) and cov.device.type == "cpu"
    assert o.rel_err(cov, cov_ref) < TOL64
    assert o.rel_err(torch.diag(cov), v_ref + noise) < TOL64
    assert o.rel_err(dist.mean, m_ref) < TOL64
    assert o.rel_err(gp.likelihood(dist).variance, v_ref + noise) < TOL64
    assert o.rel_err(dist.jacobian, j_ref) < TOL64
    mean, var, jac = gp.predict(Z, include_noise=True)
    assert o.rel_err(var, v_ref + noise) < TOL64 and o.rel_err(jac, j_ref) < TOL64
    # train() drops the cached prediction state, like gpytorch
    gp.train()
    assert gp._pred_fit is None
    prior = gp(Z)
    assert float(prior.mean.abs().max()) == 0.0
    assert o.rel_err(prior.covariance_matrix, o.se_kernel(Z, Z, ell, sf2)) < 1e-12


def test_kernel_evaluator_and_fitc_flow():
    """gp.covar_module(a[,b]).to_dense(), .add_diag(noise) and the FITC weights of gpmpc/gpmpc.py:391-398."""
    p = o.synthetic_problem(45, 3, 1, seed=3)
    X, y = p["X"], p["Y"][:, 0]
    gp, (ell, sf2, noise) = _gp_at(X, y, torch.full((3,), 0.8), 1.1, 0.02, ard=False)
    gp.K, gp.K_inv = gp.compute_covariances()
    S = X[torch.randperm(45, generator=torch.Generator().manual_seed(0))[:40]]
    K_ss = gp.covar_module(S).to_dense()
    K_xs = gp.covar_module(X, S).to_dense()
    assert o.rel_err(K_ss, o.se_kernel(S, S, ell, sf2)) < 1e-12 and o.rel_err(K_xs, o.se_kernel(X, S, ell, sf2)) < 1e-12
    Kn = gp.covar_module(X).add_diag(gp.likelihood.noise).to_dense()
    assert o.rel_err(Kn, gp.K) < 1e-13
    Gamma = torch.diagonal(gp.K - K_xs @ torch.linalg.solve(K_ss, K_xs.T))
    Gamma_inv = torch.diag_embed(1 / Gamma)
    Sigma_inv = K_ss + K_xs.T @ Gamma_inv @ K_xs
    w = torch.linalg.solve(Sigma_inv, K_xs.T) @ Gamma_inv @ y
    assert o.rel_err(w, o.fitc_posterior_mean(X, y, S, ell, sf2, noise)) < 1e-6


def test_known_answer_cases():
    from gpmpc_b200 import engine as E

    dev = torch.device("cuda", 0)
    sf2, noise, ell = 1.7, 0.3, torch.tensor([0.8, 1.1], dtype=torch.float64)
    X = torch.tensor([[0.2, -0.4]], dtype=torch.float64)
    y = torch.tensor([1.3], dtype=torch.float64)
    Z = torch.tensor([[0.5, 0.1], [0.2, -0.4], [60.0, -70.0]], dtype=torch.float64)
    fit, _ = E.factorize(X.to(dev), y.to(dev), ell, sf2, noise)
    mean, var, jac = (t[0].cpu() for t in E.predict(fit, Z.to(dev)))
    k = o.se_kernel(Z, X, ell, sf2)[:, 0]
    assert torch.allclose(mean, k / (sf2 + noise) * 1.3, rtol=1e-12, atol=1e-300)
    assert torch.allclose(var, sf2 - k * k / (sf2 + noise), rtol=1e-12)
    assert float(jac[1].abs().max()) < 1e-15                     # at the data point
    assert float(mean[2]) == 0.0 and float(var[2]) == sf2 and float(jac[2].abs().max()) == 0.0   # far away
    # query == training point, tiny noise: mean -> y_i, var_f -> ~0; duplicated rows survive
    p = o.synthetic_problem(20, 3, 1, seed=2)
    Xd = torch.cat([p["X"], p["X"][:3]])
    yd = torch.cat([p["Y"][:, 0], p["Y"][:3, 0]])
    fit, _ = E.factorize(Xd.to(dev), yd.to(dev), p["ell"], 1.0, 1e-6)
    mean, var, _ = (t[0].cpu() for t in E.predict(fit, Xd[:5].to(dev)))
    assert o.rel_err(mean, yd[:5]) < 1e-3 and float(var.abs().max()) < 1e-5


def test_full_size_properties_config2():
    """BASELINE config 2 size (P=3, N=2048, d=16, Q=8192): size-independent properties instead of an oracle
    pass: linearity of the mean/Jacobian in y, 0 <= var_f <= sf2, var independent of y, batched == single."""
    from gpmpc_b200 import engine as E

    dev = torch.device("cuda", 0)
    p = o.synthetic_problem(2048, 16, 8192, p=2, seed=1337)
    X, Y, Z = p["X"].to(dev), p["Y"].to(dev), p["Z"].to(dev)
    Y3 = torch.stack([Y[:, 0], Y[:, 1], 2.0 * Y[:, 0] - 3.0 * Y[:, 1]], dim=1)
    gb = E.fit_batch(X, Y3, p["ell"], p["sf2"], p["noise"])
    mean, var, jac = E.predict(gb, Z)
    assert o.rel_err(mean[2].cpu(), (2.0 * mean[0] - 3.0 * mean[1]).cpu()) < 1e-9
    assert o.rel_err(jac[2].cpu(), (2.0 * jac[0] - 3.0 * jac[1]).cpu()) < 1e-9
    assert torch.equal(var[0], var[1]) and torch.equal(var[0], var[2])
    assert float(var.min()) > 0.0 and float(var.max()) <= p["sf2"]
    m1, v1, j1 = E.predict(gb.fits[1], Z)
    # the training-set split over CTAs depends on the batch size, so the summation order differs
    # (also for the variance: the number of W row pieces per query tile depends on the batch size)
    assert o.rel_err(m1[0], mean[1]) < 1e-12 and o.rel_err(j1[0], jac[1]) < 1e-12 and o.rel_err(v1[0], var[1]) < 1e-12
    # a sample of the queries against the oracle
    idx = torch.arange(0, 8192, 64)
    st = o.fit_state(p["X"], p["Y"][:, 0], p["ell"], p["sf2"], p["noise"])
    m_ref, v_ref, j_ref = o.predict(p["Z"][idx], p["X"], p["ell"], p["sf2"], st["L"], st["alpha"])
    assert o.rel_err(mean[0][idx].cpu(), m_ref) < TOL64
    assert o.rel_err(var[0][idx].cpu(), v_ref) < TOL64
    assert o.rel_err(jac[0][idx].cpu(), j_ref) < TOL64
    # float32 path, same sample
    m32, v32, j32 = E.predict(gb, Z.float())
    assert o.rel_err(m32[0][idx].cpu(), m_ref) < TOL32
    assert o.rel_err(v32[0][idx].cpu(), v_ref) < TOL32
    assert o.rel_err(j32[0][idx].cpu(), j_ref) < TOL32


def test_full_size_properties_config3():
    """BASELINE config 3 training size (N=16384, d=16; the fit runs its large GEMMs on the int8 tensor pipe, the
    inverse factor on the side stream), checked through exact identities that need no O(N^3) oracle pass:
      mean at the training inputs:  K_f alpha = y - noise * alpha          (from (K_f + noise I) alpha = y)
      variance at a training input: var_f(x_i) = noise - noise^2 (K^-1)_ii  (from K_f K^-1 = I - noise K^-1)
      y^T K^-1 y = y . alpha,  W W^T-free residual:  || L (L^T alpha) - y || small
      Jacobian == central finite difference of the mean;  int8 variance path == DMMA variance path."""
    from gpmpc_b200 import engine as E

    dev = torch.device("cuda", 0)
    n, d = 16384, 16
    p = o.synthetic_problem(n, d, 512, p=1, seed=1337)
    X, y, Z = p["X"].to(dev), p["Y"][:, 0].to(dev), p["Z"].to(dev)
    noise, sf2 = p["noise"], p["sf2"]
    fit, ws = E.factorize(X, y, p["ell"], sf2, noise, want_kinv=True)
    assert ws.gemm_work is not None, "the int8 fit GEMMs must be active at this size"
    alpha = fit.alpha[:n]
    # residual of the linear system through the factor itself: K alpha = L (L^T alpha)
    Lm = torch.tril(ws.L[:n, :n])
    res = Lm @ (Lm.T @ alpha) - y
    # 6-digit int8 GEMMs: 5.8e-8 measured (|K| ~ 4e3, |alpha| ~ 10: 1.5e-12 of |K| |alpha|); DMMA only: ~1e-11
    assert float(res.abs().max()) < TOL64 * float(y.abs().max())
    assert abs(fit.quad - float(y @ alpha)) < 1e-9 * abs(fit.quad)
    assert abs(fit.logdet - 2.0 * float(torch.log(Lm.diagonal()).sum())) < 1e-10 * abs(fit.logdet)
    # predictions at 1024 training inputs
    idx = torch.arange(0, n, 16, device=dev)
    mean, var, jac = E.predict(fit, X[idx])
    assert o.rel_err(mean[0].cpu(), (y[idx] - noise * alpha[idx]).cpu()) < TOL64
    kinv_diag = ws.Kinv.diagonal()[:n][idx]
    assert o.rel_err(var[0].cpu(), (noise - noise * noise * kinv_diag).cpu()) < TOL64
    assert float(var.min()) > 0.0 and float(var.max()) <= noise
    # away from the data: both variance paths, Jacobian against finite differences of the mean
    m1, v1, j1 = E.predict(fit, Z)
    m2, v2, j2 = E.predict(fit, Z, var_path="f64")
    assert o.rel_err(v1[0].cpu(), v2[0].cpu()) < TOL64 and o.rel_err(m1[0].cpu(), m2[0].cpu()) < 1e-10
    assert float(v1.min()) > 0.0 and float(v1.max()) <= sf2
    h = 1e-5
    for j in (0, 7, 15):
        e = torch.zeros(d, dtype=torch.float64, device=dev)
        e[j] = h
        mp = E.predict(fit, Z + e, want_var=False)[0]
        mm = E.predict(fit, Z - e, want_var=False)[0]
        fd = (mp - mm)[0] / (2 * h)
        assert o.rel_err(j1[0][:, j].cpu(), fd.cpu()) < 1e-5
    E.release_scratch()


def test_full_size_properties_config4():
    """BASELINE config 4 training size (N=32768, d=16): the fit (6-digit int8 potrf / trtri, 5-digit K^-1, inverse factor
    on the side stream) and BOTH variance paths -- the tcgen05 int8 path is exact up to gpk_oz_max_npad(5) = 32768+ --
    through the same exact identities as config 3 (no O(N^3) oracle pass)."""
    from gpmpc_b200 import engine as E

    dev = torch.device("cuda", 0)
    n, d = 32768, 16
    assert E.i8_supported(n, 5) and not E.i8_supported(n + 256, 5) and not E.i8_supported(n, 6)
    p = o.synthetic_problem(n, d, 256, p=1, seed=1337)
    X, y, Z = p["X"].to(dev), p["Y"][:, 0].to(dev), p["Z"].to(dev)
    noise, sf2 = p["noise"], p["sf2"]
    fit, ws = E.factorize(X, y, p["ell"], sf2, noise, want_kinv=True, digits=5)
    assert ws.gemm_work is not None and fit.wimg is not None and fit.digits == 5
    alpha = fit.alpha[:n]
    # residual of the linear system through the factor itself, in row blocks (no second N x N temporary)
    Lm = ws.L[:n, :n]
    t = torch.zeros(n, dtype=torch.float64, device=dev)
    res = torch.empty(n, dtype=torch.float64, device=dev)
    blk = 4096
    for r0 in range(0, n, blk):          # t = L^T alpha
        t[: r0 + blk] += torch.tril(Lm[r0 : r0 + blk, : r0 + blk], diagonal=r0).T @ alpha[r0 : r0 + blk]
    for r0 in range(0, n, blk):          # res = L t - y
        res[r0 : r0 + blk] = torch.tril(Lm[r0 : r0 + blk, : r0 + blk], diagonal=r0) @ t[: r0 + blk] - y[r0 : r0 + blk]
    assert float(res.abs().max()) < TOL64 * float(y.abs().max())
    assert abs(fit.quad - float(y @ alpha)) < 1e-9 * abs(fit.quad)
    assert abs(fit.logdet - 2.0 * float(torch.log(Lm.diagonal()).sum())) < 1e-10 * abs(fit.logdet)
    # predictions at 512 training inputs: mean and variance identities, int8 variance path
    idx = torch.arange(0, n, 64, device=dev)
    mean, var, jac = E.predict(fit, X[idx])
    assert o.rel_err(mean[0].cpu(), (y[idx] - noise * alpha[idx]).cpu()) < TOL64
    kinv_diag = ws.Kinv.diagonal()[:n][idx]
    assert o.rel_err(var[0].cpu(), (noise - noise * noise * kinv_diag).cpu()) < TOL64
    assert float(var.min()) > 0.0 and float(var.max()) <= noise
    # away from the data: tcgen05 int8 path == DMMA path
    m1, v1, j1 = E.predict(fit, Z)
    m2, v2, j2 = E.predict(fit, Z, var_path="f64")
    assert o.rel_err(v1[0].cpu(), v2[0].cpu()) < TOL64 and o.rel_err(m1[0].cpu(), m2[0].cpu()) < 1e-10
    assert o.rel_err(j1[0].cpu(), j2[0].cpu()) < 1e-10
    assert float(v1.min()) > 0.0 and float(v1.max()) <= sf2
    del fit, ws, Lm
    E.release_scratch()
    torch.cuda.empty_cache()


def test_predict_host_roundtrip():
    from gpmpc_b200 import engine as E
    from gpmpc_b200.gp import predict_host

    dev = torch.device("cuda", 0)
    p = o.synthetic_problem(256, 4, 300, p=2, seed=8)
    gb = E.fit_batch(p["X"].to(dev), p["Y"].to(dev), p["ell"], 1.0, 1e-2)
    out = predict_host(gb, p["Z"].pin_memory())
    ref = E.predict(gb, p["Z"].to(dev))
    assert all(not t.is_cuda and t.is_pinned() for t in out)
    assert all(torch.equal(a, b.cpu()) for a, b in zip(out, ref))


@pytest.mark.parametrize("n,d,q,B", [(700, 16, 333, 1), (2048, 16, 1500, 3), (130, 9, 5, 2), (64, 32, 129, 1)])
def test_mean_jacobian_only_dmma_path(n, d, q, B):
    """want_var=False in float64 for d > 8 runs on the DMMA front kernel without the digit image; same numbers as the
    SIMT kernel and the oracle, also with a scratch that forces several query chunks."""
    from gpmpc_b200 import _lib
    from gpmpc_b200 import engine as E
    from gpmpc_b200._lib import call, ptr
    p = o.synthetic_problem(n, d, q, p=B, seed=n + d)
    gb = E.fit_batch(p["X"].cuda(), p["Y"].cuda(), p["ell"], p["sf2"], p["noise"])
    Z = p["Z"].cuda()
    m1, v1, j1 = E.predict(gb, Z, want_var=False)                       # DMMA
    m2, v2, j2 = E.predict(gb, Z, want_var=False, mj_path="simt")
    assert v1 is None and v2 is None
    assert o.rel_err(m1.cpu(), m2.cpu()) < 1e-12 and o.rel_err(j1.cpu(), j2.cpu()) < 1e-12
    for b in range(B):
        st = o.fit_state(p["X"], p["Y"][:, b], p["ell"], p["sf2"], p["noise"])
        mr, _, jr = o.predict(p["Z"], p["X"], p["ell"], p["sf2"], st["L"], st["alpha"])
        assert o.rel_err(m1[b].cpu(), mr) < 1e-9 and o.rel_err(j1[b].cpu(), jr) < 1e-9
    # straight through the C ABI with room for 128 queries per chunk only
    lib = _lib.load()
    wb = lib.gpk_mean_jac_dmma_work_bytes(128, d, B)
    work = torch.empty(wb, dtype=torch.uint8, device="cuda")
    mean = torch.full((B, q), float("nan"), dtype=torch.float64, device="cuda")
    jac = torch.full((B, q, d), float("nan"), dtype=torch.float64, device="cuda")
    alpha = gb.alpha
    call("gpk_predict_mean_jac_dmma", 1, ptr(Z), Z.stride(0), q, 0, d, ptr(gb.hyp), ptr(gb.X), gb.X.stride(1), n,
         gb.X.stride(0) if B > 1 else 0, ptr(alpha), alpha.stride(0), ptr(mean), q, ptr(jac), d, q * d, ptr(work), wb, B,
         torch.cuda.current_stream().cuda_stream)
    # (the k-split count follows the chunk size, so the summation order differs: equal to rounding, not bitwise)
    assert o.rel_err(mean.cpu(), m1.cpu()) < 1e-13 and o.rel_err(jac.cpu(), j1.cpu()) < 1e-13
    assert not torch.isnan(mean).any() and not torch.isnan(jac).any()


@pytest.mark.parametrize("n,d,q", [(300, 5, 77), (1000, 16, 400)])
def test_variance_paths_agree(n, d, q, monkeypatch):
    """The three variance paths -- tcgen05 int8 digits (default), FP64 DMMA (`var_path="f64"`), and the automatic
    fallback to DMMA when the training set is too large for exact int32 accumulation -- against the oracle."""
    from gpmpc_b200 import engine as E
    p = o.synthetic_problem(n, d, q, p=1, seed=n)
    X, y, Z = p["X"].cuda(), p["Y"][:, 0].cuda(), p["Z"].cuda()
    st = o.fit_state(p["X"], p["Y"][:, 0], p["ell"], p["sf2"], p["noise"])
    _, vr, _ = o.predict(p["Z"], p["X"], p["ell"], p["sf2"], st["L"], st["alpha"])
    fit, _ = E.factorize(X, y, p["ell"], p["sf2"], p["noise"])
    v_i8 = E.predict(fit, Z)[1][0].cpu()
    v_f64 = E.predict(fit, Z, var_path="f64")[1][0].cpu()
    assert o.rel_err(v_i8, vr) < 1e-7 and o.rel_err(v_f64, vr) < 1e-9
    m0, v0, j0 = E.predict(fit, Z, want_mean_jac=False)          # variance alone: the front kernel skips the Jacobian GEMM
    assert m0 is None and j0 is None and torch.equal(v0[0].cpu(), v_i8)
    assert torch.equal(E.predict(fit, Z, want_mean_jac=False, var_path="f64")[1][0].cpu(), v_f64)
    for S in (4, 6):
        fs, _ = E.factorize(X, y, p["ell"], p["sf2"], p["noise"], digits=S)
        assert o.rel_err(E.predict(fs, Z)[1][0].cpu(), vr) < {4: 1e-5, 6: 1e-9}[S]
    # a fit without digit image at a size the int8 path cannot take: predict() switches to the DMMA path by itself
    monkeypatch.setattr(E, "i8_supported", lambda npad, digits=None: False)
    fit2, _ = E.factorize(X, y, p["ell"], p["sf2"], p["noise"])
    assert fit2.wimg is None
    m2, v2, j2 = E.predict(fit2, Z)
    assert torch.equal(v2[0].cpu(), v_f64)
    monkeypatch.undo()
    assert E.i8_supported(32768) and not E.i8_supported(33024) and E.i8_supported(43904, 4) and not E.i8_supported(26496, 6)


@pytest.mark.parametrize("n", [600, 1500])
def test_graphed_mll_iteration_is_bit_identical_and_fit_gp_uses_it(n):
    """engine.GraphedMLL: the MLL + gradient iteration replayed as a CUDA graph gives the same numbers as the stream
    launches, survives changing hyper-parameters and a non-PD request; fit_gp follows the oracle loop with it."""
    from gpmpc_b200 import engine as E
    from gpmpc_b200.gp import GaussianProcess, fit_gp

    dev = torch.device("cuda", 0)
    p = o.synthetic_problem(n, 5, 1, p=1, seed=n)
    X, y = p["X"].to(dev), p["Y"][:, 0].to(dev)
    gm = E.GraphedMLL(X, y)
    for k, (ell, sf2, noise) in enumerate([(1.0, 1.0, 1e-2), (0.8, 1.3, 3e-2), (1.4, 0.7, 5e-3), (0.9, 1.1, 2e-2)]):
        ell_t = torch.full((5,), ell, dtype=torch.float64)
        got = gm(ell_t, sf2, noise)
        fit, ws = E.factorize(X, y, ell_t, sf2, noise, center=gm.center, want_kinv=True, want_image=False)
        ref = E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
        assert (gm.graph is not None) == (k >= E.GRAPH_FIT_WARMUP)
        assert got[0] == ref[0] and got[1] == ref[1] and got[2] == ref[2] and got[3] == ref[3], k
    with pytest.raises(E.NotPSDError):
        gm(torch.full((5,), 1.0, dtype=torch.float64), 1.0, -0.5)
    Xs, ys = p["X"][:520], p["Y"][:520, 0]
    gp = GaussianProcess(Xs, ys)
    fit_gp(gp, n_train=8, lr=0.05, device="cuda:0", compute_covariances=False)
    ref = o.fit_hyperparameters(Xs, ys, n_train=8, lr=0.05)
    assert len(gp.fit_losses) == len(ref["losses"])
    assert o.rel_err(torch.tensor(gp.fit_losses), torch.tensor(ref["losses"])) < 1e-5


def test_fit_gps_batched_equals_three_fit_gp_calls():
    """fit_gps: the reference's three GPs (d = 1, 3, 3, same N) fitted in lockstep, one launch per Adam iteration for all
    of them, with per-GP early stop -- same losses, parameters, K and K^-1 as three fit_gp calls and as the oracle loop."""
    from gpmpc_b200.gp import GaussianProcess, fit_gp, fit_gps

    g = torch.Generator().manual_seed(4)
    N = 45
    X = torch.rand(N, 7, generator=g, dtype=torch.float64) - 0.5
    Y = 0.3 * torch.sin(3 * X[:, :3]) + 0.01 * torch.randn(N, 3, generator=g, dtype=torch.float64)
    idx = ([0], [1, 2, 3], [4, 5, 6])
    mk = lambda: [GaussianProcess(X[:, i].contiguous(), Y[:, p].contiguous()) for p, i in enumerate(idx)]
    a, b = mk(), mk()
    fit_gps(a, n_train=40, lr=0.05, device="cpu")
    for gp in b:
        fit_gp(gp, n_train=40, lr=0.05, device="cpu")
    for p, (ga, gb) in enumerate(zip(a, b)):
        ref = o.fit_hyperparameters(X[:, idx[p]].contiguous(), Y[:, p].contiguous(), n_train=40, lr=0.05)
        assert len(ga.fit_losses) == len(gb.fit_losses) == len(ref["losses"])
        assert o.rel_err(torch.tensor(ga.fit_losses), torch.tensor(gb.fit_losses)) < 1e-9
        assert o.rel_err(torch.tensor(ga.fit_losses), torch.tensor(ref["losses"])) < 1e-5
        for pa, pb in zip(ga.parameters(), gb.parameters()):
            assert torch.allclose(pa, pb, atol=1e-6)
        assert ga.K.shape == (N, N) and o.rel_err(ga.K, gb.K) < 1e-9 and o.rel_err(ga.K_inv, gb.K_inv) < 1e-6
    # different N per GP: falls back to one fit_gp per GP
    c = [GaussianProcess(X[:30, [0]].contiguous(), Y[:30, 0].contiguous()), GaussianProcess(X[:, [1, 2, 3]].contiguous(), Y[:, 1].contiguous())]
    fit_gps(c, n_train=3, lr=0.05)
    assert all(len(gp.fit_losses) >= 1 and gp.K is not None for gp in c)
