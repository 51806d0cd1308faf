"""GPU parity against outputs of the REFERENCE'S OWN CODE (tests/golden/make_golden_covse.py ran the unmodified
gpmpc/gp.py:12-21 covSE_single / covSE_vectorized and gpmpc/gp.py:72-85 gpytorch_predict2casadi over a NumPy casadi
shim).  Rows a1, a2, a5 (K), a6 of SURVEY.md 8(a).  Run with -m gpu."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
COVSE = [f"covse_d{d}_{k}" for d in (1, 3, 16) for k in ("shared", "ard")]
DEV = torch.device("cuda", 0)


def _ell_vec(g, d):
    ell = torch.from_numpy(g["ell"]).reshape(-1)
    return ell.expand(d).clone() if ell.numel() == 1 else ell


@pytest.mark.parametrize("name", COVSE)
def test_cross_kernel_matches_reference_covse(name):
    """gpk_cross_kernel (the kernel every CUDA path shares) against covSE_single / covSE_vectorized, element-wise."""
    from gpmpc_b200 import engine as E

    g = np.load(os.path.join(GOLD, name + ".npz"))
    x, Z = torch.from_numpy(g["x"])[None].to(DEV), torch.from_numpy(g["Z"]).to(DEV)
    d, sf2 = Z.shape[1], float(g["sf2"])
    ell = _ell_vec(g, d)
    center = Z.mean(dim=0).cpu()                      # the kernels centre their inputs (translation invariant)
    fit = E.GPFit(X=Z, y=None, n=Z.shape[0], d=d, npad=E.round_up(Z.shape[0], E.TILE), center=center, ell=ell, sf2=sf2,
                  noise=0.0, hyp=E.make_hyp(center, ell, sf2, 0.0, d, DEV), W=None, alpha=None)
    k = E.cross_K(fit, x, Z).cpu().numpy()
    for ref in (g["k_single"], g["k_vec"]):
        assert k.shape == ref.shape
        assert (np.abs(k - ref) / ref).max() < 5e-13, name       # element-wise relative (no cancellation in the kernel)


@pytest.mark.parametrize("name", ["pred_d3_n45", "pred_d16_n300"])
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-6), (torch.float32, 1e-4)])
def test_fused_mean_matches_reference_casadi_mean(name, dtype, tol):
    """The fused CUDA mean (every prediction path) against the reference's exported CasADi "pred" function."""
    from gpmpc_b200 import engine as E

    g = np.load(os.path.join(GOLD, name + ".npz"))
    X, y, Z = (torch.from_numpy(g[k]).to(DEV) for k in ("X", "y", "Z"))
    ell, sf2, noise = _ell_vec(g, X.shape[1]), float(g["sf2"]), float(g["noise"])
    fit, _ = E.factorize(X, y, ell, sf2, noise)
    ref = g["mean"]
    for kw in ({}, {"want_var": False}, {"want_var": False, "mj_path": "simt"}, {"var_path": "f64"}):
        mean = E.predict(fit, Z.to(dtype), **kw)[0][0].double().cpu().numpy()
        err = np.abs(mean - ref).max() / np.abs(ref).max()
        assert err < (1e-9 if dtype == torch.float64 else tol), (name, kw, err)


def test_repo_casadi_export_matches_reference_export(monkeypatch):
    """gpmpc_b200.gp.gpytorch_predict2casadi (alpha from the GPU fit) evaluated over the NumPy casadi shim against the
    reference's own function on the same GP; gp.K against K_f + noise I of the reference's kernel (gp.py:45)."""
    import casadi_shim
    from gpmpc_b200 import gp as G

    monkeypatch.setattr(G, "cs", casadi_shim)
    g = np.load(os.path.join(GOLD, "pred_d3_n45.npz"))
    X, y = torch.from_numpy(g["X"]), torch.from_numpy(g["y"])
    gp = G.GaussianProcess(X, y).double()             # float64 parameters: the fixture's values are not float32 numbers
    gp.set_hyperparameters(torch.from_numpy(g["ell"]).reshape(-1), float(g["sf2"]), float(g["noise"]))
    gp.eval()
    pred = G.gpytorch_predict2casadi(gp)
    assert pred.name() == "pred"
    mean = np.array([float(np.asarray(pred(z=z)["mean"]).reshape(-1)[0]) for z in g["Z"]])
    assert np.abs(mean - g["mean"]).max() / np.abs(g["mean"]).max() < 1e-9
    K, K_inv = gp.compute_covariances()
    assert np.abs(K.numpy() - g["K"]).max() < 1e-13 * np.abs(g["K"]).max()
    assert np.abs(K_inv.numpy() - np.linalg.inv(g["K"])).max() < 1e-8 * np.abs(np.linalg.inv(g["K"])).max()
    # the CasADi-free callable serves the same mean
    call = G.predict_callable(gp)
    assert np.abs(call(g["Z"])["mean"] - g["mean"]).max() / np.abs(g["mean"]).max() < 1e-9


def test_casadi_callback_bridge_serves_mean_and_jacobian(monkeypatch):
    """gp_casadi_callback over the shim's Callback protocol: value == the reference's exported mean,
    Jacobian == the oracle's analytic input-Jacobian (what CasADi AD of gp.py:12-14, 85 yields)."""
    import casadi_shim
    from gpmpc_b200 import gp as G
    from oracle import gp_oracle as o

    monkeypatch.setattr(G, "cs", casadi_shim)
    g = np.load(os.path.join(GOLD, "pred_d3_n45.npz"))
    X, y = torch.from_numpy(g["X"]), torch.from_numpy(g["y"])
    ell, sf2, noise = torch.from_numpy(g["ell"]).reshape(-1), float(g["sf2"]), float(g["noise"])
    gp = G.GaussianProcess(X, y).double()
    gp.set_hyperparameters(ell, sf2, noise)
    gp.eval()
    cb = G.gp_casadi_callback(gp)
    assert cb.name() == "pred_gpu" and cb.get_n_in() == 1 and cb.has_jacobian()
    jac_fn = cb.jacobian()
    st = o.fit_state(X, y, ell.expand(3), sf2, noise)
    _, _, j_ref = o.predict(torch.from_numpy(g["Z"]), X, ell.expand(3), sf2, st["L"], st["alpha"])
    for i, z in enumerate(g["Z"][:8]):
        m = float(cb(z=z)["mean"].reshape(-1)[0])
        assert abs(m - g["mean"][i]) < 1e-9 * np.abs(g["mean"]).max()
        J = jac_fn(z, m)
        assert J.shape == (1, 3) and np.abs(J.reshape(-1) - j_ref[i].numpy()).max() < 1e-8 * float(j_ref.abs().max())


@pytest.mark.parametrize("name", ["sk_d1_n15", "sk_d3_n45", "sk_d16_n300"])
def test_cuda_fit_and_prediction_match_scikit_learn(name):
    """Rows a7-a9 against an independent third-party implementation (tests/golden/make_golden_sklearn.py:
    scikit-learn's GaussianProcessRegressor, sf2 * RBF(ARD) + White): MLL within 1e-6 relative, its gradient, the
    posterior mean and the predictive variance var_f + noise (element-wise) through the CUDA fit and every prediction
    path, float64 1e-6 / float32 1e-4 (BASELINE.json)."""
    from gpmpc_b200 import engine as E

    g = np.load(os.path.join(GOLD, name + ".npz"))
    X, y, Z = (torch.from_numpy(g[k]).to(DEV) for k in ("X", "y", "Z"))
    ell, sf2, noise = torch.from_numpy(g["ell"]), float(g["sf2"]), float(g["noise"])
    n, d = X.shape
    fit, ws = E.factorize(X, y, ell, sf2, noise, want_kinv=True)
    lml = fit.mll_per_n() * n
    assert abs(lml - float(g["lml"])) < 1e-6 * abs(float(g["lml"]))
    loss, g_noise, g_sf2, g_ell = E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
    grad_log = np.array([-n * g_sf2 * sf2] + [-n * ge * float(e) for ge, e in zip(g_ell, ell)] + [-n * g_noise * noise])
    assert np.abs(grad_log - g["grad_log"]).max() < 1e-6 * np.abs(g["grad_log"]).max()
    if n <= 128:                                     # the fused one-launch kernel the reference's sizes go through
        l2, gn2, gs2, ge2, _ = E.small_loss_and_grad(X, y, ell, sf2, noise)
        grad2 = np.array([-n * gs2 * sf2] + [-n * ge * float(e) for ge, e in zip(ge2, ell)] + [-n * gn2 * noise])
        assert abs(-n * l2 - float(g["lml"])) < 1e-6 * abs(float(g["lml"]))
        assert np.abs(grad2 - g["grad_log"]).max() < 1e-6 * np.abs(g["grad_log"]).max()
    for dtype, tol in ((torch.float64, 1e-6), (torch.float32, 1e-4)):
        for kw in ({}, {"var_path": "f64"}):
            mean, var, _ = E.predict(fit, Z.to(dtype), **kw)
            m = mean[0].double().cpu().numpy()
            v = var[0].double().cpu().numpy() + noise
            assert np.abs(m - g["mean"]).max() < tol * np.abs(g["mean"]).max(), (name, dtype, kw)
            assert (np.abs(v - g["var_y"]) / g["var_y"]).max() < tol, (name, dtype, kw)
