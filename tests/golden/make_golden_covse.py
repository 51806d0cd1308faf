"""Golden vectors for SURVEY.md 8(a) rows a1, a2 and a6  --  produced by the REFERENCE'S OWN CODE.

`gpmpc/gp.py` cannot be imported as is (casadi and gpytorch are absent), but `covSE_single`, `covSE_vectorized`
(gpmpc/gp.py:12-21) and `gpytorch_predict2casadi` (gpmpc/gp.py:72-85) are plain arithmetic over the `cs`
namespace.  This script loads that file from /root/reference with
  * `casadi`   replaced by tests/casadi_shim.py (NumPy float64 with CasADi's shape conventions), and
  * `gpytorch` replaced by an empty stub (only touched at import time and as the base class of GaussianProcess),
and calls the UNMODIFIED functions:

  covse_*.npz   x (d,), Z (M,d), ell, sf2  ->  k_single = covSE_single(x, Z^T, ell^T, sf2)       (the call at gp.py:82)
                                               k_vec    = covSE_vectorized(x, Z, ell^T, sf2)      (the call at gpmpc.py:417)
                for d in {1, 3, 16}, the reference's shared (1,1) lengthscale and an ARD (1,d) one
  pred_*.npz    a reference GaussianProcess instance (built without gpytorch: attributes set by hand, K_inv from
                torch.linalg.inv like gp.py:46)  ->  mean(z) of the CasADi function "pred" at Q queries

What is pinned: the kernel formula and the literal mean formula k(z,X) @ K_inv @ y of the reference; the gpytorch
internals (Cholesky, MLL, prediction strategy) stay unpinned (oracle/gp_oracle.py header).
/root/reference exists only in the build container: the fixtures are committed.

Run from the repo root:  python tests/golden/make_golden_covse.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import casadi_shim  # noqa: E402

REF = "/root/reference/gpmpc/gp.py"


def load_reference():
    """gpmpc/gp.py with casadi -> the NumPy shim and gpytorch -> an inert stub."""
    class _Anything(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return object
    stubs = {"casadi": casadi_shim}
    for m in ["gpytorch", "gpytorch.distributions", "gpytorch.kernels", "gpytorch.likelihoods", "gpytorch.means",
              "gpytorch.models", "gpytorch.constraints", "gpytorch.mlls"]:
        stubs[m] = _Anything(m)
    stubs["gpytorch"].models = stubs["gpytorch.models"]
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    try:
        spec = importlib.util.spec_from_file_location("_ref_gp", REF)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod


def covse_case(ref, name, seed, d, m, ard):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1.0, 1.0, size=d)
    Z = rng.uniform(-1.0, 1.0, size=(m, d))
    ell = rng.uniform(0.3, 1.7, size=(1, d)) if ard else np.array([[0.8 + 0.1 * d]])   # gpytorch: (1,d) ARD, (1,1) shared
    sf2 = float(rng.uniform(0.5, 2.0))
    k_single = np.asarray(ref.covSE_single(x[:, None], Z.T, ell.T, sf2))              # gp.py:82 passes X^T, ell^T
    k_vec = np.asarray(ref.covSE_vectorized(x[:, None], Z, ell.T, sf2))               # gpmpc.py:417
    assert k_single.shape == (1, m) and k_vec.shape == (1, m)
    np.savez(os.path.join(HERE, name + ".npz"), x=x, Z=Z, ell=ell, sf2=np.float64(sf2), k_single=k_single, k_vec=k_vec)
    print(name, "d", d, "M", m, "ard", ard, "max|k_single - k_vec|", float(np.abs(k_single - k_vec).max()))


def pred_case(ref, name, seed, n, d, q, noise):
    """The reference's CasADi mean function on a GaussianProcess whose gpytorch parts are set by hand."""
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1.0, 1.0, size=(n, d))
    w = rng.normal(size=d)
    y = np.sin(X @ w) + 0.05 * rng.normal(size=n)
    Zq = rng.uniform(-1.2, 1.2, size=(q, d))
    ell, sf2 = np.array([[0.5 * np.sqrt(d) + 0.2]]), 1.3
    gp = object.__new__(ref.GaussianProcess)
    gp.train_inputs, gp.train_targets = (torch.tensor(X),), torch.tensor(y)
    gp.covar_module = types.SimpleNamespace(base_kernel=types.SimpleNamespace(lengthscale=torch.tensor(ell)),
                                            outputscale=torch.tensor(sf2, dtype=torch.float64))
    # K = K_f + noise I with the reference's own kernel, K_inv = torch.linalg.inv(K) as at gp.py:45-46
    K = np.vstack([np.asarray(ref.covSE_single(X[i][:, None], X.T, ell.T, sf2)) for i in range(n)]) + noise * np.eye(n)
    gp.K, gp.K_inv = torch.tensor(K), torch.linalg.inv(torch.tensor(K))
    pred = ref.gpytorch_predict2casadi(gp)
    mean = np.array([float(np.asarray(pred(z=Zq[i])["mean"]).reshape(-1)[0]) for i in range(q)])
    extra = {"K": K} if n <= 64 else {}        # K = K_f + noise I of the reference's own kernel (pins gp.K, gp.py:45)
    np.savez(os.path.join(HERE, name + ".npz"), X=X, y=y, Z=Zq, ell=ell, sf2=np.float64(sf2), noise=np.float64(noise),
             mean=mean, **extra)
    print(name, "N", n, "d", d, "Q", q, "max|mean|", float(np.abs(mean).max()))


def main():
    ref = load_reference()
    for d in (1, 3, 16):
        covse_case(ref, f"covse_d{d}_shared", seed=100 + d, d=d, m=45 if d < 16 else 300, ard=False)
        covse_case(ref, f"covse_d{d}_ard", seed=200 + d, d=d, m=45 if d < 16 else 300, ard=True)
    pred_case(ref, "pred_d3_n45", seed=31, n=45, d=3, q=25, noise=1e-2)       # the reference's shapes (yaml: N <= 45, T = 25)
    pred_case(ref, "pred_d16_n300", seed=32, n=300, d=16, q=64, noise=1e-2)


if __name__ == "__main__":
    main()
