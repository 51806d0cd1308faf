"""Golden vectors of the exact-GP core from an INDEPENDENT third-party implementation: scikit-learn's
GaussianProcessRegressor (Rasmussen & Williams, Algorithm 2.1) with the kernel  sf2 * RBF(ell, ARD) + White(noise).

gpytorch (the reference's dependency, gpmpc/gp.py:24-46) is not installable here, so the gpytorch internals behind
SURVEY.md 8(a) rows a7-a9 stay unpinned to the REFERENCE; these fixtures pin the same published formulas -- exact
log marginal likelihood, its gradient, posterior mean and predictive variance -- to a second, widely used library that
shares no code with this repository or its oracle.  sklearn has no input-Jacobian; that stays with the oracle's
autograd / finite-difference checks.

    lml           log p(y | X)                     = N * (gpytorch's ExactMarginalLogLikelihood value, gp.py:57)
    grad_log      d lml / d log(sf2, ell_1..d, noise)
    mean, var_y   posterior mean and predictive variance var_f + noise at Z   (gpmpc.py:441-445 keeps this diagonal)

Run from the repo root:  python tests/golden/make_golden_sklearn.py   (needs scikit-learn; the tests only read the .npz)
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import gp_oracle as o  # noqa: E402  (only for the seeded synthetic inputs)

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [  # name, N, d, Q, seed, ell scale, sf2, noise
    ("sk_d1_n15", 15, 1, 25, 11, 0.9, 0.8, 2e-2),
    ("sk_d3_n45", 45, 3, 25, 12, 0.7, 0.5, 1e-3),
    ("sk_d16_n300", 300, 16, 40, 13, 2.0, 1.3, 1e-2),
]


def sklearn_gp(X, y, Z, ell, sf2, noise):
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel, WhiteKernel

    kernel = ConstantKernel(sf2) * RBF(length_scale=ell) + WhiteKernel(noise)
    gpr = GaussianProcessRegressor(kernel=kernel, alpha=0.0, optimizer=None, normalize_y=False).fit(X, y)
    theta = gpr.kernel_.theta                       # log(sf2), log(ell_1..d), log(noise)
    lml, grad = gpr.log_marginal_likelihood(theta, eval_gradient=True)
    mean, std = gpr.predict(Z, return_std=True)     # std of y* (the White kernel's level is part of k(z, z))
    return float(lml), np.asarray(grad), mean, std ** 2


def main():
    for name, n, d, q, seed, es, sf2, noise in CASES:
        p = o.synthetic_problem(n, d, q, p=1, seed=seed)
        X, y, Z = p["X"].numpy(), p["Y"][:, 0].numpy(), p["Z"].numpy()
        g = torch.Generator().manual_seed(200 + seed)
        ell = (es * (0.75 + 0.5 * torch.rand(d, generator=g, dtype=torch.float64))).numpy()
        lml, grad, mean, var_y = sklearn_gp(X, y, Z, ell, sf2, noise)
        np.savez(os.path.join(HERE, name + ".npz"), X=X, y=y, Z=Z, ell=ell, sf2=np.float64(sf2), noise=np.float64(noise),
                 lml=np.float64(lml), grad_log=grad, mean=mean, var_y=var_y)
        print(name, "lml/N", lml / n)


if __name__ == "__main__":
    main()
