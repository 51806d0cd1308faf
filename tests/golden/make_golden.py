"""Writes the golden fixtures tests/golden/*.npz from the CPU oracle (oracle/gp_oracle.py).

The reference (gpmpc/gp.py) cannot be imported here: gpytorch / linear_operator / casadi are absent
and not installable, so these vectors come from the torch-float64 restatement,
not from gpytorch itself -- parity stays "unpinned" and the fixtures only freeze the oracle's
arithmetic so the CUDA path, the oracle and later rounds are all compared with the same numbers.
Shapes follow the reference run (P=3 GPs with d=(1,3,3), N in {15,30,45}, Q=T=25:
gpmpc/gpmpc.py:59, scripts/gp_mpc_config.yaml:6,24,30) plus two synthetic d=16 cases.

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import gp_oracle as o  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [  # name, N, d, Q, ard, seed, (ell scale, sf2, noise)
    ("ref_d1_n15", 15, 1, 25, False, 1, (0.9, 0.8, 2e-2)),
    ("ref_d3_n30", 30, 3, 25, False, 2, (1.3, 1.2, 1e-2)),
    ("ref_d3_n45", 45, 3, 25, False, 3, (0.7, 0.5, 1e-3)),
    ("syn_d16_n256", 256, 16, 64, True, 4, (2.0, 1.0, 1e-2)),
    ("syn_d16_n700", 700, 16, 48, True, 5, (2.0, 1.0, 1e-2)),
]


def main():
    for name, n, d, q, ard, seed, (es, sf2, noise) in CASES:
        p = o.synthetic_problem(n, d, q, p=1, seed=seed)
        X, y, Z = p["X"], p["Y"][:, 0], p["Z"]
        g = torch.Generator().manual_seed(100 + seed)
        if ard:
            ell = es * (0.75 + 0.5 * torch.rand(d, generator=g, dtype=torch.float64))
        else:
            ell = torch.full((d,), es, dtype=torch.float64)
        st = o.fit_state(X, y, ell, sf2, noise)
        mean, var, jac = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
        raw_noise = o.inv_softplus(torch.tensor([noise - o.NOISE_FLOOR], dtype=torch.float64))
        raw_sf2 = o.inv_softplus(torch.tensor(sf2, dtype=torch.float64))
        raw_ell = o.inv_softplus(ell.reshape(1, d) if ard else ell[:1].reshape(1, 1))
        loss, gn, gs, gl = o.loss_and_grad(X, y, raw_noise, raw_sf2, raw_ell)
        np.savez(
            os.path.join(HERE, name + ".npz"),
            X=X.numpy(), y=y.numpy(), Z=Z.numpy(), ell=ell.numpy(), sf2=np.float64(sf2), noise=np.float64(noise),
            alpha=st["alpha"].numpy(), logdet=st["logdet"].numpy(), quad=st["quad"].numpy(),
            mll_per_n=o.mll_per_n(X, y, ell, sf2, noise).numpy(),
            mean=mean.numpy(), var=var.numpy(), jac=jac.numpy(),
            raw_noise=raw_noise.numpy(), raw_sf2=raw_sf2.numpy(), raw_ell=raw_ell.numpy(),
            loss_raw=loss.numpy(), g_raw_noise=gn.numpy(), g_raw_sf2=gs.numpy(), g_raw_ell=gl.numpy(),
        )
        print(name, "N", n, "d", d, "loss", float(loss))


if __name__ == "__main__":
    main()
