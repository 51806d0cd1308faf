"""gpmpc_b200 -- B200-native (sm_100a) Gaussian-process fit/predict path of gp-mpc.

Drop-in for the GP layer of the reference (`gpmpc/gp.py`): `GaussianProcess`, `fit_gp`,
`gpytorch_predict2casadi`, `covSE_single`, `covSE_vectorized`, plus the batched `predict` entry.
"""

from .gp import (  # noqa: F401
    GaussianProcess,
    covSE_single,
    covSE_vectorized,
    fit_gp,
    fit_gps,
    gp_casadi_callback,
    gpytorch_predict2casadi,
    predict_batch,
    predict_callable,
    predict_host,
)
