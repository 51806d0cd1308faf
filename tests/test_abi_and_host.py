"""CPU-only tests (-m "not gpu"): the C-ABI library loads and exports every symbol the header declares,
the host-side GP surface mirrors the reference, and the product path fails loudly without a GPU."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from gpmpc_b200 import _lib

    hdr = open(os.path.join(ROOT, "include", "gpmpc_b200.h")).read()
    declared = set(re.findall(r"\b(gpk_\w+)\s*\(", hdr))
    assert declared, "no declarations found in include/gpmpc_b200.h"
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    lib = _lib.load()           # raises if a declared symbol is not exported
    for name in declared:
        assert hasattr(lib, name)
    assert lib.gpk_abi_version() == 2
    # size queries do no GPU work
    assert lib.gpk_pad_dim(1) == 1 and lib.gpk_pad_dim(5) == 8 and lib.gpk_pad_dim(16) == 16 and lib.gpk_pad_dim(33) == -1
    assert lib.gpk_hyp_stride() == 72
    assert lib.gpk_pack_train_elems(2048, 16) == 2048 * 18
    assert lib.gpk_pack_train_elems(15, 3) == 256 * 5
    assert lib.gpk_trtri_scratch_elems(128) == 0 and lib.gpk_trtri_scratch_elems(2048) == 1024 * 1024
    assert lib.gpk_predict_var_work_bytes(100, 256, 2) == (128 * 256 + 2 * 128) * 8 * 2


def test_header_cites_reference_call_sites():
    hdr = open(os.path.join(ROOT, "include", "gpmpc_b200.h")).read()
    for cite in ("gpmpc/gp.py:45", "gpmpc/gp.py:62", "gpmpc/gp.py:46", "gpmpc/gp.py:85", "gpmpc/gpmpc.py:443-445"):
        assert cite in hdr


def test_bad_arguments_return_status_not_crash():
    from gpmpc_b200 import _lib

    lib = _lib.load()
    # null pointers / bad sizes are rejected before any CUDA call
    assert lib.gpk_potrf(None, 100, 100, None, None, None) == -1
    assert b"npad" in lib.gpk_last_error()
    assert lib.gpk_dgemm_nt(None, 16, None, 16, None, 16, 100, 128, 16, 1.0, 0.0, 0, 0, None) == -1
    with pytest.raises(_lib.GpkError):
        _lib.call("gpk_pack_train", 7, None, 1, 1, 1, None, None, None, 1, None)


def test_gp_surface_matches_reference():
    """Attribute / method surface that gpmpc/gpmpc.py touches, checked without a GPU."""
    import inspect

    import gpmpc_b200 as G
    from gpmpc_b200.gp import GaussianProcess, fit_gp, gpytorch_predict2casadi

    x = torch.rand(15, 3, dtype=torch.float64)
    y = torch.rand(15, dtype=torch.float64)
    gp = GaussianProcess(x, y)
    assert gp.train_inputs[0] is x and gp.train_targets is y
    assert gp.n_ind_points == 15 and gp.input_dimension == 3
    assert gp.K is None and gp.K_inv is None
    names = [n for n, _ in gp.named_parameters()]
    assert names == ["likelihood.noise_covar.raw_noise", "covar_module.raw_outputscale",
                     "covar_module.base_kernel.raw_lengthscale"]
    assert all(p.dtype == torch.float32 for p in gp.parameters())       # gpytorch default dtype
    # raw = 0 -> noise = ln2 + 1e-6, sf2 = ell = ln2
    ell, sf2, noise = gp.hyperparameters()
    ln2 = 0.6931471805599453
    assert abs(sf2 - ln2) < 1e-7 and abs(noise - ln2 - 1e-6) < 1e-7 and torch.allclose(ell, torch.full((3,), ln2, dtype=torch.float64), atol=1e-7)
    assert gp.covar_module.base_kernel.lengthscale.shape == (1, 1)
    assert gp.covar_module.base_kernel.lengthscale.to_dense().shape == (1, 1)   # gp.py:78 calls .to_dense()
    assert gp.covar_module.outputscale.dim() == 0 and gp.likelihood.noise.shape == (1,)
    assert gp.train() is gp and gp.eval() is gp and gp.to("cpu") is gp
    sig = inspect.signature(fit_gp)
    assert list(sig.parameters)[:4] == ["gp", "n_train", "lr", "device"]
    assert sig.parameters["n_train"].default == 500 and sig.parameters["lr"].default == 0.01
    assert sig.parameters["device"].default == "cpu"
    assert list(inspect.signature(gpytorch_predict2casadi).parameters) == ["gp"]
    for name in ("GaussianProcess", "fit_gp", "gpytorch_predict2casadi", "covSE_single", "covSE_vectorized"):
        assert hasattr(G, name)
    with pytest.raises(AssertionError):
        GaussianProcess([[1.0]], y)            # reference asserts tensor inputs (gp.py:29-30)
    with pytest.raises(AssertionError):
        fit_gp("not a gp")


def test_set_hyperparameters_roundtrip_and_ard():
    from gpmpc_b200.gp import GaussianProcess

    gp = GaussianProcess(torch.rand(10, 4, dtype=torch.float64), torch.rand(10, dtype=torch.float64), ard=True)
    assert gp.covar_module.base_kernel.lengthscale.shape == (1, 4)
    gp.set_hyperparameters([0.5, 1.0, 1.5, 2.0], 1.3, 0.02)
    ell, sf2, noise = gp.hyperparameters()
    assert torch.allclose(ell, torch.tensor([0.5, 1.0, 1.5, 2.0], dtype=torch.float64), rtol=1e-6)
    assert abs(sf2 - 1.3) < 1e-6 and abs(noise - 0.02) < 1e-7


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_product_path_fails_loudly_without_gpu():
    from gpmpc_b200.gp import GaussianProcess, fit_gp

    gp = GaussianProcess(torch.rand(10, 2, dtype=torch.float64), torch.rand(10, dtype=torch.float64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fit_gp(gp, n_train=2)
    gp.eval()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        gp(torch.rand(3, 2, dtype=torch.float64)).mean
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        gp.compute_covariances()


def test_casadi_exports_are_import_guarded():
    from gpmpc_b200 import gp as G

    if G.cs is None:
        with pytest.raises(ImportError, match="casadi"):
            G.covSE_single(None, None, None, None)
        with pytest.raises(ImportError, match="casadi"):
            G.covSE_vectorized(None, None, None, None)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gpmpc_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("the oracle", "").replace("The oracle", "") or "import" not in \
                "".join(l for l in src.splitlines() if "oracle" in l), fn
