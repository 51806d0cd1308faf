"""The accuracy policy of the int8 variance path over the fit's conditioning (engine.select_digits).

The digit images quantise W = L^-1 relative to its row maxima (~ 1/sqrt(noise)), so the ABSOLUTE variance error of a
fixed digit count grows as the noise shrinks while the predictive variance var_f + noise the reference consumes
(gpmpc/gpmpc.py:444-445) shrinks with it.  These tests hold the path predict() actually takes -- 5 digits, 6 digits or
the FP64 DMMA fallback, chosen per fit -- to BASELINE.json's tolerances ELEMENT-WISE on the predictive variance, at the
reference's input dimensions (d = 1, 3: gpmpc.py:59) and down to its noise floor 1e-6 (gp.py:31).  Run with -m gpu."""
import math

import pytest
import torch

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _problem(d, n, noise, sf2=1.0):
    g = torch.Generator().manual_seed(100 * d + n)
    X = torch.rand(n, d, generator=g, dtype=torch.float64) * 2 - 1
    y = math.sqrt(sf2) * torch.sin(X @ torch.randn(d, generator=g, dtype=torch.float64)) \
        + math.sqrt(noise) * torch.randn(n, generator=g, dtype=torch.float64)
    Zr = (torch.rand(300, d, generator=g, dtype=torch.float64) * 2 - 1) * 1.2
    Z = torch.cat([X[:: n // 200][:200], Zr])            # at training inputs (smallest variance) and away from them
    ell = torch.full((d,), 0.5 if d == 1 else 0.8, dtype=torch.float64)
    return X, y, Z, ell


@pytest.mark.parametrize("noise", [1e-2, 1e-4, 1e-5, 1e-6])
@pytest.mark.parametrize("d", [1, 3])
def test_selected_path_meets_elementwise_tolerance(d, noise):
    from gpmpc_b200 import engine as E

    n, sf2 = 1024, 1.0
    X, y, Z, ell = _problem(d, n, noise)
    fit, _ = E.factorize(X.to(DEV), y.to(DEV), ell, sf2, noise)          # digits chosen by the policy
    assert fit.digits in (0, 5, 6) and (fit.digits == 0) == (fit.wimg is None)
    if fit.digits:
        assert E.VAR_PROBE_SAFETY * fit.var_err[fit.digits] <= E.VAR_RTOL_F64 * noise
    Zd = Z.to(DEV)
    v_ref = E.predict(fit, Zd, var_path="f64")[1][0]
    # (1) against the GPU's own FP64 path: the policy's guarantee, element-wise on var_f + noise
    v64 = E.predict(fit, Zd)[1][0]
    err64 = float(((v64 - v_ref).abs() / (v_ref.clamp_min(0.0) + noise)).max())
    assert err64 < 1e-6, (d, noise, fit.digits, err64)
    Z32 = Zd.float()
    v_ref32 = E.predict(fit, Z32.double(), var_path="f64")[1][0]       # same (float32-rounded) queries
    v32 = E.predict(fit, Z32)[1][0].double()
    # float32 output rounding alone is 6e-8 relative to var_f <= sf2: allow it on top of the 1e-4 relative tolerance
    err32 = float(((v32 - v_ref32).abs() / (v_ref32.clamp_min(0.0) + noise + 1e-3 * sf2)).max())
    assert err32 < 1e-4, (d, noise, fit.digits_f32, err32)
    # (2) against the CPU oracle.  Two float64 formulations of the reference itself (Cholesky solve vs the literal
    # k K^-1 k with torch.linalg.inv, gp.py:46) already disagree by the conditioning of K; that disagreement is the
    # resolution of the comparison
    st = o.fit_state(X, y, ell, sf2, noise)
    _, v_chol, _ = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    _, v_inv = o.predict_explicit_inverse(Z, X, y, ell, sf2, noise)
    inherent = float((v_chol - v_inv).abs().max())
    gap = float(((v64.cpu() - v_chol).abs() - 4.0 * inherent).clamp_min(0.0).div(v_chol.clamp_min(0.0) + noise).max())
    assert gap < 1e-6, (d, noise, fit.digits, gap, inherent)
    m64, _, j64 = E.predict(fit, Zd)
    m_ref, _, j_ref = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    assert o.rel_err(m64[0].cpu(), m_ref) < 1e-6 and o.rel_err(j64[0].cpu(), j_ref) < 1e-6


def test_policy_picks_more_digits_as_noise_shrinks():
    """5 digits at the benchmark's noise, fewer guarantees -> more digits -> DMMA fallback as the noise goes to the
    reference's floor; explicit digit counts bypass the policy."""
    from gpmpc_b200 import engine as E

    picked = {}
    for noise in (1e-2, 1e-4, 1e-6):
        X, y, Z, ell = _problem(1, 1024, noise)
        fit, _ = E.factorize(X.to(DEV), y.to(DEV), ell, 1.0, noise)
        picked[noise] = fit.digits
        assert 5 in fit.var_err
    assert picked[1e-2] == 5
    rank = {5: 0, 6: 1, 0: 2}
    assert rank[picked[1e-2]] <= rank[picked[1e-4]] <= rank[picked[1e-6]] and picked[1e-6] != 5
    X, y, Z, ell = _problem(1, 1024, 1e-6)
    forced, _ = E.factorize(X.to(DEV), y.to(DEV), ell, 1.0, 1e-6, digits=5)
    assert forced.digits == 5 and forced.wimg is not None and forced.digits_f32 == 4


def test_batch_shares_one_digit_count():
    """A GPBatch has one image format: the policy must satisfy every GP of the batch (here one well and one badly
    conditioned GP on the same inputs)."""
    from gpmpc_b200 import engine as E

    X, y, Z, ell = _problem(3, 512, 1e-2)
    Y = torch.stack([y, y * 0.5], dim=1)
    gb = E.fit_batch(X.to(DEV), Y.to(DEV), ell, [1.0, 1.0], [1e-2, 1e-6])
    assert all(f.digits == gb.digits for f in gb.fits)
    v = E.predict(gb, Z.to(DEV))[1]
    for p, noise in enumerate((1e-2, 1e-6)):
        st = o.fit_state(X, Y[:, p], ell, 1.0, noise)
        _, v_ref, _ = o.predict(Z, X, ell, 1.0, st["L"], st["alpha"])
        _, v_inv = o.predict_explicit_inverse(Z, X, Y[:, p], ell, 1.0, noise)
        inherent = float((v_ref - v_inv).abs().max())
        gap = float(((v[p].cpu() - v_ref).abs() - 4.0 * inherent).clamp_min(0.0).div(v_ref.clamp_min(0.0) + noise).max())
        assert gap < 1e-6, (p, gb.digits, gap)
