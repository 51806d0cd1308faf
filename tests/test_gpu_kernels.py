"""GPU parity tests of the individual C-ABI entry points against the CPU oracle (run with -m gpu).

Tolerances follow BASELINE.json: 1e-6 relative (normwise max|a-b|/max|b|) on the float64 path, 1e-4 on
the float32 path.  All calls go through the C ABI of include/gpmpc_b200.h via ctypes.
"""
import math

import pytest
import torch

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu

TOL64 = 1e-6
TOL32 = 1e-4


def _dev():
    return torch.device("cuda", 0)


def _lib():
    from gpmpc_b200 import _lib as L

    return L


@pytest.mark.parametrize("M,N,K,a_tri,c_lower,alpha,beta", [
    (128, 128, 16, 0, 0, 1.0, 0.0),
    (128, 128, 128, 0, 0, 1.0, 0.0),
    (256, 384, 512, 0, 0, -1.0, 1.0),
    (384, 128, 384, 1, 0, -1.0, 0.0),
    (512, 256, 512, 2, 0, 1.0, 0.0),
    (640, 640, 256, 0, 1, -1.0, 1.0),
    (2560, 2560, 2560, 2, 1, 1.0, 0.0),   # > one super-tile row (20 tile rows)
    (3072, 1280, 3072, 1, 0, 1.0, 0.0),
])
def test_dgemm_nt(M, N, K, a_tri, c_lower, alpha, beta):
    L = _lib()
    g = torch.Generator().manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g, dtype=torch.float64)
    B = torch.randn(N, K, generator=g, dtype=torch.float64)
    C0 = torch.randn(M, N, generator=g, dtype=torch.float64)
    if a_tri == 1:
        A = torch.tril(A)
    if a_tri == 2:
        A = torch.triu(A)
    Ad, Bd, Cd = A.to(_dev()), B.to(_dev()), C0.to(_dev())
    L.call("gpk_dgemm_nt", L.ptr(Ad), K, L.ptr(Bd), K, L.ptr(Cd), N, M, N, K, alpha, beta, a_tri, c_lower, L.stream_ptr())
    ref = beta * C0 + alpha * (A @ B.T)
    out = Cd.cpu()
    if c_lower:
        # only tiles that touch the lower triangle are computed (tile size depends on the problem size):
        # the lower triangle is defined, everything above the diagonal is unspecified
        low = torch.arange(M)[:, None] >= torch.arange(N)[None, :]
        out, ref = torch.where(low, out, 0.0), torch.where(low, ref, 0.0)
    assert o.rel_err(out, ref) < 1e-12


@pytest.mark.parametrize("n,d,q", [(15, 1, 25), (45, 3, 25), (200, 5, 70), (300, 16, 130)])
def test_cross_kernel_and_assembly(n, d, q):
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(n, d, q, seed=n)
    X, y, Z = p["X"], p["Y"][:, 0], p["Z"]
    ell = p["ell"] * torch.linspace(0.8, 1.3, d, dtype=torch.float64)
    fit, _ = E.factorize(X.to(_dev()), y.to(_dev()), ell, 1.3, 0.02)
    K = E.dense_K(fit).cpu()
    Kref = o.assemble_K(X, ell, 1.3, 0.02)
    assert o.rel_err(K, Kref) < 1e-13
    assert torch.equal(K, K.T), "assembled K must be exactly symmetric"
    assert torch.equal(K.diagonal(), torch.full((n,), 1.3 + 0.02, dtype=torch.float64))
    Kzx = E.cross_K(fit, Z.to(_dev()), X.to(_dev())).cpu()
    assert o.rel_err(Kzx, o.se_kernel(Z, X, ell, 1.3)) < 1e-13


@pytest.mark.parametrize("n,d", [(15, 1), (45, 3), (128, 3), (129, 5), (300, 16), (1000, 16), (2048, 16)])
def test_factorize_matches_oracle(n, d):
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(n, d, 1, seed=7 + n)
    X, y, ell, sf2, noise = p["X"], p["Y"][:, 0], p["ell"], p["sf2"], p["noise"]
    fit, ws = E.factorize(X.to(_dev()), y.to(_dev()), ell, sf2, noise, want_kinv=True)
    st = o.fit_state(X, y, ell, sf2, noise)
    Lg = torch.tril(ws.L[:n, :n]).cpu()
    assert o.rel_err(Lg, st["L"]) < 1e-11
    Wref = torch.linalg.solve_triangular(st["L"], torch.eye(n, dtype=torch.float64), upper=False)
    assert o.rel_err(fit.W[:n, :n].cpu(), Wref) < 1e-9
    assert o.rel_err(ws.U[:n, :n].cpu(), Wref.T) < 1e-9
    assert o.rel_err(fit.alpha[:n].cpu(), st["alpha"]) < TOL64 * 1e-2
    assert float(fit.alpha[n:].abs().max()) == 0.0 if fit.npad > n else True
    assert abs(fit.quad - float(st["quad"])) < 1e-9 * abs(float(st["quad"]))
    assert abs(fit.logdet - float(st["logdet"])) < 1e-10 * max(1.0, abs(float(st["logdet"])))
    mll_ref = float(o.mll_per_n(X, y, ell, sf2, noise))
    assert abs(fit.mll_per_n() - mll_ref) < TOL64 * abs(mll_ref)
    Kinv = ws.Kinv[:n, :n].cpu()
    assert o.rel_err(Kinv, torch.cholesky_inverse(st["L"])) < 1e-8
    assert torch.equal(Kinv, Kinv.T)


def test_potrf_reports_non_pd_and_jitter_retry():
    from gpmpc_b200 import engine as E

    # duplicated rows with (almost) no noise: exactly singular K -> jitter retries kick in
    X = torch.rand(40, 3, dtype=torch.float64)
    X = torch.cat([X, X[:10]])
    y = torch.randn(50, dtype=torch.float64)
    fit, _ = E.factorize(X.to(_dev()), y.to(_dev()), torch.ones(3, dtype=torch.float64), 1.0, 0.0)
    assert fit.jitter > 0.0 and math.isfinite(fit.quad)
    # a genuinely indefinite request (negative noise) must raise, like gpytorch's NotPSDError
    with pytest.raises(E.NotPSDError):
        E.factorize(X.to(_dev()), y.to(_dev()), torch.ones(3, dtype=torch.float64), 1.0, -0.5)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, TOL64), (torch.float32, TOL32)])
@pytest.mark.parametrize("n,d,q", [(15, 1, 25), (30, 3, 25), (45, 3, 25), (300, 2, 77), (513, 8, 1000), (700, 16, 333),
                                    (2048, 16, 4096), (1000, 20, 64)])
def test_predict_matches_oracle(dtype, tol, n, d, q):
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(n, d, q, seed=3 * n + d)
    X, y, Z, ell, sf2, noise = p["X"], p["Y"][:, 0], p["Z"], p["ell"], p["sf2"], p["noise"]
    fit, _ = E.factorize(X.to(_dev()), y.to(_dev()), ell, sf2, noise)
    mean, var, jac = E.predict(fit, Z.to(_dev()).to(dtype))
    st = o.fit_state(X, y, ell, sf2, noise)
    m_ref, v_ref, j_ref = o.predict(Z, X, ell, sf2, st["L"], st["alpha"])
    assert mean.dtype == dtype and mean.shape == (1, q) and jac.shape == (1, q, d)
    assert o.rel_err(mean[0].cpu(), m_ref) < tol
    assert o.rel_err(jac[0].cpu(), j_ref) < tol
    assert o.rel_err(var[0].cpu(), v_ref) < tol


def test_predict_batched_three_gps():
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(640, 16, 500, p=3, seed=99)
    X, Y, Z = p["X"], p["Y"], p["Z"]
    ell = torch.stack([p["ell"] * s for s in (0.8, 1.0, 1.3)])
    sf2 = [0.7, 1.0, 1.6]
    noise = [1e-2, 3e-2, 5e-3]
    gb = E.fit_batch(X.to(_dev()), Y.to(_dev()), ell, sf2, noise)
    mean, var, jac = E.predict(gb, Z.to(_dev()))
    for i in range(3):
        st = o.fit_state(X, Y[:, i], ell[i], sf2[i], noise[i])
        m_ref, v_ref, j_ref = o.predict(Z, X, ell[i], sf2[i], st["L"], st["alpha"])
        assert o.rel_err(mean[i].cpu(), m_ref) < TOL64
        assert o.rel_err(var[i].cpu(), v_ref) < TOL64
        assert o.rel_err(jac[i].cpu(), j_ref) < TOL64


def test_predict_var_chunking_and_empty():
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(300, 4, 1000, seed=5)
    X, y, Z, ell = p["X"], p["Y"][:, 0], p["Z"], p["ell"]
    fit, _ = E.factorize(X.to(_dev()), y.to(_dev()), ell, 1.0, 1e-2)
    small = _lib().load().gpk_predict_var_work_bytes(256, fit.npad, 1)    # forces 4 chunks of 256 queries
    _, v_small, _ = E.predict(fit, Z.to(_dev()), var_work_bytes=small)
    _, v_big, _ = E.predict(fit, Z.to(_dev()))
    assert torch.equal(v_small, v_big)
    m, v, j = E.predict(fit, torch.empty(0, 4, dtype=torch.float64, device=_dev()))
    assert m.shape == (1, 0) and j.shape == (1, 0, 4)


@pytest.mark.parametrize("n,d,ard", [(30, 1, False), (45, 3, False), (200, 5, True), (600, 16, True)])
def test_mll_gradient_matches_oracle(n, d, ard):
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(n, d, 1, seed=31 + n)
    X, y = p["X"], p["Y"][:, 0]
    g = torch.Generator().manual_seed(n)
    rn = torch.randn(1, generator=g, dtype=torch.float64) - 2.0
    rs = torch.randn((), generator=g, dtype=torch.float64)
    rl = torch.randn(1, d if ard else 1, generator=g, dtype=torch.float64) + 1.0
    noise, sf2, ell = o.constrained(rn, rs, rl)
    fit, ws = E.factorize(X.to(_dev()), y.to(_dev()), ell.reshape(-1), float(sf2), float(noise), want_kinv=True)
    # row-split in two halves and summed == what two GPUs would all-reduce
    h = n // 2
    part = E.grad_partials(fit, ws, 0, h).clone() + E.grad_partials(fit, ws, h, n)
    loss, g_noise, g_sf2, g_ell = E.loss_and_grad_from_partials(fit, part)
    l_ref, gn_ref, gs_ref, gl_ref = o.loss_and_grad(X, y, rn, rs, rl)
    assert abs(loss - float(l_ref)) < TOL64 * abs(float(l_ref))
    # chain to raw parameters exactly as the host side of fit_gp does
    gn = g_noise * float(torch.sigmoid(rn))
    gs = g_sf2 * float(torch.sigmoid(rs))
    gl = torch.tensor(g_ell, dtype=torch.float64)
    gl = (gl.sum().reshape(1, 1) if not ard else gl.reshape(1, d)) * torch.sigmoid(rl)
    assert abs(gn - float(gn_ref)) < TOL64 * max(abs(float(gn_ref)), 1e-3)
    assert abs(gs - float(gs_ref)) < TOL64 * max(abs(float(gs_ref)), 1e-3)
    assert o.rel_err(gl, gl_ref) < TOL64


@pytest.mark.parametrize("n,d", [(15, 1), (30, 3), (45, 3), (100, 16), (128, 5)])
def test_fused_small_mll_matches_blocked_path_and_oracle(n, d):
    """gpk_mll_small (one CTA per GP, one launch per Adam iteration at the reference's sizes) against the blocked
    factorize + grad_partials path and the oracle's closed-form gradient; batch of three GPs in one launch."""
    from gpmpc_b200 import engine as E
    from gpmpc_b200._lib import call, ptr

    p = o.synthetic_problem(n, d, 1, p=3, seed=7 * n + d)
    X, Y = p["X"], p["Y"]
    g = torch.Generator().manual_seed(n)
    rn = torch.randn(1, generator=g, dtype=torch.float64) - 2.0
    rs = torch.randn((), generator=g, dtype=torch.float64)
    rl = torch.randn(1, d, generator=g, dtype=torch.float64) + 1.0
    noise, sf2, ell = o.constrained(rn, rs, rl)
    ell = ell.reshape(-1)
    Xd, Yd = X.to(_dev()), Y.to(_dev())
    loss, g_noise, g_sf2, g_ell, jit = E.small_loss_and_grad(Xd, Yd[:, 0], ell, float(sf2), float(noise))
    fit, ws = E.factorize(Xd, Yd[:, 0], ell, float(sf2), float(noise), want_kinv=True, want_image=False)
    l2, gn2, gs2, ge2 = E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
    assert jit == 0.0 and abs(loss - l2) < 1e-10 * abs(l2)
    assert abs(g_noise - gn2) < 1e-8 * max(abs(gn2), 1e-3) and abs(g_sf2 - gs2) < 1e-8 * max(abs(gs2), 1e-3)
    assert o.rel_err(torch.tensor(g_ell), torch.tensor(ge2)) < 1e-8
    lo, gno, gso, glo = o.loss_and_grad(X, Y[:, 0], rn, rs, rl)
    assert abs(loss - float(lo)) < 1e-6 * abs(float(lo))
    assert abs(g_noise * float(torch.sigmoid(rn)) - float(gno)) < 1e-6 * max(abs(float(gno)), 1e-3)
    assert o.rel_err(torch.tensor(g_ell).reshape(1, -1) * torch.sigmoid(rl), glo) < 1e-6
    # three GPs in one launch straight through the C ABI, with alpha
    lib = _lib().load()
    hyp = torch.stack([E.make_hyp(X.mean(0), ell, float(sf2), float(noise), d, _dev())] * 3)
    Yt = Yd.T.contiguous()
    out = torch.empty(3, lib.gpk_mll_small_out_stride(), dtype=torch.float64, device=_dev())
    alpha = torch.empty(3, n, dtype=torch.float64, device=_dev())
    Kd = torch.empty(3, n, n, dtype=torch.float64, device=_dev())
    Kinv = torch.empty(3, n, n, dtype=torch.float64, device=_dev())
    call("gpk_mll_small", ptr(Xd), Xd.stride(0), 0, ptr(Yt), Yt.stride(0), n, d, ptr(hyp), ptr(out), ptr(alpha), n,
         ptr(Kd), ptr(Kinv), n, n * n, 3, torch.cuda.current_stream().cuda_stream)
    K_ref = o.assemble_K(X, ell, sf2, noise)
    assert o.rel_err(Kd[2].cpu(), K_ref) < 1e-13 and o.rel_err(Kinv[1].cpu(), torch.linalg.inv(K_ref)) < 1e-7
    for k in range(3):
        st = o.fit_state(X, Y[:, k], ell, sf2, noise)
        assert o.rel_err(alpha[k].cpu(), st["alpha"]) < 1e-8
        assert abs(float(out[k, 1]) - float(st["quad"])) < 1e-9 * abs(float(st["quad"]))
        assert abs(float(out[k, 2]) - float(st["logdet"])) < 1e-9 * max(abs(float(st["logdet"])), 1.0)
    assert abs(float(out[0, 1]) - fit.quad) < 1e-10 * abs(fit.quad)
    # not positive definite -> info > 0 surfaces as NotPSDError after the jitter retries
    with pytest.raises(E.NotPSDError):
        E.small_loss_and_grad(Xd, Yd[:, 0], ell, float(sf2), -0.5)
