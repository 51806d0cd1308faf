"""GPU parity tests of the tcgen05 int8 digit GEMM (gpk_oz_dgemm_nt) and of the fit path that routes its large
GEMMs through it (run with -m gpu).  The GEMM is checked against a float64 matmul; the fit (Cholesky, W = L^-1,
alpha, MLL, MLL gradient) against the CPU oracle and against the DMMA-only path at BASELINE.json's 1e-6 tolerance.
"""
import pytest
import torch

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu

# |C_oz - C_ref| / max|C_ref| for random operands: the truncated digit products leave ~2^(-8 S + 4)
GEMM_TOL = {5: 2e-9, 6: 1e-11, 7: 1e-13}


def _dev():
    return torch.device("cuda", 0)


def _lib():
    from gpmpc_b200 import _lib as L

    return L


@pytest.mark.parametrize("digits", [5, 6, 7])
@pytest.mark.parametrize("M,N,K,a_tri,c_lower,same,alpha,beta", [
    (128, 128, 64, 0, 0, False, 1.0, 0.0),          # one tile, one k-block
    (256, 384, 1024, 0, 0, False, -1.0, 1.0),
    (384, 128, 384, 1, 0, False, -1.0, 0.0),        # lower-triangular A: k-range grows with the tile row
    (512, 256, 512, 2, 0, False, 1.0, 0.0),         # upper-triangular A
    (640, 640, 4160, 0, 1, True, -1.0, 1.0),        # SYRK-like: shared image, two scale chunks (K > 4096)
    (4352, 4352, 4352, 2, 1, True, 1.0, 0.0),       # K^-1 = U U^T shape, > one super-tile, chunk boundary inside a tile row
    (1280, 1152, 8256, 0, 0, False, 0.5, -2.0),     # three chunks, general alpha / beta
])
def test_oz_dgemm_nt(digits, M, N, K, a_tri, c_lower, same, alpha, beta):
    L = _lib()
    g = torch.Generator().manual_seed(M + N + K + digits)
    A = torch.randn(M, K, generator=g, dtype=torch.float64)
    # rows that decay along k like the rows of a Cholesky factor (exercises the per-chunk scales)
    A = A * torch.logspace(0, -5, K, dtype=torch.float64)[None, :] * torch.logspace(0, 3, M, dtype=torch.float64)[:, None]
    if a_tri == 1:
        A = torch.tril(A)
    if a_tri == 2:
        A = torch.triu(A)
    B = A if same else torch.randn(N, K, generator=g, dtype=torch.float64)
    C0 = torch.randn(M, N, generator=g, dtype=torch.float64)
    Ad = A.to(_dev())
    Bd = Ad if same else B.to(_dev())
    Cd = C0.to(_dev())
    wb = L.load().gpk_oz_dgemm_nt_work_bytes(M, N, K, int(same), digits)
    guard = 4096
    work = torch.full((wb + guard,), 0x5A, dtype=torch.uint8, device=_dev())
    L.call("gpk_oz_dgemm_nt", L.ptr(Ad), K, L.ptr(Bd), K, L.ptr(Cd), N, M, N, K, alpha, beta, a_tri, c_lower, digits,
           L.ptr(work), wb, L.stream_ptr())
    torch.cuda.synchronize()
    assert bool((work[wb:] == 0x5A).all()), "scratch overrun"
    prod = A @ B.T
    ref = beta * C0 + alpha * prod
    out = Cd.cpu()
    if c_lower:
        # tiles (128 x 128) strictly above the diagonal are untouched, the rest of the lower tiles is computed
        tile_low = (torch.arange(M)[:, None] // 128) >= (torch.arange(N)[None, :] // 128)
        assert torch.equal(out[~tile_low], C0[~tile_low])
        out, ref = torch.where(tile_low, out, 0.0), torch.where(tile_low, ref, 0.0)
    # error model: relative to the row scale of A times the row scale of B (random B: O(1)), i.e. per ROW of C
    row_mag = prod.abs().amax(dim=1, keepdim=True).clamp_min(1e-300)
    err = ((out - ref).abs() / (row_mag * abs(alpha) + (beta * C0).abs().amax())).max().item()
    assert err < GEMM_TOL[digits], f"digits={digits}: {err:.2e}"


def test_oz_dgemm_nt_rejects_bad_arguments():
    L = _lib()
    A = torch.zeros(128, 64, dtype=torch.float64, device=_dev())
    C = torch.zeros(128, 128, dtype=torch.float64, device=_dev())
    work = torch.zeros(1 << 20, dtype=torch.uint8, device=_dev())
    args = (L.ptr(A), 64, L.ptr(A), 64, L.ptr(C), 128, 128, 128, 64, 1.0, 0.0, 0, 0)
    with pytest.raises(L.GpkError):
        L.call("gpk_oz_dgemm_nt", *args, 4, L.ptr(work), work.numel(), L.stream_ptr())       # unsupported digit count
    with pytest.raises(L.GpkError):
        L.call("gpk_oz_dgemm_nt", *args, 6, L.ptr(work), 16, L.stream_ptr())                 # scratch too small
    with pytest.raises(L.GpkError):                                                          # a_tri needs M == K
        L.call("gpk_oz_dgemm_nt", L.ptr(A), 64, L.ptr(A), 64, L.ptr(C), 128, 128, 128, 64, 1.0, 0.0, 1, 0, 6,
               L.ptr(work), work.numel(), L.stream_ptr())


@pytest.fixture
def fit_gemm_settings():
    from gpmpc_b200 import engine as E

    saved = (E.FIT_GEMM_DIGITS, E.FIT_GEMM_MIN_NPAD, E.FIT_GEMM_MIN_WORK_G)
    yield E
    E.FIT_GEMM_DIGITS, E.FIT_GEMM_MIN_NPAD, E.FIT_GEMM_MIN_WORK_G = saved
    E.release_scratch()


@pytest.mark.parametrize("digits", [6, 7])
def test_fit_through_int8_gemms_matches_oracle(fit_gemm_settings, digits):
    """N = 2304 with the int8 path forced onto every GEMM of at least 1e9 multiply-adds: factor, inverse factor,
    alpha, MLL and the MLL gradient against the oracle (float64 LAPACK + closed-form gradient)."""
    E = fit_gemm_settings
    n, d = 2304, 16
    p = o.synthetic_problem(n, d, 1, seed=11)
    X, y, ell, sf2, noise = p["X"], p["Y"][:, 0], p["ell"], p["sf2"], p["noise"]
    E.FIT_GEMM_DIGITS, E.FIT_GEMM_MIN_NPAD, E.FIT_GEMM_MIN_WORK_G = digits, 1024, 1
    E.release_scratch()
    launches0 = _lib().load().gpk_launch_count()
    fit, ws = E.factorize(X.to(_dev()), y.to(_dev()), ell, sf2, noise, want_kinv=True)
    assert ws.gemm_work is not None and ws.gemm_digits == digits
    assert _lib().load().gpk_launch_count() > launches0
    st = o.fit_state(X, y, ell, sf2, noise)
    assert o.rel_err(torch.tril(ws.L[:n, :n]).cpu(), st["L"]) < 1e-9
    Wref = torch.linalg.solve_triangular(st["L"], torch.eye(n, dtype=torch.float64), upper=False)
    assert o.rel_err(fit.W[:n, :n].cpu(), Wref) < 1e-7
    assert o.rel_err(fit.alpha[:n].cpu(), st["alpha"]) < 1e-7
    mll_ref = float(o.mll_per_n(X, y, ell, sf2, noise))
    assert abs(fit.mll_per_n() - mll_ref) < 1e-8 * abs(mll_ref)          # BASELINE.json asks for 1e-6
    assert o.rel_err(ws.Kinv[:n, :n].cpu(), torch.cholesky_inverse(st["L"])) < 1e-7
    loss, g_noise, g_sf2, g_ell = E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
    # oracle gradient w.r.t. the constrained parameters through its raw-parameter chain rule
    raw = [o.inv_softplus(torch.tensor(v, dtype=torch.float64)) for v in (noise - 1e-6, sf2)]
    raw_ell = o.inv_softplus(torch.as_tensor(ell, dtype=torch.float64).reshape(-1))
    loss_ref, gn_ref, gs_ref, ge_ref = o.loss_and_grad(X, y, raw[0], raw[1], raw_ell)
    assert abs(loss - float(loss_ref)) < 1e-8 * abs(float(loss_ref))
    got = torch.cat([torch.tensor([g_noise * float(torch.sigmoid(raw[0])), g_sf2 * float(torch.sigmoid(raw[1]))],
                                  dtype=torch.float64),
                     torch.tensor(g_ell, dtype=torch.float64) * torch.sigmoid(raw_ell)])
    ref = torch.cat([gn_ref.reshape(1), gs_ref.reshape(1), ge_ref.reshape(-1)])
    assert o.rel_err(got, ref) < 1e-6


def test_fit_int8_gemms_agree_with_dmma_path(fit_gemm_settings):
    """Same fit at N = 4096 with the large GEMMs on the int8 pipe (default settings) and on the DMMA pipe only."""
    E = fit_gemm_settings
    n, d = 4096, 16
    p = o.synthetic_problem(n, d, 1, seed=5)
    X, y, ell, sf2, noise = p["X"].to(_dev()), p["Y"][:, 0].to(_dev()), p["ell"], p["sf2"], p["noise"]
    res = {}
    for digits in (0, E.FIT_GEMM_DIGITS or 6):
        E.FIT_GEMM_DIGITS = digits
        E.release_scratch()
        fit, ws = E.factorize(X, y, ell, sf2, noise, want_kinv=True, want_image=False)
        assert (ws.gemm_work is None) == (digits == 0)
        part = E.grad_partials(fit, ws).clone()
        res[digits] = (fit.mll_per_n(), fit.alpha.clone(), fit.W.clone(), part)
    (m0, a0, w0, g0), (m1, a1, w1, g1) = res.values()
    assert abs(m1 - m0) < 1e-9 * abs(m0)
    assert o.rel_err(a1.cpu(), a0.cpu()) < 1e-7
    assert o.rel_err(w1.cpu(), w0.cpu()) < 1e-7
    assert o.rel_err(g1.cpu(), g0.cpu()) < 1e-7


@pytest.mark.parametrize("digits", [0, 6])
def test_kinv_rows_match_full_inverse(fit_gemm_settings, digits):
    """gpk_kinv_rows (row-split gradient) against the rows of the full K^-1, on the DMMA and on the int8 path (where
    the window of U shares one digit image with U)."""
    E = fit_gemm_settings
    n, d = 4096, 16
    p = o.synthetic_problem(n, d, 1, seed=3)
    X, y = p["X"].to(_dev()), p["Y"][:, 0].to(_dev())
    E.FIT_GEMM_DIGITS, E.FIT_GEMM_MIN_NPAD, E.FIT_GEMM_MIN_WORK_G = digits, 1024, 1
    E.release_scratch()
    fit, ws = E.factorize(X, y, p["ell"], p["sf2"], p["noise"], want_kinv=True, want_image=False)
    full = ws.Kinv.clone()
    gfull = E.grad_partials(fit, ws).clone()
    acc = torch.zeros_like(gfull)
    for r0, r1 in ((0, 1000), (1000, 2304), (2304, 4096)):      # unaligned cuts: the 128-hull is formed
        ws.Kinv.fill_(float("nan"))
        fit2, ws2 = E.factorize(X, y, p["ell"], p["sf2"], p["noise"], want_kinv=(r0, r1), want_image=False)
        k0, k1 = E.kinv_row_hull(r0, r1, fit2.npad)
        assert E.FIT_SYMMETRIC_TRACE            # default: only columns [0, k1) of the rows are formed and read
        assert o.rel_err(ws2.Kinv[k0:k1, :k1].cpu(), full[k0:k1, :k1].cpu()) < (1e-12 if digits == 0 else 1e-8)
        assert bool(torch.isnan(ws2.Kinv[:k0]).all()) and bool(torch.isnan(ws2.Kinv[k1:]).all())
        assert bool(torch.isnan(ws2.Kinv[k0:k1, k1:]).all())
        acc += E.grad_partials(fit2, ws2, r0, r1)
    assert o.rel_err(acc.cpu(), gfull.cpu()) < 1e-9


@pytest.mark.parametrize("n", [300, 1000, 2304])
def test_overlapped_inverse_is_bit_identical(fit_gemm_settings, n):
    """gpk_potrf_trtri (inverse factor on a side stream beside the factorisation) against gpk_potrf + gpk_trtri on one
    stream: same kernels on the same data, so L, W = L^-1, U = W^T, alpha and the scalars must agree bit for bit --
    any difference is a missing dependency between the two streams.  Repeated to give a race a chance to show."""
    E = fit_gemm_settings
    p = o.synthetic_problem(n, 16, 1, seed=n)
    X, y = p["X"].to(_dev()), p["Y"][:, 0].to(_dev())
    saved = E.FIT_OVERLAP_INVERSE
    try:
        E.FIT_OVERLAP_INVERSE = False
        fit0, ws0 = E.factorize(X, y, p["ell"], p["sf2"], p["noise"], want_kinv=True)
        ref = [t.clone() for t in (ws0.L, fit0.W, ws0.U, fit0.alpha, ws0.Kinv)] + [fit0.quad, fit0.logdet]
        E.FIT_OVERLAP_INVERSE = True
        for _ in range(5):
            fit1, ws1 = E.factorize(X, y, p["ell"], p["sf2"], p["noise"], want_kinv=True)
            got = [ws1.L, fit1.W, ws1.U, fit1.alpha, ws1.Kinv]
            for a, b in zip(got, ref[:5]):
                assert torch.equal(a, b)
            assert (fit1.quad, fit1.logdet) == (ref[5], ref[6])
    finally:
        E.FIT_OVERLAP_INVERSE = saved


def test_symmetric_trace_equals_full_trace(fit_gemm_settings):
    """gpk_mll_grad_reduce_sym (lower triangle of K^-1, off-diagonal entries counted twice) against gpk_mll_grad_reduce
    (all entries): same d+2 sums, also when the rows are cut at unaligned places; the upper triangle is never read."""
    E = fit_gemm_settings
    n, d = 1000, 5
    p = o.synthetic_problem(n, d, 1, seed=9)
    ell = p["ell"] * torch.linspace(0.7, 1.4, d, dtype=torch.float64)
    fit, ws = E.factorize(p["X"].to(_dev()), p["Y"][:, 0].to(_dev()), ell, 1.3, 0.02, want_kinv=True)
    saved = E.FIT_SYMMETRIC_TRACE
    try:
        E.FIT_SYMMETRIC_TRACE = False
        full = E.grad_partials(fit, ws).clone()
        E.FIT_SYMMETRIC_TRACE = True
        ws.Kinv.copy_(torch.tril(ws.Kinv) + torch.triu(torch.full_like(ws.Kinv, float("nan")), diagonal=1))
        sym = E.grad_partials(fit, ws).clone()
        assert o.rel_err(sym.cpu(), full.cpu()) < 1e-12
        acc = torch.zeros_like(full)
        for r0, r1 in ((0, 1), (1, 333), (333, 334), (334, 1000)):
            acc += E.grad_partials(fit, ws, r0, r1)
        assert o.rel_err(acc.cpu(), full.cpu()) < 1e-12
    finally:
        E.FIT_SYMMETRIC_TRACE = saved
