"""GPU parity tests of the tcgen05 int8 ("Ozaki") variance path: digit images, the exact int32 tensor-core
accumulators, and the recombined column sums of squares, each against an independent host computation."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

F64 = torch.float64


def _decode(img: torch.Tensor, tiles: int, kb_total: int, S: int, rows: int) -> torch.Tensor:
    """digit image [tile][kb][s][rows x 64 B core-matrix layout] -> int tensor (S, tiles*rows, kb_total*64)."""
    a = img.cpu().view(torch.int8).reshape(tiles, kb_total, S, rows // 8, 4, 8, 16)   # rg, kchunk, r%8, k%16
    a = a.permute(2, 0, 3, 5, 1, 4, 6).reshape(S, tiles * rows, kb_total * 64)
    return a.to(torch.int64)


def _digits_ref(x: torch.Tensor, S: int) -> torch.Tensor:
    """balanced base-256 digits (most significant first) of round(x * 2^(6+8(S-1))), |x| <= 1."""
    I = torch.round(x * float(2 ** (6 + 8 * (S - 1)))).to(torch.int64)
    out = []
    for _ in range(S - 1):
        low = I & 255
        d = torch.where(low >= 128, low - 256, low)
        out.append(d)
        I = (I - d) >> 8
    out.append(I)
    return torch.stack(out[::-1])


@pytest.fixture(scope="module")
def env():
    from gpmpc_b200 import _lib, engine as E
    from oracle import gp_oracle as o

    dev = _lib.require_cuda()
    return _lib, E, o, dev


@pytest.mark.parametrize("S", [4, 5, 6])
def test_w_image_digits(env, S):
    _lib, E, o, dev = env
    npad = 256
    g = torch.Generator().manual_seed(3)
    W = torch.tril(torch.randn(npad, npad, generator=g, dtype=F64) * torch.logspace(-3, 1, npad, dtype=F64)[:, None])
    Wd = W.to(dev)
    img = torch.zeros(_lib.load().gpk_oz_w_image_bytes(npad, S), dtype=torch.int8, device=dev)
    scale = torch.empty(npad, dtype=F64, device=dev)
    _lib.call("gpk_oz_slice_w", _lib.ptr(Wd), npad, npad, S, _lib.ptr(img), _lib.ptr(scale), _lib.stream_ptr(dev))
    sc = scale.cpu()
    m = W.abs().max(dim=1).values
    assert torch.all(sc > m) and torch.all(sc <= 2 * m), "row scale must be the next power of two above max|row|"
    assert torch.all(torch.log2(sc) == torch.round(torch.log2(sc)))
    got = _decode(img, npad // 64, npad // 64, S, 64)
    ref = _digits_ref(W / sc[:, None], S)
    # blocks above the diagonal block are never written / read
    mask = (torch.arange(npad)[None, :] // 64) <= ((torch.arange(npad)[:, None] // 64) | 1)
    assert torch.equal(got[:, mask], ref[:, mask])
    assert got.abs().max() <= 128 and got[0].abs().max() <= 64


@pytest.mark.parametrize("S,dtype", [(4, F64), (5, F64), (6, F64), (5, torch.float32)])
def test_e_image_reconstructs_kernel(env, S, dtype):
    _lib, E, o, dev = env
    n, d, q = 300, 16, 100
    p = o.synthetic_problem(n, d, q, p=1, seed=5)
    X, Z, ell = p["X"], p["Z"].to(dtype), p["ell"]
    npad, qpad = 384, 128
    hyp = E.make_hyp(X.mean(0), ell, p["sf2"], p["noise"], d, dev)
    img = torch.zeros(_lib.load().gpk_oz_e_image_bytes(qpad, npad, S), dtype=torch.int8, device=dev)
    Xd, Zd = X.to(dev), Z.to(dev)
    _lib.call("gpk_oz_cross", E.DTYPE_CODE[dtype], _lib.ptr(Zd), d, q, 0, _lib.ptr(Xd), d, n, 0, d, _lib.ptr(hyp),
              _lib.ptr(img), img.numel(), npad, qpad, S, 1, _lib.stream_ptr(dev))
    dig = _decode(img, qpad // 128, npad // 64, S, 128).to(F64)
    rec = sum(dig[s] * 2.0 ** (-6 - 8 * s) for s in range(S))
    ref = o.se_kernel(Z.to(F64), X, ell, 1.0)
    assert dig.abs().max() <= 128 and dig[0].abs().max() <= 64
    assert (rec[:q, :n] - ref).abs().max() < 2.0 ** (-6 - 8 * (S - 1)) + 2e-15
    assert rec[q:, :].abs().max() == 0 and rec[:, n:].abs().max() == 0


@pytest.mark.parametrize("S", [4, 5, 6])
def test_tensor_core_accumulators_exact(env, S):
    """raw int32 TMEM accumulators D_g = sum_{s+t=g} a_s b_t^T against an int64 host matmul of the decoded digits"""
    _lib, E, o, dev = env
    lib = _lib.load()
    npad, qpad = 384, 128
    g = torch.Generator().manual_seed(11)
    W = torch.tril(torch.randn(npad, npad, generator=g, dtype=F64))
    n, d, q = 380, 3, 120
    X = torch.rand(n, d, generator=g, dtype=F64)
    Z = torch.rand(q, d, generator=g, dtype=F64)
    hyp = E.make_hyp(X.mean(0), torch.full((d,), 0.7, dtype=F64), 1.3, 1e-2, d, dev)
    wimg = torch.zeros(lib.gpk_oz_w_image_bytes(npad, S), dtype=torch.int8, device=dev)
    eimg = torch.zeros(lib.gpk_oz_e_image_bytes(qpad, npad, S), dtype=torch.int8, device=dev)
    scale = torch.empty(npad, dtype=F64, device=dev)
    st = _lib.stream_ptr(dev)
    Wd, Xd, Zd = W.to(dev), X.to(dev), Z.to(dev)
    _lib.call("gpk_oz_slice_w", _lib.ptr(Wd), npad, npad, S, _lib.ptr(wimg), _lib.ptr(scale), st)
    _lib.call("gpk_oz_cross", 1, _lib.ptr(Zd), d, q, 0, _lib.ptr(Xd), d, n, 0, d, _lib.ptr(hyp), _lib.ptr(eimg),
              eimg.numel(), npad, qpad, S, 1, st)
    nib, qtiles = npad // 64, qpad // 128
    pieces = 3
    part = torch.zeros(2 * pieces, qpad, dtype=F64, device=dev)
    dbg = torch.full((qtiles, nib, S, 128, 64), -7, dtype=torch.int32, device=dev)
    _lib.call("gpk_oz_var_partials", _lib.ptr(eimg), 0, _lib.ptr(wimg), 0, _lib.ptr(scale), 0, _lib.ptr(hyp), npad, qpad, S,
              _lib.ptr(part), qpad, 0, pieces, 1, _lib.ptr(dbg), st)
    torch.cuda.synchronize()
    a = _decode(wimg, nib, npad // 64, S, 64)               # (S, npad, npad) digits of W
    b = _decode(eimg, qtiles, npad // 64, S, 128)           # (S, qpad, npad) digits of E
    a = a * ((torch.arange(npad)[None, :] // 64) <= ((torch.arange(npad)[:, None] // 64) | 1))   # unread blocks
    D = torch.zeros(S, qpad, npad, dtype=torch.int64)
    for s_ in range(S):
        for t in range(S - s_):
            D[s_ + t] += b[t] @ a[s_].T
    got = dbg.cpu().to(torch.int64).permute(2, 0, 3, 1, 4).reshape(S, qpad, npad)   # (S, qt, row, ib, col)
    assert torch.equal(got, D)
    # recombination: V = sf2 * scale_i / 4096 * sum_g 256^-g D_g ; part = per-piece sums over W rows of V^2
    R = sum(D[gi].to(F64) * 256.0 ** (-gi) for gi in range(S))
    V = 1.3 * scale.cpu()[None, :] / 4096.0 * R             # (qpad, npad)
    tot = part.cpu().sum(0)
    ref = (V * V).sum(1)
    assert (tot - ref).abs().max() <= 1e-13 * ref.abs().max()
    # and against plain fp64:  V = K_ZX W^T
    Vd = o.se_kernel(Z, X, torch.full((d,), 0.7, dtype=F64), 1.3) @ W[:, :n].T
    refd = (Vd * Vd).sum(1)
    tol = {4: 1e-5, 5: 1e-7, 6: 1e-10}[S]
    assert (tot[:q] - refd).abs().max() <= tol * refd.abs().max()


@pytest.mark.parametrize("dtype,want_mj", [(F64, True), (torch.float32, True), (F64, False)])
def test_pipelined_step_matches_sequential(env, dtype, want_mj, monkeypatch):
    """gpk_predict_i8's pipeline (front end of sub-chunk c+1 as co-resident 4-warp CTAs on the side stream while
    oz_var_kernel contracts sub-chunk c) against the one-chunk sequential step and the oracle; ragged last sub-chunk,
    3 GPs, sub-chunks of 1, 2 and 3 query tiles."""
    _lib, E, o, dev = env
    n, d, q, P = 700, 16, 1100, 3
    p = o.synthetic_problem(n, d, q, p=P, seed=11)
    gb = E.fit_batch(p["X"].to(dev), p["Y"].to(dev), p["ell"], p["sf2"], p["noise"], digits=5)
    Z = p["Z"].to(dev).to(dtype)
    monkeypatch.setenv("GPK_OZ_PIPELINE", "0")
    ref = E.predict(gb, Z, want_mean_jac=want_mj)
    torch.cuda.synchronize()
    tol = 1e-12 if dtype == F64 else 1e-5
    # sequential chunks that fit a small scratch budget (3 query tiles per chunk)
    small = _lib.load().gpk_predict_i8_work_bytes(3 * 128, gb.npad, d, 5, P)
    got = E.predict(gb, Z, want_mean_jac=want_mj, var_work_bytes=small)
    torch.cuda.synchronize()
    for a, b in zip(got, ref):
        assert (a is None and b is None) or o.rel_err(a.cpu().double(), b.cpu().double()) < tol
    for tsub in (1, 2, 3):
        monkeypatch.setenv("GPK_OZ_PIPELINE", "2")
        monkeypatch.setenv("GPK_OZ_PIPE_TSUB", str(tsub))
        for rep in range(2):                                   # twice: the side stream's events are reused
            got = E.predict(gb, Z, want_mean_jac=want_mj)
            torch.cuda.synchronize()
            for a, b in zip(got, ref):
                if b is None:
                    assert a is None
                    continue
                assert o.rel_err(a.cpu().double(), b.cpu().double()) < tol
    if want_mj and dtype == F64:
        st = o.fit_state(p["X"], p["Y"][:, 0], p["ell"], p["sf2"], p["noise"])
        m_ref, v_ref, j_ref = o.predict(p["Z"], p["X"], p["ell"], p["sf2"], st["L"], st["alpha"])
        assert o.rel_err(got[0][0].cpu(), m_ref) < 1e-6 and o.rel_err(got[1][0].cpu(), v_ref) < 1e-6
        assert o.rel_err(got[2][0].cpu(), j_ref) < 1e-6
