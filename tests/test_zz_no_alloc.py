"""gpk_mll_grad_reduce_static: the allocation-free variant of the MLL gradient trace (its per-CTA partial sums live in a
file-scope device array of the library) against the existing entry points gpk_mll_grad_reduce / gpk_mll_grad_reduce_sym
called with a caller-supplied workspace.  Called through the C ABI (ctypes).  No counterpart in the reference: the
entry was added for callers that cannot call cudaMalloc.  Collected LAST on purpose (pytest -x).  Run with -m gpu."""
import pytest
import torch

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
TOL = 1e-12     # the tolerance of tests/test_gpu_ozgemm.py::test_symmetric_trace_equals_full_trace


def _fit(n, d, seed):
    from gpmpc_b200 import engine as E

    p = o.synthetic_problem(n, d, 1, seed=seed)
    ell = p["ell"] * torch.linspace(0.7, 1.4, d, dtype=torch.float64)
    fit, ws = E.factorize(p["X"].to(DEV), p["Y"][:, 0].to(DEV), ell, 1.3, 0.02, want_kinv=True)
    # private copies: the engine's scratch (ws) is reused by the next factorize
    return fit, fit.X.clone(), fit.hyp.clone(), ws.Kinv.clone(), fit.alpha.clone()


def _existing(lib, sym, X, n, d, hyp, Kinv, alpha, r0, r1):
    from gpmpc_b200 import _lib as L

    elems = lib.gpk_grad_trace_work_elems(n, d, r0, r1)
    work = torch.full((elems,), float("nan"), dtype=torch.float64, device=DEV)
    out = torch.zeros(lib.gpk_pad_dim(d) + 2, dtype=torch.float64, device=DEV)
    L.call("gpk_mll_grad_reduce_sym" if sym else "gpk_mll_grad_reduce", L.ptr(X), X.stride(0), n, d, L.ptr(hyp),
           L.ptr(Kinv), Kinv.stride(0), L.ptr(alpha), r0, r1, L.ptr(work), elems, L.ptr(out), L.stream_ptr(DEV))
    return out


def _static(sym, X, n, d, hyp, Kinv, alpha, r0, r1, out):
    from gpmpc_b200 import _lib as L

    L.call("gpk_mll_grad_reduce_static", L.ptr(X), X.stride(0), n, d, L.ptr(hyp), L.ptr(Kinv), Kinv.stride(0),
           L.ptr(alpha), r0, r1, int(sym), L.ptr(out), L.stream_ptr(DEV))
    return out


def test_static_scratch_trace_matches_workspace_trace_twice_in_a_row():
    from gpmpc_b200 import _lib as L

    lib = L.load()
    assert lib.gpk_mll_grad_reduce_static_max_elems() == 4096 * 34
    big = _fit(1000, 5, seed=9)              # 16 x 16 CTAs x 10 partials
    small = _fit(300, 3, seed=4)             # 5 x 5 CTAs x 5 partials: leaves most of the first call's partials behind
    outs = [torch.full((34,), float("nan"), dtype=torch.float64, device=DEV) for _ in range(4)]
    refs, got = [], []
    # two calls in a row per form, the second on a different GP and row range, outputs preallocated: between the first
    # static call and the last nothing is allocated by this test either
    cases = [(big, 0, 1000, 1), (small, 37, 290, 1), (big, 333, 1000, 0), (small, 0, 300, 0)]
    for (fit, X, hyp, Kinv, alpha), r0, r1, sym in cases:
        refs.append(_existing(lib, sym, X, fit.n, fit.d, hyp, Kinv, alpha, r0, r1))
    torch.cuda.synchronize(DEV)
    free_before = None
    for k, ((fit, X, hyp, Kinv, alpha), r0, r1, sym) in enumerate(cases):
        got.append(_static(sym, X, fit.n, fit.d, hyp, Kinv, alpha, r0, r1, outs[k]))
        torch.cuda.synchronize(DEV)
        if k == 0:
            free_before = torch.cuda.mem_get_info(DEV)[0]    # after the first call: symbol lookup and module load done
    assert torch.cuda.mem_get_info(DEV)[0] >= free_before, "device memory was allocated by a static-scratch call"
    for k, ((fit, *_), r0, r1, sym) in enumerate(cases):
        w = lib.gpk_pad_dim(fit.d) + 2
        a, b = got[k][:w].cpu(), refs[k][:w].cpu()
        assert torch.isfinite(a).all()
        assert o.rel_err(a, b) < TOL, (k, a, b)
        assert torch.equal(a, b), (k, a, b)              # same kernels, same order of summation: identical bits
    # and the sums are the right ones: the symmetric and the full form agree on the whole matrix
    full = _static(0, big[1], 1000, 5, big[2], big[3], big[4], 0, 1000, outs[2])
    assert o.rel_err(full[:10].cpu(), refs[0][:10].cpu()) < TOL


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from gpmpc_b200 import _lib as L

    lib = L.load()
    n, d = 4160, 32                          # 65 x 65 CTAs x 34 partials > 4096 x 34
    assert lib.gpk_grad_trace_work_elems(n, d, 0, n) > 0
    X = torch.zeros(n, d, dtype=torch.float64, device=DEV)
    hyp = torch.zeros(lib.gpk_hyp_stride(), dtype=torch.float64, device=DEV)
    Kinv = torch.zeros(8, dtype=torch.float64, device=DEV)       # never read: the call must refuse before launching
    alpha = torch.zeros(n, dtype=torch.float64, device=DEV)
    out = torch.full((34,), 7.0, dtype=torch.float64, device=DEV)
    launches = lib.gpk_launch_count()
    with pytest.raises(L.GpkError, match="built-in scratch holds"):
        _static(1, X, n, d, hyp, Kinv, alpha, 0, n, out)
    assert lib.gpk_launch_count() == launches
    torch.cuda.synchronize(DEV)
    assert bool((out == 7.0).all())
    # a row slice of the same GP that fits is accepted: 63 row tiles x 65 column tiles = 4095 CTAs x 34 partials
    Kfull = torch.eye(n, dtype=torch.float64, device=DEV)
    hyp[lib.gpk_hyp_stride() - 8] = 1.0      # sf2
    hyp[32:64] = 1.0                         # 1 / ell
    ok = _static(1, X, n, d, hyp, Kfull, alpha, 0, 63 * 64, out)
    torch.cuda.synchronize(DEV)
    assert abs(float(ok[0]) + 63 * 64) < 1e-9        # tr(alpha alpha^T - I) over 4032 rows with alpha = 0
