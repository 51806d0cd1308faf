"""Device-side engine of the GP hot path: owns the HBM buffers and drives the C ABI (include/gpmpc_b200.h).

Everything numeric happens in libgpmpc_b200.so (hand-written sm_100a CUDA); torch is used for device
memory, streams and the tiny host<->device scalar traffic only.  There is no CPU path.

HBM layout (row-major, float64 unless noted; npad = N rounded up to 128, identity on the padding):
  X      (N, d)          training inputs, raw (centred / scaled on the fly inside the kernels)
  hyp    (72,)           centre[32] | 1/ell[32] | sf2 | noise | pad      (per GP)
  L      (npad, npad)    K = K_f + noise I, overwritten by its lower Cholesky factor   (fit scratch)
  W, U   (npad, npad)    W = L^-1 (lower, kept for the variance), U = W^T (fit scratch)
  Kinv   (npad, npad)    K^-1 = U U^T, only when the MLL gradient or gp.K_inv is wanted (fit scratch)
  alpha  (npad,)         K^-1 y, zero on the padding
  pack   tiles of 256 training points [x~ | -1/2|x~|^2 | sf2 alpha] in the prediction dtype
"""

from __future__ import annotations

import functools
import math
from dataclasses import dataclass, field

import torch

from . import _lib
from ._lib import call, ptr

TILE = 128
LOG_2PI = math.log(2.0 * math.pi)
F64 = torch.float64
DTYPE_CODE = {torch.float32: 0, torch.float64: 1}


def _on_arg_device(fn):
    """Run `fn` with the CUDA device of its first device-carrying argument current (kernels launch on the current
    device; cudaFuncSetAttribute, the side stream and the registered scratch are per device).  A no-op in the usual
    one-process-per-GPU setting where the caller already selected the device."""

    @functools.wraps(fn)
    def wrapper(*args, **kw):
        dev = None
        for a in args:
            d = getattr(a, "device", None)
            if isinstance(d, torch.device) and d.type == "cuda":
                dev = d
                break
        if dev is None or dev.index is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kw)
        with torch.cuda.device(dev):
            return fn(*args, **kw)

    return wrapper


_initialised: set = set()


def init(device=None) -> torch.device:
    """Create the library's per-device state up front (gpk_init: exp table, side stream, event pool) so that no later
    compute entry allocates or synchronises for it; called by factorize() / predict() on first use of a device."""
    dev = _lib.require_cuda(device)
    if dev.index not in _initialised:
        with torch.cuda.device(dev):
            call("gpk_init", _lib.stream_ptr(dev))
        _initialised.add(dev.index)
    return dev


def shutdown():
    """Free the per-device library state and the engine's scratch caches (gpk_shutdown per initialised device)."""
    release_scratch()
    for idx in list(_initialised):
        with torch.cuda.device(idx):
            call("gpk_shutdown")
    _initialised.clear()


class NotPSDError(RuntimeError):
    """K + noise I was not positive definite even after the jitter retries (gpytorch's NotPSDError)."""


def round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


def make_hyp(center, ell, sf2: float, noise: float, d: int, device) -> torch.Tensor:
    """Hyper-parameter block for the kernels (layout in include/gpmpc_b200.h), built on the host."""
    stride = _lib.load().gpk_hyp_stride()
    h = torch.zeros(stride, dtype=F64)
    ell = torch.as_tensor(ell, dtype=F64).reshape(-1).cpu()
    if ell.numel() == 1:
        ell = ell.expand(d)
    assert ell.numel() == d and d <= 32, "lengthscale must have 1 or d entries, d <= 32"
    h[:d] = torch.as_tensor(center, dtype=F64).reshape(-1).cpu()
    h[32:64] = 1.0
    h[32 : 32 + d] = 1.0 / ell
    h[64] = float(sf2)
    h[65] = float(noise)
    return h.to(device, non_blocking=True)


# Large fit GEMMs (potrf / trtri / K^-1 at npad >= FIT_GEMM_MIN_NPAD) run on the tcgen05 int8 pipe with this many
# balanced base-256 digits per operand (csrc/ozgemm.cu): 7 digits reproduce the DMMA result to ~2e-15 of max|C|,
# 6 to ~5e-13, 5 to ~1e-10; 0 keeps every GEMM on the FP64 DMMA pipe.
FIT_GEMM_DIGITS = 6
FIT_GEMM_MIN_NPAD = 2048
# K^-1 = U U^T only feeds the MLL gradient (and gp.K_inv): one digit less there costs ~1e-9 on the gradient (nothing on
# the MLL, alpha or the predictions) and removes 6 of its 21 digit products.  0 = same digits as the factorisation.
FIT_GEMM_DIGITS_KINV = 5
FIT_SYMMETRIC_TRACE = True  # gradient trace from the lower triangle of K^-1 only (gpk_mll_grad_reduce_sym)
FIT_OVERLAP_INVERSE = True  # W = L^-1 on a side stream beside the Cholesky recursion (gpk_potrf_trtri)
FIT_GEMM_MIN_WORK_G = 0     # tuning hook: smallest GEMM (1e9 multiply-adds) sent to the int8 path; 0 = library default


def _register_fit_gemm_workspace(buf: torch.Tensor | None, digits: int = 0):
    """(Un)register the per-device scratch of the int8 fit GEMMs with the library (it only borrows the pointer)."""
    if buf is None:
        call("gpk_oz_gemm_set_workspace", None, 0, 0, 0)
    else:
        call("gpk_oz_gemm_set_workspace", ptr(buf), buf.numel(), digits, -int(FIT_GEMM_MIN_WORK_G))


@dataclass
class FitScratch:
    """Scratch for fitting one GP of padded size npad; reused across GPs and MLL iterations."""

    npad: int
    device: torch.device
    L: torch.Tensor = None
    U: torch.Tensor = None
    T: torch.Tensor = None
    dinv: torch.Tensor = None
    t_work: torch.Tensor = None
    scal: torch.Tensor = None
    info: torch.Tensor = None
    Kinv: torch.Tensor = None
    grad_work: torch.Tensor = None
    grad_out: torch.Tensor = None
    y_pad: torch.Tensor = None
    gemm_work: torch.Tensor = None     # digit images of the int8 fit GEMMs (None: DMMA only)
    gemm_digits: int = 0

    def __post_init__(self):
        lib = _lib.load()
        n, dev = self.npad, self.device
        self.L = torch.empty(n, n, dtype=F64, device=dev)
        self.U = torch.empty(n, n, dtype=F64, device=dev)
        self.T = torch.empty(max(1, lib.gpk_trtri_scratch_elems(n)), dtype=F64, device=dev)
        self.dinv = torch.empty(n * TILE, dtype=F64, device=dev)
        self.t_work = torch.empty(n, dtype=F64, device=dev)
        self.scal = torch.zeros(2, dtype=F64, device=dev)
        self.info = torch.zeros(1, dtype=torch.int32, device=dev)
        self.grad_out = torch.zeros(34, dtype=F64, device=dev)
        self.y_pad = torch.zeros(n, dtype=F64, device=dev)
        if FIT_GEMM_DIGITS and n >= FIT_GEMM_MIN_NPAD:
            self.gemm_digits = int(FIT_GEMM_DIGITS)
            self.gemm_work = torch.empty(lib.gpk_oz_gemm_work_bytes(n, self.gemm_digits), dtype=torch.uint8, device=dev)

    def activate(self):
        """Point the library's GEMM dispatcher at this scratch (or at nothing) on this device."""
        with torch.cuda.device(self.device):
            _register_fit_gemm_workspace(self.gemm_work, self.gemm_digits)

    def need_kinv(self):
        if self.Kinv is None:
            self.Kinv = torch.empty(self.npad, self.npad, dtype=F64, device=self.device)
        return self.Kinv


_scratch_cache: dict = {}


SCRATCH_KEEP_BYTES = 2 << 30     # several fit-scratch sizes stay cached per device while they total less than this


def _scratch_bytes(npad: int) -> int:
    return 4 * npad * npad * 8       # L, U, K^-1, T/gemm images: order of magnitude


def get_fit_scratch(npad: int, device: torch.device) -> FitScratch:
    """Fit scratch of one padded size on one device, reused across GPs and MLL iterations.  Small sizes are kept side
    by side (a FITC set-up alternates between M inducing points and N training points); a size is evicted (oldest
    first) only when the cached scratch of the device would exceed SCRATCH_KEEP_BYTES -- the big configs need the
    memory back.  Single-stream contract: the scratch is keyed by device, not by stream (include/gpmpc_b200.h)."""
    key = (npad, device.index, int(FIT_GEMM_DIGITS), int(FIT_GEMM_MIN_NPAD))
    ws = _scratch_cache.pop(key, None)
    if ws is None:
        budget = SCRATCH_KEEP_BYTES - _scratch_bytes(npad)
        mine = [k for k in _scratch_cache if k[1] == device.index]
        while mine and sum(_scratch_bytes(k[0]) for k in mine) > max(budget, 0):
            del _scratch_cache[mine.pop(0)]          # dict order = least recently used first
        ws = FitScratch(npad, device)
    _scratch_cache[key] = ws                          # most recently used last
    ws.activate()                                     # the library's GEMM dispatcher follows the scratch in use
    return ws


def release_scratch():
    if torch.cuda.is_available():
        for k in _scratch_cache:
            with torch.cuda.device(k[1]):
                _register_fit_gemm_workspace(None)
    _scratch_cache.clear()
    _work_cache.clear()


_work_cache: dict = {}


def get_work(name: str, nbytes: int, device: torch.device) -> torch.Tensor:
    """Grow-only byte scratch per (name, device)."""
    key = (name, device.index)
    t = _work_cache.get(key)
    if t is None or t.numel() < nbytes:
        _work_cache[key] = t = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
    return t


@dataclass
class GPFit:
    """Fitted state of one exact GP at fixed hyper-parameters (what gpytorch's prediction strategy caches)."""

    X: torch.Tensor            # (N, d) float64, device
    y: torch.Tensor            # (N,)
    n: int
    d: int
    npad: int
    center: torch.Tensor       # (d,) host float64
    ell: torch.Tensor          # (d,) host float64
    sf2: float
    noise: float
    hyp: torch.Tensor          # (72,) device
    W: torch.Tensor            # (npad, npad) L^-1
    alpha: torch.Tensor        # (npad,)
    wimg: torch.Tensor = None  # (digits * npad * npad,) int8 digit image of W for the tcgen05 variance path
    wscale: torch.Tensor = None  # (npad,) power-of-two row scales of the image
    digits: int = 0
    digits_f32: int = 0        # leading digits multiplied for float32 queries (0: all of them)
    var_err: dict = field(default_factory=dict)   # probed |var_i8 - var_f64| per digit count (select_digits)
    quad: float = float("nan")
    logdet: float = float("nan")
    jitter: float = 0.0
    packs: dict = field(default_factory=dict)

    @property
    def device(self):
        return self.X.device

    def mll_per_n(self) -> float:
        """log p(y) / N  (what ExactMarginalLogLikelihood returns, gpmpc/gp.py:57,62)."""
        return -0.5 * (self.quad + self.logdet + self.n * LOG_2PI) / self.n

    def pack(self, dtype: torch.dtype) -> torch.Tensor:
        """Resident training pack for the fused mean/Jacobian kernel in the prediction dtype."""
        if dtype not in self.packs:
            lib = _lib.load()
            elems = lib.gpk_pack_train_elems(self.n, self.d)
            p = torch.empty(elems, dtype=dtype, device=self.device)
            call("gpk_pack_train", DTYPE_CODE[dtype], ptr(self.X), self.X.stride(0), self.n, self.d, ptr(self.alpha),
                 ptr(self.hyp), ptr(p), elems, _lib.stream_ptr(self.device))
            self.packs[dtype] = p
        return self.packs[dtype]


def _as_device_f64(t: torch.Tensor, device) -> torch.Tensor:
    return t.detach().to(device=device, dtype=F64).contiguous()


def i8_supported(npad: int, digits: int | None = None) -> bool:
    """The int32 accumulators of the tcgen05 variance path provably cannot overflow up to gpk_oz_max_npad(digits)
    training points (32768 at the default 5 digits: the leading digits are <= 64 and E > 0, see csrc/ozaki.cu); larger
    training sets use the DMMA path (var_path="f64")."""
    lib = _lib.load()
    digits = lib.gpk_oz_default_digits() if digits is None else digits
    return npad <= lib.gpk_oz_max_npad(digits)


@_on_arg_device
def slice_w(W: torch.Tensor, npad: int, digits: int | None = None, wimg_out=None, wscale_out=None):
    """int8 digit image + power-of-two row scales of W = L^-1 (operand of the tcgen05 variance contraction)."""
    lib = _lib.load()
    digits = lib.gpk_oz_default_digits() if digits is None else digits
    dev = W.device
    nbytes = lib.gpk_oz_w_image_bytes(npad, digits)
    assert nbytes > 0, "digits must be 4, 5 or 6"
    wimg = wimg_out if wimg_out is not None else torch.empty(nbytes, dtype=torch.int8, device=dev)
    wscale = wscale_out if wscale_out is not None else torch.empty(npad, dtype=F64, device=dev)
    call("gpk_oz_slice_w", ptr(W), W.stride(0), npad, digits, ptr(wimg), ptr(wscale), _lib.stream_ptr(dev))
    return wimg, wscale, digits


# Accuracy policy of the int8 variance path.  The S-digit images quantise W = L^-1 (per row, relative to the row
# maximum) and E = exp(-r^2/2) (absolute) to 8S-2 bits and drop the digit products below 2^-8S, so the ABSOLUTE error
# of var_f grows with max|W| ~ 1/sqrt(noise) and sqrt(N) while the quantity the reference consumes, the predictive
# variance var_f + noise (gpmpc/gpmpc.py:444-445), can be as small as the noise.  The digit count is therefore chosen
# PER FIT by measurement: the int8 path is compared with the FP64 DMMA path on a probe set (training inputs, where
# ||W k|| and with it the error is largest, plus perturbed copies) and S is the smallest count whose probed error,
# times a safety factor, stays below rtol * noise -- i.e. 1e-6 (float64 queries) / 1e-4 (float32 queries) ELEMENT-WISE
# relative on the predictive variance.  If no digit count qualifies (noise <= ~1e-4 with dense data) the fit carries no
# image and predict() takes the DMMA path (var_path="f64").
VAR_RTOL_F64 = 1e-6
VAR_RTOL_F32 = 1e-4
VAR_PROBE_SAFETY = 2.0
VAR_PROBE_POINTS = 256
# below this padded size (N <= 128: the reference's own shapes, N <= 45) the policy builds no image at all: one
# 128-tile of DMMA work is as fast as the tensor-pipe path and the probing would dominate the fit
VAR_IMAGE_MIN_NPAD = 256


def _probe_queries(fit: "GPFit") -> torch.Tensor:
    """Probe set of select_digits: <= 256 training inputs spread over the set + the same points shifted by a third
    of a lengthscale in every coordinate (deterministic)."""
    n = fit.n
    idx = torch.linspace(0, n - 1, min(n, VAR_PROBE_POINTS), device=fit.device).round().long()
    Xp = fit.X[idx]
    shift = (fit.ell.to(fit.device) / 3.0) * (1.0 - 2.0 * (torch.arange(fit.d, device=fit.device) % 2))
    return torch.cat([Xp, Xp + shift]).contiguous()


@_on_arg_device
def select_digits(fits: list, candidates=(5, 6), wimg_out=None, wscale_out=None) -> int:
    """Pick ONE digit count for the GPs in `fits` (same npad; a GPBatch shares its image format) by probing, build the
    images, and record the probed errors in fit.var_err.  Returns the digit count (0: no image, DMMA path)."""
    npad = fits[0].npad
    if npad < VAR_IMAGE_MIN_NPAD:
        return 0
    wimg_out = wimg_out or [None] * len(fits)
    wscale_out = wscale_out or [None] * len(fits)
    refs = []
    for f in fits:
        Zp = _probe_queries(f)
        saved = (f.wimg, f.wscale, f.digits)
        f.wimg = f.wscale = None
        f.digits = 0
        refs.append((Zp, predict(f, Zp, var_path="f64")[1][0]))
        f.wimg, f.wscale, f.digits = saved
    for S in candidates:
        if not i8_supported(npad, S):
            continue
        lib = _lib.load()
        ok = True
        errs = []
        for f, (Zp, v_ref), wo, so in zip(fits, refs, wimg_out, wscale_out):
            if wo is not None and wo.numel() != lib.gpk_oz_w_image_bytes(npad, S):
                wo = None                                         # caller's buffer was sized for another digit count
            f.wimg, f.wscale, f.digits = slice_w(f.W, npad, S, wo, so)
            f.digits_f32 = 0
            e = [(predict(f, Zp, want_mean_jac=False)[1][0] - v_ref).abs().max()]
            if S > 4:
                e.append((predict(f, Zp, want_mean_jac=False, mul_digits=4)[1][0] - v_ref).abs().max())
            errs.append(torch.stack(e))
        errs = torch.stack(errs).tolist()                          # one host round trip per candidate
        for f, e in zip(fits, errs):
            f.var_err[S] = e[0]
            if len(e) > 1:
                f.var_err[4] = e[1]
            ok = ok and VAR_PROBE_SAFETY * e[0] <= VAR_RTOL_F64 * f.noise
        if ok:
            f32_ok = all(len(e) > 1 and VAR_PROBE_SAFETY * e[1] <= VAR_RTOL_F32 * f.noise for f, e in zip(fits, errs))
            for f in fits:
                f.digits_f32 = 4 if f32_ok else 0
            return S
    for f in fits:
        f.wimg = f.wscale = None
        f.digits = f.digits_f32 = 0
    return 0


def kinv_row_hull(r0: int, r1: int, npad: int) -> tuple[int, int]:
    """128-aligned hull of the row range [r0, r1) inside [0, npad] (the granularity of gpk_kinv_rows)."""
    if r1 <= r0:
        return 0, 0
    return r0 // TILE * TILE, min(npad, round_up(r1, TILE))


@_on_arg_device
def factorize(X: torch.Tensor, y: torch.Tensor, ell, sf2: float, noise: float, *, center=None,
              want_kinv: bool = False, W_out: torch.Tensor | None = None, max_tries: int = 3,
              digits: int | None = None, wimg_out=None, wscale_out=None,
              want_image: bool = True) -> tuple[GPFit, FitScratch]:
    """Assemble K, Cholesky, W = L^-1, alpha, y^T K^-1 y and log det K on the GPU (float64).
    want_kinv: True also forms K^-1 (ws.Kinv); a (r0, r1) tuple forms only those rows of it (row-split gradient).

    Replaces the gpytorch work behind gpmpc/gp.py:61-62 (train mode) and the prediction-strategy cache
    behind gpmpc/gpmpc.py:444 (eval mode).  On a failed pivot the diagonal jitter policy of gpytorch's
    psd_safe_cholesky is applied (1e-8 for float64, x10 per retry, `max_tries` retries).
    """
    dev = init(X.device if X.is_cuda else None)
    X = _as_device_f64(X, dev)
    y = _as_device_f64(y, dev).reshape(-1)
    n, d = X.shape
    assert y.numel() == n, "x and y disagree on the number of training points"
    npad = round_up(n, TILE)
    ws = get_fit_scratch(npad, dev)
    st = _lib.stream_ptr(dev)
    ell_t = torch.as_tensor(ell, dtype=F64).reshape(-1).cpu()
    if ell_t.numel() == 1:
        ell_t = ell_t.expand(d).clone()
    if center is None:
        center = X.mean(dim=0)
    center_h = torch.as_tensor(center, dtype=F64).reshape(-1).cpu()
    W = W_out if W_out is not None else torch.empty(npad, npad, dtype=F64, device=dev)
    alpha = torch.empty(npad, dtype=F64, device=dev)
    ws.y_pad.zero_()
    ws.y_pad[:n] = y

    jitter = 0.0
    for attempt in range(max_tries + 1):
        hyp = make_hyp(center_h, ell_t, sf2, noise + jitter, d, dev)
        call("gpk_cross_kernel", 1, ptr(X), X.stride(0), n, 0, ptr(X), X.stride(0), n, 0, d, ptr(hyp), ptr(ws.L),
             npad, 0, npad, npad, 1, 1, st)
        # the rest of the chain is enqueued before the pivot check is read back (one host round trip per fit instead of
        # two; a failed factorisation is rare and only costs the wasted launches)
        if FIT_OVERLAP_INVERSE:
            call("gpk_potrf_trtri", ptr(ws.L), npad, npad, ptr(ws.dinv), ptr(ws.info), ptr(W), npad, ptr(ws.U), npad,
                 ptr(ws.T), ws.T.numel(), st)
        else:
            call("gpk_potrf", ptr(ws.L), npad, npad, ptr(ws.dinv), ptr(ws.info), st)
            call("gpk_trtri", ptr(ws.L), npad, npad, ptr(ws.dinv), ptr(W), npad, ptr(ws.U), npad, ptr(ws.T), ws.T.numel(), st)
        call("gpk_solve_alpha", ptr(W), npad, ptr(ws.U), npad, ptr(ws.L), npad, npad, ptr(ws.y_pad), ptr(ws.t_work),
             ptr(alpha), ptr(ws.scal), st)
        kinv_digits = int(FIT_GEMM_DIGITS_KINV) if ws.gemm_work is not None else 0
        switch_digits = bool(want_kinv) and 5 <= kinv_digits < ws.gemm_digits      # the scratch is sized for gemm_digits
        if switch_digits:
            _register_fit_gemm_workspace(ws.gemm_work, kinv_digits)
        if isinstance(want_kinv, tuple) and kinv_row_hull(want_kinv[0], want_kinv[1], npad) == (0, npad):
            want_kinv = True            # all rows: the symmetric product (lower tiles + mirror) is 3x cheaper
        if isinstance(want_kinv, tuple):
            # row-split gradient: only the (128-aligned hull of the) rows this rank's share of the trace reads
            k0, k1 = kinv_row_hull(want_kinv[0], want_kinv[1], npad)
            if k1 > k0:   # the symmetric trace reads columns <= row only
                call("gpk_kinv_rows", ptr(ws.U), npad, npad, k0, k1, k1 if FIT_SYMMETRIC_TRACE else npad,
                     ptr(ws.need_kinv()), npad, st)
        elif want_kinv:
            call("gpk_kinv", ptr(ws.U), npad, npad, ptr(ws.need_kinv()), npad, st)
        if switch_digits:
            _register_fit_gemm_workspace(ws.gemm_work, ws.gemm_digits)
        info = int(ws.info.item())          # the reference syncs here too (loss value / NotPSD check)
        if info == 0:
            break
        if attempt == max_tries:
            raise NotPSDError(f"K + noise I not positive definite after jitter {jitter:g} (potrf info={info})")
        jitter = 1e-8 if jitter == 0.0 else jitter * 10.0
    quad, logdet = ws.scal.tolist()
    fit = GPFit(X=X, y=y, n=n, d=d, npad=npad, center=center_h, ell=ell_t, sf2=float(sf2), noise=float(noise),
                hyp=make_hyp(center_h, ell_t, sf2, noise, d, dev) if jitter else hyp, W=W, alpha=alpha,
                quad=quad, logdet=logdet, jitter=jitter)
    if want_image:
        if digits is None:
            select_digits([fit], wimg_out=[wimg_out], wscale_out=[wscale_out])
        elif i8_supported(npad, digits):
            fit.wimg, fit.wscale, fit.digits = slice_w(W, npad, digits, wimg_out, wscale_out)
            fit.digits_f32 = min(fit.digits, 4)     # explicit digit count: the caller owns the accuracy decision
    return fit, ws


@_on_arg_device
def grad_partials(fit: GPFit, ws: FitScratch, r0: int = 0, r1: int | None = None) -> torch.Tensor:
    """(d+2) partial sums of the MLL gradient trace over rows [r0, r1) (device tensor, float64).
    Needs ws.Kinv from factorize(want_kinv=True) (or the rows [r0, r1) of it from want_kinv=(r0, r1)).  With
    FIT_SYMMETRIC_TRACE the partial of a row block counts its columns b < a twice and skips b > a: partials of a
    partition of the rows still add up to the full sums, a single block's partial is not the full-row sum."""
    lib = _lib.load()
    r1 = fit.n if r1 is None else r1
    assert ws.Kinv is not None, "factorize(..., want_kinv=True) first"
    elems = lib.gpk_grad_trace_work_elems(fit.n, fit.d, r0, r1)
    work = get_work("grad", elems * 8, fit.device)
    call("gpk_mll_grad_reduce_sym" if FIT_SYMMETRIC_TRACE else "gpk_mll_grad_reduce", ptr(fit.X), fit.X.stride(0), fit.n,
         fit.d, ptr(fit.hyp), ptr(ws.Kinv), fit.npad,
         ptr(fit.alpha), r0, r1, ptr(work), elems, ptr(ws.grad_out), _lib.stream_ptr(fit.device))
    return ws.grad_out[: lib.gpk_pad_dim(fit.d) + 2]


def loss_and_grad_from_partials(fit: GPFit, partials: torch.Tensor):
    """loss = -MLL/N and d loss / d (noise, sf2, ell_j) from the reduced trace partials (host floats).
    Closed form of SURVEY.md 8c: dMLL/dnoise = tr(G)/2, dMLL/dsf2 = sum(G K_f)/(2 sf2),
    dMLL/dell_j = sum(G K_f (x_aj-x_bj)^2/ell_j^2) / (2 ell_j)."""
    p = partials.tolist()
    n = fit.n
    loss = -fit.mll_per_n()
    g_noise = -(0.5 * p[0]) / n
    g_sf2 = -(0.5 * p[1] / fit.sf2) / n
    g_ell = [-(0.5 * p[2 + j] / float(fit.ell[j])) / n for j in range(fit.d)]
    return loss, g_noise, g_sf2, g_ell


def make_hyp_host(center, ell, sf2: float, noise: float, d: int) -> torch.Tensor:
    """make_hyp without the copy to the device."""
    stride = _lib.load().gpk_hyp_stride()
    h = torch.zeros(stride, dtype=F64)
    h[:d] = torch.as_tensor(center, dtype=F64).reshape(-1)
    h[32:64] = 1.0
    h[32 : 32 + d] = 1.0 / torch.as_tensor(ell, dtype=F64).reshape(-1)
    h[64] = float(sf2)
    h[65] = float(noise)
    return h


GRAPH_FIT_MIN_N = 512       # below: the launch chain is short enough that a replay is no faster (measured)
GRAPH_FIT_WARMUP = 2        # plain iterations before the capture (scratch, module loads, int8 item lists exist by then)


class GraphedMLL:
    """One MLL + gradient iteration of fit_gp (gpmpc/gp.py:59-63) as a CUDA GRAPH: assembly, Cholesky with the inverse
    factor on the side stream, alpha / quad / logdet, K^-1 and the gradient trace are ~170 dependent launches on two
    streams at N = 2048; replayed as one graph their launch and cross-stream latencies disappear (measured on B200,
    tools/graph_fit_experiment.py: N = 1024 1.09 -> 0.78 ms, 2048 2.16 -> 1.81 ms, 4096 5.04 -> 4.48 ms per iteration;
    bit-identical results).  The hyper-parameters live in a device block that is rewritten before every replay; the
    results come back in one host round trip.  The first GRAPH_FIT_WARMUP calls run the plain path."""

    def __init__(self, X: torch.Tensor, y: torch.Tensor, center=None):
        self.dev = init(X.device if X.is_cuda else None)
        self.key = X                                     # identity of the training inputs this graph was built for
        self.X = _as_device_f64(X, self.dev)
        self.y = _as_device_f64(y, self.dev).reshape(-1)
        self.n, self.d = self.X.shape
        self.npad = round_up(self.n, TILE)
        self.center = torch.as_tensor(self.X.mean(dim=0) if center is None else center, dtype=F64).reshape(-1).cpu()
        self.calls = 0
        self.graph = None

    def _capture(self):
        lib = _lib.load()
        dev, n, d, npad = self.dev, self.n, self.d, self.npad
        self.ws = ws = get_fit_scratch(npad, dev)        # the reference keeps the scratch alive for the graph's lifetime
        self.W = torch.empty(npad, npad, dtype=F64, device=dev)
        self.alpha = torch.empty(npad, dtype=F64, device=dev)
        self.hyp = torch.zeros(lib.gpk_hyp_stride(), dtype=F64, device=dev)
        self.hyp.copy_(self._hyp_host)
        self.ypad = torch.zeros(npad, dtype=F64, device=dev)
        self.ypad[:n] = self.y
        self.Kinv = ws.need_kinv()
        self.elems = lib.gpk_grad_trace_work_elems(n, d, 0, n)
        self.work = torch.empty(max(self.elems, 1), dtype=F64, device=dev)
        self.width = lib.gpk_pad_dim(d) + 2
        self.result = torch.zeros(3 + self.width, dtype=F64, device=dev)
        X, W, alpha, hyp = self.X, self.W, self.alpha, self.hyp
        kinv_digits = int(FIT_GEMM_DIGITS_KINV) if ws.gemm_work is not None else 0
        switch = 5 <= kinv_digits < ws.gemm_digits
        name = "gpk_mll_grad_reduce_sym" if FIT_SYMMETRIC_TRACE else "gpk_mll_grad_reduce"

        def enqueue():
            st = _lib.stream_ptr(dev)
            call("gpk_cross_kernel", 1, ptr(X), X.stride(0), n, 0, ptr(X), X.stride(0), n, 0, d, ptr(hyp), ptr(ws.L),
                 npad, 0, npad, npad, 1, 1, st)
            call("gpk_potrf_trtri", ptr(ws.L), npad, npad, ptr(ws.dinv), ptr(ws.info), ptr(W), npad, ptr(ws.U), npad,
                 ptr(ws.T), ws.T.numel(), st)
            call("gpk_solve_alpha", ptr(W), npad, ptr(ws.U), npad, ptr(ws.L), npad, npad, ptr(self.ypad), ptr(ws.t_work),
                 ptr(alpha), ptr(ws.scal), st)
            if switch:
                _register_fit_gemm_workspace(ws.gemm_work, kinv_digits)
            call("gpk_kinv", ptr(ws.U), npad, npad, ptr(self.Kinv), npad, st)
            if switch:
                _register_fit_gemm_workspace(ws.gemm_work, ws.gemm_digits)
            call(name, ptr(X), X.stride(0), n, d, ptr(hyp), ptr(self.Kinv), npad, ptr(alpha), 0, n, ptr(self.work),
                 self.elems, ptr(ws.grad_out), st)
            self.result[0] = ws.info[0]
            self.result[1:3] = ws.scal
            self.result[3:] = ws.grad_out[: self.width]

        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            enqueue()                                   # everything the capture must not allocate exists after this
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            enqueue()
        self.graph = g

    def __call__(self, ell, sf2: float, noise: float, max_tries: int = 3):
        """(loss, g_noise, g_sf2, g_ell, jitter) at the given hyper-parameters."""
        n, d = self.n, self.d
        ell_t = torch.as_tensor(ell, dtype=F64).reshape(-1).cpu()
        if ell_t.numel() == 1:
            ell_t = ell_t.expand(d).clone()
        self.calls += 1
        if self.graph is None and self.calls <= GRAPH_FIT_WARMUP:
            fit, ws = factorize(self.X, self.y, ell_t, sf2, noise, center=self.center, want_kinv=True, want_image=False,
                                max_tries=max_tries)
            return (*loss_and_grad_from_partials(fit, grad_partials(fit, ws)), fit.jitter)
        jitter = 0.0
        for attempt in range(max_tries + 1):
            self._hyp_host = make_hyp_host(self.center, ell_t, sf2, noise + jitter, d)
            if self.graph is None:
                self._capture()
            self.hyp.copy_(self._hyp_host)
            self.graph.replay()
            r = self.result.tolist()                    # the one host round trip of the iteration
            if int(r[0]) == 0:
                break
            if attempt == max_tries:
                raise NotPSDError(f"K + noise I not positive definite after jitter {jitter:g} (potrf info={int(r[0])})")
            jitter = 1e-8 if jitter == 0.0 else jitter * 10.0
        quad, logdet, p = r[1], r[2], r[3:]
        loss = 0.5 * (quad + logdet + n * LOG_2PI) / n
        g_noise = -(0.5 * p[0]) / n
        g_sf2 = -(0.5 * p[1] / float(sf2)) / n
        g_ell = [-(0.5 * p[2 + j] / float(ell_t[j])) / n for j in range(d)]
        return loss, g_noise, g_sf2, g_ell, jitter


@_on_arg_device
def small_loss_and_grad(X: torch.Tensor, y: torch.Tensor, ell, sf2: float, noise: float, *, center=None,
                        max_tries: int = 3, covariances: dict | None = None):
    """loss = -MLL/N and d loss / d (noise, sf2, ell_j) for ONE small GP (N <= 128) in one launch and one host round
    trip (gpk_mll_small): what fit_gp's iteration needs at the reference's own sizes (gpmpc/gp.py:59-63, N <= 45).
    Same jitter policy as factorize().  Returns (loss, g_noise, g_sf2, g_ell, jitter); a dict passed as `covariances`
    receives the dense K = K_f + noise I and K^-1 (device tensors) of the same launch."""
    lib = _lib.load()
    dev = init(X.device if X.is_cuda else None)
    X = _as_device_f64(X, dev)
    y = _as_device_f64(y, dev).reshape(-1)
    n, d = X.shape
    assert n <= lib.gpk_mll_small_max_n(), "small_loss_and_grad handles N <= 128"
    ell_t = torch.as_tensor(ell, dtype=F64).reshape(-1).cpu()
    if ell_t.numel() == 1:
        ell_t = ell_t.expand(d).clone()
    center_h = torch.as_tensor(X.mean(dim=0) if center is None else center, dtype=F64).reshape(-1).cpu()
    out = torch.empty(lib.gpk_mll_small_out_stride(), dtype=F64, device=dev)
    jitter = 0.0
    for attempt in range(max_tries + 1):
        hyp = make_hyp(center_h, ell_t, sf2, noise + jitter, d, dev)
        K = Kinv = None
        if covariances is not None:         # gp.K, gp.K_inv from the same launch (compute_covariances, gpmpc/gp.py:43-46)
            K = torch.empty(n, n, dtype=F64, device=dev)
            Kinv = torch.empty(n, n, dtype=F64, device=dev)
        call("gpk_mll_small", ptr(X), X.stride(0), 0, ptr(y), 0, n, d, ptr(hyp), ptr(out), None, 0, ptr(K), ptr(Kinv), n, 0,
             1, _lib.stream_ptr(dev))
        o = out.tolist()                    # the one host round trip of the iteration
        if int(o[0]) == 0:
            break
        if attempt == max_tries:
            raise NotPSDError(f"K + noise I not positive definite after jitter {jitter:g} (potrf info={int(o[0])})")
        jitter = 1e-8 if jitter == 0.0 else jitter * 10.0
    if covariances is not None:
        covariances["K"], covariances["K_inv"] = K, Kinv
    quad, logdet, p = o[1], o[2], o[4:]
    loss = 0.5 * (quad + logdet + n * LOG_2PI) / n
    g_noise = -(0.5 * p[0]) / n
    g_sf2 = -(0.5 * p[1] / float(sf2)) / n
    g_ell = [-(0.5 * p[2 + j] / float(ell_t[j])) / n for j in range(d)]
    return loss, g_noise, g_sf2, g_ell, jitter


@_on_arg_device
def small_batch_loss_and_grad(Xpad: torch.Tensor, Y: torch.Tensor, centers: torch.Tensor, ells: list, sf2s: list,
                              noises: list, covariances: dict | None = None):
    """small_loss_and_grad for B small GPs of the same N in ONE launch (gpk_mll_small, one CTA per GP): the reference's
    thrust / roll / pitch GPs (gpmpc/gpmpc.py:159-162 fits them one after the other).  Xpad (B, n, dp): inputs zero-padded
    to a common dimension (a zero column with lengthscale 1 leaves every kernel value unchanged); ells[b] has the GP's own
    d_b entries.  Returns a list of (loss, g_noise, g_sf2, g_ell[:d_b], info); info > 0 = that GP's K was not positive
    definite (the caller retries it alone, with jitter).  `covariances` receives K, K_inv as (B, n, n) device tensors."""
    lib = _lib.load()
    dev = init(Xpad.device)
    B, n, dp = Xpad.shape
    assert n <= lib.gpk_mll_small_max_n() and Y.shape == (B, n)
    stride = lib.gpk_hyp_stride()
    h = torch.zeros(B, stride, dtype=F64)
    for b in range(B):
        e = torch.ones(dp, dtype=F64)
        e[: len(ells[b])] = torch.as_tensor(ells[b], dtype=F64).reshape(-1)
        h[b] = make_hyp_host(centers[b], e, sf2s[b], noises[b], dp)
    hyp = h.to(dev)
    so = lib.gpk_mll_small_out_stride()
    out = torch.empty(B, so, dtype=F64, device=dev)
    K = Kinv = None
    if covariances is not None:
        K = torch.empty(B, n, n, dtype=F64, device=dev)
        Kinv = torch.empty(B, n, n, dtype=F64, device=dev)
    call("gpk_mll_small", ptr(Xpad), Xpad.stride(1), Xpad.stride(0), ptr(Y), Y.stride(0), n, dp, ptr(hyp), ptr(out), None, 0,
         ptr(K), ptr(Kinv), n, n * n, B, _lib.stream_ptr(dev))
    o = out.tolist()                        # the one host round trip for all B GPs
    if covariances is not None:
        covariances["K"], covariances["K_inv"] = K, Kinv
    res = []
    for b in range(B):
        quad, logdet, p = o[b][1], o[b][2], o[b][4:]
        d_b = len(ells[b])
        res.append((0.5 * (quad + logdet + n * LOG_2PI) / n, -(0.5 * p[0]) / n, -(0.5 * p[1] / float(sf2s[b])) / n,
                    [-(0.5 * p[2 + j] / float(ells[b][j])) / n for j in range(d_b)], int(o[b][0])))
    return res


@_on_arg_device
def dense_K(fit: GPFit) -> torch.Tensor:
    """K_f(X,X) + noise I as a dense (N,N) tensor (gpmpc/gp.py:45)."""
    n, npad = fit.n, round_up(fit.n, 64)
    out = torch.empty(npad, npad, dtype=F64, device=fit.device)
    call("gpk_cross_kernel", 1, ptr(fit.X), fit.X.stride(0), n, 0, ptr(fit.X), fit.X.stride(0), n, 0, fit.d,
         ptr(fit.hyp), ptr(out), npad, 0, npad, npad, 1, 1, _lib.stream_ptr(fit.device))
    return out[:n, :n]


@_on_arg_device
def cross_K(fit: GPFit, A: torch.Tensor, B: torch.Tensor | None = None) -> torch.Tensor:
    """Noise-free K_f(A, B) with this GP's hyper-parameters (gp.covar_module(a, b).to_dense(), gpmpc.py:392-393)."""
    dev = fit.device
    A = _as_device_f64(A, dev)
    B = A if B is None else _as_device_f64(B, dev)
    na, nb = A.shape[0], B.shape[0]
    rp, cp = round_up(na, 64), round_up(nb, 64)
    out = torch.empty(rp, cp, dtype=F64, device=dev)
    call("gpk_cross_kernel", 1, ptr(A), A.stride(0), na, 0, ptr(B), B.stride(0), nb, 0, fit.d, ptr(fit.hyp), ptr(out),
         cp, 0, rp, cp, 0, 1, _lib.stream_ptr(dev))
    return out[:na, :nb]


@dataclass
class GPBatch:
    """`batch` fitted GPs of identical (N, d) stacked for one-launch prediction (grid.z = GP)."""

    fits: list
    X: torch.Tensor        # (B, N, d)
    W: torch.Tensor        # (B, npad, npad)
    hyp: torch.Tensor      # (B, 72)
    wimg: torch.Tensor = None     # (B, digits * npad * npad) int8
    wscale: torch.Tensor = None   # (B, npad)
    digits: int = 0
    packs: dict = field(default_factory=dict)
    _alpha: torch.Tensor = None   # (B, npad), stacked lazily

    @property
    def digits_f32(self) -> int:
        """Leading digits multiplied for float32 queries: 4 only if every GP of the batch qualified for it."""
        d = [f.digits_f32 for f in self.fits]
        return 0 if (0 in d or not d) else max(d)

    @property
    def alpha(self) -> torch.Tensor:
        if self._alpha is None:
            self._alpha = torch.stack([f.alpha for f in self.fits])
        return self._alpha

    @classmethod
    def from_fits(cls, fits: list) -> "GPBatch":
        f0 = fits[0]
        assert all(f.n == f0.n and f.d == f0.d and f.device == f0.device for f in fits), "GPs must share (N, d, device)"
        has_img = all(f.wimg is not None and f.digits == f0.digits for f in fits)
        if len(fits) == 1:
            return cls(fits, f0.X[None], f0.W[None] if f0.W is not None else None, f0.hyp[None], f0.wimg[None] if has_img else None,
                       f0.wscale[None] if has_img else None, f0.digits if has_img else 0)
        # W tensors allocated by fit_batch are already slices of one (B, npad, npad) buffer
        base = fits[0].W._base if fits[0].W._base is not None else None
        if base is not None and base.dim() == 3 and base.shape[0] == len(fits) and all(
            f.W._base is base and f.W.data_ptr() == base[i].data_ptr() for i, f in enumerate(fits)
        ):
            W = base
        else:
            W = torch.stack([f.W for f in fits])
        X = torch.stack([f.X for f in fits])
        hyp = torch.stack([f.hyp for f in fits])
        wimg = wscale = None
        if has_img:
            ib = fits[0].wimg._base
            if ib is not None and ib.dim() == 2 and ib.shape[0] == len(fits) and all(
                f.wimg._base is ib and f.wimg.data_ptr() == ib[i].data_ptr() for i, f in enumerate(fits)
            ):
                wimg = ib
            else:
                wimg = torch.stack([f.wimg for f in fits])
            wscale = torch.stack([f.wscale for f in fits])
        return cls(fits, X, W, hyp, wimg, wscale, f0.digits if has_img else 0)

    @property
    def n(self):
        return self.fits[0].n

    @property
    def d(self):
        return self.fits[0].d

    @property
    def npad(self):
        return self.fits[0].npad

    @property
    def device(self):
        return self.X.device

    def pack(self, dtype):
        if dtype not in self.packs:
            self.packs[dtype] = torch.stack([f.pack(dtype) for f in self.fits])
        return self.packs[dtype]


@_on_arg_device
def fit_batch(X: torch.Tensor, Y: torch.Tensor, ell, sf2, noise, digits: int | None = None) -> GPBatch:
    """Fit P independent GPs that share the inputs X (N,d) with targets Y (N,P) at fixed hyper-parameters.
    ell: (P,d) or (d,) ; sf2, noise: (P,) or scalars.  The factors are written straight into one stacked buffer."""
    dev = _lib.require_cuda(X.device if X.is_cuda else None)
    X = _as_device_f64(X, dev)
    Y = _as_device_f64(Y, dev)
    n, d = X.shape
    P = Y.shape[1]
    npad = round_up(n, TILE)
    Wall = torch.empty(P, npad, npad, dtype=F64, device=dev)
    lib = _lib.load()
    ell = torch.as_tensor(ell, dtype=F64)
    fits = []
    for p in range(P):
        ell_p = ell[p] if ell.dim() == 2 else ell
        sf2_p = float(sf2[p]) if hasattr(sf2, "__len__") else float(sf2)
        noise_p = float(noise[p]) if hasattr(noise, "__len__") else float(noise)
        f, _ = factorize(X, Y[:, p], ell_p, sf2_p, noise_p, W_out=Wall[p], want_image=False)
        fits.append(f)
    # one image format for the batch: explicit digits, or the accuracy policy over all P GPs (select_digits)
    S0 = lib.gpk_oz_default_digits() if digits is None else digits
    if i8_supported(npad, S0):
        Iall = torch.empty(P, lib.gpk_oz_w_image_bytes(npad, S0), dtype=torch.int8, device=dev)
        if digits is None:
            select_digits(fits, wimg_out=list(Iall))
        else:
            for p, f in enumerate(fits):
                f.wimg, f.wscale, f.digits = slice_w(f.W, npad, digits, Iall[p])
                f.digits_f32 = min(digits, 4)
    elif digits is None:
        select_digits(fits)
    return GPBatch.from_fits(fits)


VAR_WORK_BYTES = 4 << 30   # default scratch budget for the K_ZX chunks of the variance path


@_on_arg_device
def predict(gp: GPBatch | GPFit, Z: torch.Tensor, *, want_var: bool = True, var_work_bytes: int | None = None,
            var_path: str = "i8", mj_event: torch.cuda.Event | None = None, mj_path: str = "auto",
            want_mean_jac: bool = True, mul_digits: int | None = None):
    """Posterior mean (B,Q), latent variance var_f (B,Q) and input-Jacobian (B,Q,d) at queries Z.

    Z: (Q,d) shared by all GPs or (B,Q,d); float64 or float32 selects the prediction path.  The batched
    entry configs 2-5 time; per stage it is what acados evaluates through gpytorch_predict2casadi
    (gpmpc/gp.py:72-85) plus the variance diagonal of gpmpc/gpmpc.py:443-445.
    """
    if isinstance(gp, GPFit):
        gp = GPBatch.from_fits([gp])
    lib = _lib.load()
    dev = init(gp.device)
    assert Z.dtype in DTYPE_CODE, "queries must be float32 or float64"
    dt = Z.dtype
    code = DTYPE_CODE[dt]
    Z = Z.to(dev).contiguous()
    B, n, d, npad = len(gp.fits), gp.n, gp.d, gp.npad
    if Z.dim() == 2:
        q, strideZ = Z.shape[0], 0
    else:
        assert Z.shape[0] == B
        q, strideZ = Z.shape[1], Z.stride(0)
    assert Z.shape[-1] == d
    st = _lib.stream_ptr(dev)
    assert want_var or want_mean_jac, "nothing to compute"
    mean = torch.empty(B, q, dtype=dt, device=dev) if want_mean_jac else None
    jac = torch.empty(B, q, d, dtype=dt, device=dev) if want_mean_jac else None
    var = torch.empty(B, q, dtype=dt, device=dev) if want_var else None
    if q == 0:
        return mean, var, jac
    if want_var and var_path == "i8" and gp.wimg is None:
        # no digit image: the training set is too large for exact int32 accumulation, or no digit count met the
        # accuracy policy at this fit's conditioning (select_digits) -> FP64 DMMA path
        assert gp.W is not None, "fit carries neither a digit image nor W"
        var_path = "f64"
    if not want_mean_jac and var_path != "i8":
        # variance alone exists on the int8 path only; the DMMA path computes mean and Jacobian on the way
        mean = torch.empty(B, q, dtype=dt, device=dev)
        jac = torch.empty(B, q, d, dtype=dt, device=dev)
    if want_var and var_path == "i8":
        # fused front end (kernel evaluated once for mean, Jacobian and the digit image) + variance contraction
        # on the tcgen05 int8 pipe (exact digit products, fp64 recombination)
        if mul_digits is None:               # float32 queries: the leading digits the fit qualified for
            mul_digits = gp.digits_f32 if dt == torch.float32 else 0
        if mul_digits >= gp.digits:
            mul_digits = 0
        budget = VAR_WORK_BYTES if var_work_bytes is None else var_work_bytes
        need = lib.gpk_predict_i8_work_bytes(q, npad, d, gp.digits, B)
        vbytes = max(min(need, budget), lib.gpk_predict_i8_work_bytes(128, npad, d, gp.digits, B))
        vwork = get_work("predict_i8", vbytes, dev)
        strideX = gp.X.stride(0) if B > 1 else 0
        alpha = gp.alpha
        call("gpk_predict_i8", code, ptr(Z), Z.stride(-2), q, strideZ, d, ptr(gp.hyp), ptr(gp.X), gp.X.stride(1), n,
             strideX, ptr(alpha), alpha.stride(0), ptr(gp.wimg), gp.wimg.stride(0) if B > 1 else 0, ptr(gp.wscale),
             gp.wscale.stride(0) if B > 1 else 0, npad, gp.digits, ptr(mean), q, ptr(jac), d, q * d, ptr(var), q,
             ptr(vwork), vbytes, B, int(mul_digits), mj_event.cuda_event if mj_event is not None else None, st)
        return mean, var, jac
    if not want_var and dt == F64 and d > 8 and mj_path != "simt":
        # mean + Jacobian alone in float64: the DMMA front kernel without the digit image (1.4x the SIMT kernel at d=16)
        wbytes = min(lib.gpk_mean_jac_dmma_work_bytes(q, d, B), 1 << 30)
        work = get_work("mean_jac_dmma", wbytes, dev)
        alpha = gp.alpha
        call("gpk_predict_mean_jac_dmma", code, ptr(Z), Z.stride(-2), q, strideZ, d, ptr(gp.hyp), ptr(gp.X),
             gp.X.stride(1), n, gp.X.stride(0) if B > 1 else 0, ptr(alpha), alpha.stride(0), ptr(mean), q, ptr(jac), d,
             q * d, ptr(work), wbytes, B, st)
        return mean, var, jac
    pack = gp.pack(dt)
    wbytes = min(lib.gpk_mean_jac_work_bytes(code, q, n, d, B), 1 << 30)
    work = get_work("mean_jac", wbytes, dev)
    call("gpk_predict_mean_jac", code, ptr(Z), Z.stride(-2), q, strideZ, d, ptr(gp.hyp), ptr(pack), pack.stride(0), n,
         ptr(mean), q, ptr(jac), d, q * d, ptr(work), wbytes, B, st)
    if want_var:
        assert var_path == "f64", "var_path must be 'i8' (tcgen05 int8 digits) or 'f64' (DMMA)"
        budget = VAR_WORK_BYTES if var_work_bytes is None else var_work_bytes
        need = lib.gpk_predict_var_work_bytes(q, npad, B)
        vbytes = max(min(need, budget), lib.gpk_predict_var_work_bytes(TILE, npad, B))
        vwork = get_work("var", vbytes, dev)
        strideX = gp.X.stride(0) if B > 1 else 0
        call("gpk_predict_var", code, ptr(Z), Z.stride(-2), q, strideZ, d, ptr(gp.hyp), ptr(gp.X), gp.X.stride(1), n,
             strideX, ptr(gp.W), npad, gp.W.stride(0) if B > 1 else 0, npad, ptr(var), q, ptr(vwork), vbytes, B, st)
    return mean, var, jac


@_on_arg_device
def posterior_covariance(fit: GPFit, Z: torch.Tensor) -> torch.Tensor:
    """Dense latent posterior covariance K_zz - K_zX K^-1 K_Xz (Q,Q) for modest Q: the matrix
    gp(z).covariance_matrix returns at gpmpc/gpmpc.py:444 (the caller keeps only its diagonal, for which
    `predict` is the fast path).  Three tensor-pipe GEMMs: V^T = K_zX W^T, cov = K_zz - V^T V."""
    dev = fit.device
    Z = _as_device_f64(Z, dev)
    q, npad = Z.shape[0], fit.npad
    qp = round_up(max(q, 1), TILE)
    st = _lib.stream_ptr(dev)
    kzx = torch.empty(qp, npad, dtype=F64, device=dev)
    vt = torch.empty(qp, npad, dtype=F64, device=dev)
    cov = torch.empty(qp, qp, dtype=F64, device=dev)
    X = fit.X
    call("gpk_cross_kernel", 1, ptr(Z), Z.stride(0), q, 0, ptr(X), X.stride(0), fit.n, 0, fit.d, ptr(fit.hyp), ptr(kzx),
         npad, 0, qp, npad, 0, 1, st)
    call("gpk_cross_kernel", 1, ptr(Z), Z.stride(0), q, 0, ptr(Z), Z.stride(0), q, 0, fit.d, ptr(fit.hyp), ptr(cov),
         qp, 0, qp, qp, 0, 1, st)
    call("gpk_dgemm_nt", ptr(kzx), npad, ptr(fit.W), npad, ptr(vt), npad, qp, npad, npad, 1.0, 0.0, 0, 0, st)
    call("gpk_dgemm_nt", ptr(vt), npad, ptr(vt), npad, ptr(cov), qp, qp, qp, npad, -1.0, 1.0, 0, 0, st)
    return cov[:q, :q].clone()
