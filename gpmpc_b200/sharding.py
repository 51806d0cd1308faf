"""Sharding layer (north-star subsystem d): one process per GPU under torch.distributed.

The reference has no counterpart (single process, a Python loop over its three GPs, gpmpc/gpmpc.py:159-162);
the three axes below are the ones SURVEY.md 8e names (plus rollouts for the MPC-side rows of 8f):

  queries      embarrassingly parallel given the fitted state: contiguous Q/G blocks per rank, fitted state
               (X, hyp, W, alpha) replicated once after the fit -> NO collective during prediction;
               an optional all_gather returns the full result on every rank.
  output GPs   GP p is owned by rank p mod G; each rank assembles / factorises / fits its own GPs -> no
               collective for the fit, one all_gather of the (d+2) fitted hyper-parameters afterwards.
  gradient rows  one GP's MLL-gradient trace G (.) dK/dtheta is split by rows [gN/G, (g+1)N/G) over the ranks;
               the (d+2) float64 partial sums are all-reduced (NCCL over NVLink; latency-bound, 18 doubles
               at d=16).  The Cholesky itself stays on one GPU per GP ("replicas only").

Every function takes the compute step as a callable so the same host logic runs over NCCL on GPUs and over
gloo in the CPU tests (where the callable is a test double; the product callables are in engine.py).
"""

from __future__ import annotations

from typing import Callable, Sequence

import torch
import torch.distributed as dist


def world_info(group=None) -> tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def shard_bounds(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank` when n items are split over `world` ranks; the first
    n % world ranks get one extra item, empty shards are allowed (n < world)."""
    assert world >= 1 and 0 <= rank < world and n >= 0
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def weighted_bounds(n: int, weights: Sequence[float], granularity: int = 1) -> list[tuple[int, int]]:
    """Contiguous blocks [lo, hi) of n items for ranks whose measured speeds are `weights` (items/s, any scale): rank r
    gets a share proportional to weights[r], cut at multiples of `granularity` (the 128-query tile of the prediction
    kernels) except for the last block.  Power-capped GPUs of one box hold different clocks on the same kernel
    (1.17-1.34 GHz at 8 GPUs, profiles/README_r02.md); with equal blocks the step ends when the slowest rank does."""
    world = len(weights)
    assert world >= 1 and n >= 0 and granularity >= 1 and all(w > 0 for w in weights)
    total = float(sum(weights))
    cuts = [0]
    acc = 0.0
    for r in range(world - 1):
        acc += weights[r]
        c = int(round(n * acc / total / granularity)) * granularity
        cuts.append(min(max(c, cuts[-1]), n))
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def measured_weights(items_local: int, seconds_local: float, group=None, clamp: float = 0.15) -> list[float]:
    """All-gather of every rank's measured speed (items/s) on its current block: one small collective OUTSIDE the
    prediction step, e.g. after a warm-up step; feed the result to weighted_bounds.  Speeds are clamped to within
    `clamp` of their median (the clocks of one box differ by ~15 %; a larger deviation is a hiccup of the measurement,
    which must not skew the blocks of every later step)."""
    rank, world = world_info(group)
    mine = float(items_local) / max(float(seconds_local), 1e-12)
    if world == 1:
        return [mine]
    out = [None] * world
    dist.all_gather_object(out, mine, group=group)
    w = [float(x) for x in out]
    med = sorted(w)[len(w) // 2]
    return [min(max(x, med * (1.0 - clamp)), med * (1.0 + clamp)) for x in w]


def gp_owner(p: int, world: int) -> int:
    """Rank that owns output GP p (round robin)."""
    return p % world


def owned_gps(n_gps: int, world: int, rank: int) -> list[int]:
    return [p for p in range(n_gps) if gp_owner(p, world) == rank]


# --------------------------------------------------------------------------------------------------
# query sharding
# --------------------------------------------------------------------------------------------------
def predict_sharded(predict_fn: Callable, Z: torch.Tensor, *, gather: bool = False, group=None,
                    bounds: Sequence[tuple[int, int]] | None = None):
    """Each rank predicts its contiguous block of the global query batch Z (Q,d): equal blocks (shard_bounds), or the
    blocks `bounds` (one (lo, hi) per rank, e.g. weighted_bounds from measured speeds; must be identical on all ranks).

    predict_fn(Z_local) -> tuple of tensors whose query axis is dim 1 for 2-D+ outputs shaped (B,q,...)
    (what engine.predict returns: mean (B,q), var (B,q), jac (B,q,d)).  Returns (lo, hi, outputs).  With
    gather=True the outputs are all_gathered along the query axis so that every rank holds all Q results
    in the original order; otherwise nothing is communicated.
    """
    rank, world = world_info(group)
    sizes = [shard_bounds(Z.shape[0], world, r) for r in range(world)] if bounds is None else [tuple(b) for b in bounds]
    assert len(sizes) == world and sizes[0][0] == 0 and sizes[-1][1] == Z.shape[0]
    assert all(sizes[r][1] == sizes[r + 1][0] for r in range(world - 1)), "blocks must tile the batch in rank order"
    lo, hi = sizes[rank]
    outs = predict_fn(Z[lo:hi])
    if not gather or world == 1:
        return lo, hi, outs
    gathered = []
    qmax = max(h - l for l, h in sizes)
    for t in outs:
        if t is None:
            gathered.append(None)
            continue
        # pad the query axis to the largest shard so that all_gather sees equal shapes
        pad_shape = list(t.shape)
        pad_shape[1] = qmax
        buf = t.new_zeros(pad_shape)
        buf[:, : t.shape[1]] = t
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf, group=group)
        gathered.append(torch.cat([p[:, : h - l] for p, (l, h) in zip(parts, sizes)], dim=1))
    return 0, Z.shape[0], tuple(gathered)


def rollouts_sharded(fn: Callable, inputs: Sequence[torch.Tensor], *, gather: bool = False, group=None):
    """Rollout / stage sharding for the MPC-side rows (mpc.ConstraintTightening.propagate_constraint_limits,
    mpc.StageDynamics.disc_dyn): every input has the independent items (rollouts, or (x, u) pairs) on dim 0; each rank
    processes its contiguous block, fn(*local_inputs) -> tuple of tensors with the items on dim 0.  No collective
    unless gather=True (all_gather along dim 0, original order).  Returns (lo, hi, outputs)."""
    rank, world = world_info(group)
    n = inputs[0].shape[0]
    assert all(t.shape[0] == n for t in inputs), "inputs disagree on the number of items"
    lo, hi = shard_bounds(n, world, rank)
    outs = fn(*[t[lo:hi] for t in inputs])
    if not gather or world == 1:
        return lo, hi, outs
    sizes = [shard_bounds(n, world, r) for r in range(world)]
    nmax = max(h - l for l, h in sizes)
    gathered = []
    for t in outs:
        buf = t.new_zeros((nmax,) + tuple(t.shape[1:]))
        buf[: t.shape[0]] = t
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf, group=group)
        gathered.append(torch.cat([p[: h - l] for p, (l, h) in zip(parts, sizes)], dim=0))
    return 0, n, tuple(gathered)


def broadcast_tensors(tensors: Sequence[torch.Tensor], src: int = 0, group=None) -> None:
    """Replicate fitted state (W, alpha, hyp, X ...) from `src` into the same-shaped buffers of every rank:
    the one-time state broadcast after a fit (about 1.07 GB per GP at N=16k float32 W)."""
    _, world = world_info(group)
    if world == 1:
        return
    for t in tensors:
        dist.broadcast(t, src=src, group=group)


def replicate_batch(batch, src: int = 0, *, device=None, with_W: bool = False, group=None):
    """Fitted prediction state of `src` (an engine.GPBatch; None on the other ranks) -> an equivalent GPBatch on
    every rank: the one-time state replication after a fit that query sharding needs.  One small
    object broadcast (sizes, hyper-parameter scalars) and one broadcast per tensor: X, y, hyp, alpha and -- when the
    fit carries them -- the int8 digit image of W = L^-1 with its row scales (what the tcgen05 variance path reads;
    S N^2 bytes per GP) ; the float64 W itself (8 N^2 bytes per GP, only the DMMA variance path reads it) travels
    only with with_W=True.  Returns (batch, bytes received per rank)."""
    from . import engine as E

    rank, world = world_info(group)
    if world == 1:
        return batch, 0
    meta = [None]
    if rank == src:
        f0 = batch.fits[0]
        has_img = batch.wimg is not None
        meta[0] = {"B": len(batch.fits), "n": f0.n, "d": f0.d, "npad": f0.npad, "digits": batch.digits if has_img else 0,
                   "img_bytes": int(batch.wimg.shape[1]) if has_img else 0,
                   "with_W": bool(with_W and batch.W is not None),
                   "fits": [{"center": f.center, "ell": f.ell, "sf2": f.sf2, "noise": f.noise, "quad": f.quad,
                             "logdet": f.logdet, "jitter": f.jitter, "digits_f32": f.digits_f32,
                             "var_err": dict(f.var_err)} for f in batch.fits]}
        device = batch.device
    dist.broadcast_object_list(meta, src=src, group=group, device=device if device is not None and
                               torch.device(device).type == "cuda" else None)
    m = meta[0]
    B, n, d, npad = m["B"], m["n"], m["d"], m["npad"]
    F64 = torch.float64
    if rank == src:
        X, hyp, alpha = batch.X.contiguous(), batch.hyp.contiguous(), batch.alpha.contiguous()
        y = torch.stack([f.y for f in batch.fits])
        wimg, wscale = batch.wimg, batch.wscale
        W = batch.W if m["with_W"] else None
    else:
        X = torch.empty(B, n, d, dtype=F64, device=device)
        y = torch.empty(B, n, dtype=F64, device=device)
        hyp = torch.empty(B, batch_hyp_stride(), dtype=F64, device=device)
        alpha = torch.empty(B, npad, dtype=F64, device=device)
        wimg = torch.empty(B, m["img_bytes"], dtype=torch.int8, device=device) if m["digits"] else None
        wscale = torch.empty(B, npad, dtype=F64, device=device) if m["digits"] else None
        W = torch.empty(B, npad, npad, dtype=F64, device=device) if m["with_W"] else None
    tensors = [t for t in (X, y, hyp, alpha, wimg, wscale, W) if t is not None]
    broadcast_tensors(tensors, src=src, group=group)
    nbytes = sum(t.numel() * t.element_size() for t in tensors)
    if rank == src:
        return batch, nbytes
    fits = [E.GPFit(X=X[p], y=y[p], n=n, d=d, npad=npad, center=fm["center"], ell=fm["ell"], sf2=fm["sf2"],
                    noise=fm["noise"], hyp=hyp[p], W=W[p] if W is not None else None, alpha=alpha[p],
                    wimg=wimg[p] if wimg is not None else None, wscale=wscale[p] if wscale is not None else None,
                    digits=m["digits"], digits_f32=fm["digits_f32"], var_err=fm["var_err"], quad=fm["quad"],
                    logdet=fm["logdet"], jitter=fm["jitter"])
            for p, fm in enumerate(m["fits"])]
    out = E.GPBatch(fits, X, W, hyp, wimg, wscale, m["digits"])
    out._alpha = alpha
    return out, nbytes


def batch_hyp_stride() -> int:
    """Doubles per GP in the hyper-parameter block (include/gpmpc_b200.h: centre[32] | 1/ell[32] | sf2 | noise | pad)."""
    return 72


# --------------------------------------------------------------------------------------------------
# output-GP sharding
# --------------------------------------------------------------------------------------------------
def fit_gps_sharded(fit_one: Callable[[int], torch.Tensor], n_gps: int, width: int, *, device=None, group=None):
    """Rank r fits the GPs it owns with fit_one(p) -> 1-D tensor of `width` float64 values (the fitted
    hyper-parameters noise, sf2, ell_1..d, and anything else the caller wants shared, e.g. the final loss).
    Returns an (n_gps, width) tensor available on every rank (one all_gather)."""
    rank, world = world_info(group)
    mine = owned_gps(n_gps, world, rank)
    per_rank = (n_gps + world - 1) // world
    local = torch.zeros(per_rank, width, dtype=torch.float64, device=device)
    for slot, p in enumerate(mine):
        v = fit_one(p)
        local[slot] = v.to(local.device, torch.float64)
    if world == 1:
        return local[:n_gps]
    parts = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(parts, local, group=group)
    out = torch.zeros(n_gps, width, dtype=torch.float64, device=local.device)
    for r in range(world):
        for slot, p in enumerate(owned_gps(n_gps, world, r)):
            out[p] = parts[r][slot]
    return out


# --------------------------------------------------------------------------------------------------
# row-split MLL gradient
# --------------------------------------------------------------------------------------------------
def allreduce_row_partials(partials_fn: Callable[[int, int], torch.Tensor], n_rows: int, *, group=None) -> torch.Tensor:
    """partials_fn(r0, r1) -> (d+2,) float64 partial sums of the gradient trace over rows [r0, r1)
    (engine.grad_partials).  Every rank reduces its row block; one all_reduce(sum) of d+2 doubles makes the
    full sums available everywhere.  A rank with an empty block contributes zeros."""
    rank, world = world_info(group)
    r0, r1 = shard_bounds(n_rows, world, rank)
    if r1 > r0:
        part = partials_fn(r0, r1).clone()
    else:
        part = torch.zeros_like(partials_fn(0, min(1, n_rows)))
    if world > 1:
        dist.all_reduce(part, op=dist.ReduceOp.SUM, group=group)
    return part


def balanced_row_bounds(n: int, world: int, rank: int, align: int = 128, lower_only: bool = True) -> tuple[int, int]:
    """Row block [r0, r1) of `rank` for the row-split K^-1 + gradient trace of one GP.  Row i of K^-1 = U U^T costs
    (n - i) multiply-adds per column (upper-triangular U).  With the symmetric trace (lower_only) a block only forms
    columns [0, r1): work(r0, r1) = r1 * integral_{r0}^{r1} (n - i) di; with all columns: n * the same integral, i.e.
    r_g = n (1 - sqrt(1 - g / world)).  Cuts are rounded to `align` (the tile of gpk_kinv_rows); the last block ends
    at n."""
    assert world >= 1 and 0 <= rank < world and n >= 0

    def snap(r: float) -> int:
        return min(n, int(round(r / align)) * align)

    if not lower_only or world == 1 or n == 0:
        cuts = [snap(n * (1.0 - (1.0 - g / world) ** 0.5)) for g in range(world)] + [n]
    else:
        def integral(a: float, b: float) -> float:      # of (n - i) over [a, b]
            return (b - a) * (n - 0.5 * (a + b))

        def cuts_for(w: float) -> list[float]:           # greedy cuts so that every block has work w (the last: the rest)
            c = [0.0]
            for _ in range(world - 1):
                lo, hi = c[-1], float(n)
                for _ in range(50):
                    mid = 0.5 * (lo + hi)
                    if mid * integral(c[-1], mid) < w:
                        lo = mid
                    else:
                        hi = mid
                c.append(0.5 * (lo + hi))
            return c + [float(n)]

        lo_w, hi_w = 0.0, float(n) * integral(0.0, float(n))
        for _ in range(60):                               # the work per block for which the last block matches too
            w = 0.5 * (lo_w + hi_w)
            c = cuts_for(w)
            if n * integral(c[-2], float(n)) > w:
                lo_w = w
            else:
                hi_w = w
        cuts = [snap(x) for x in cuts_for(0.5 * (lo_w + hi_w))[:-1]] + [n]
    cuts = [max(cuts[i], max(cuts[:i], default=0)) for i in range(len(cuts))]      # monotone after rounding
    return cuts[rank], cuts[rank + 1]


class RowSplitGradReduce:
    """grad_reduce callback for gp.fit_gp(..., grad_reduce=...): every rank factorises the same GP (the Cholesky is not
    distributed), forms ONLY its rows of K^-1 (`rows(n)`, picked up by the fit loop) and reduces the
    gradient trace over them; one all_reduce(sum) of d+2 doubles completes the gradient on every rank."""

    def __init__(self, group=None, time_allreduce: bool = False):
        self.group = group
        # with time_allreduce the all_reduce of every call is bracketed by CUDA events on the caller's stream
        # (`allreduce_events`, consumed by `allreduce_ms()`): wait for the slowest rank + the collective itself
        self.time_allreduce = time_allreduce
        self.allreduce_events: list = []

    def rows(self, n: int) -> tuple[int, int]:
        from . import engine as E

        rank, world = world_info(self.group)
        return balanced_row_bounds(n, world, rank, lower_only=bool(E.FIT_SYMMETRIC_TRACE))

    def __call__(self, fit, ws):
        from . import engine as E

        rank, world = world_info(self.group)
        r0, r1 = self.rows(fit.n)
        if r1 > r0:
            part = E.grad_partials(fit, ws, r0, r1).clone()
        else:   # empty block (small n): contributes zeros, and must not touch K^-1 rows it never formed
            part = torch.zeros(E._lib.load().gpk_pad_dim(fit.d) + 2, dtype=torch.float64, device=fit.device)
        if world > 1:
            timed = self.time_allreduce and part.is_cuda
            if timed:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            dist.all_reduce(part, op=dist.ReduceOp.SUM, group=self.group)
            if timed:
                e1.record()
                self.allreduce_events.append((e0, e1))
        return part

    def allreduce_ms(self) -> list:
        """Durations (ms) of the timed all_reduces since the last call; synchronises the recorded events."""
        out = []
        for e0, e1 in self.allreduce_events:
            e1.synchronize()
            out.append(e0.elapsed_time(e1))
        self.allreduce_events = []
        return out


def rowsplit_grad_reduce(group=None, time_allreduce: bool = False) -> Callable:
    """See RowSplitGradReduce."""
    return RowSplitGradReduce(group, time_allreduce)
