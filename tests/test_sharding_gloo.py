"""world_size-2 gloo tests (CPU) of the sharding layer's host logic: partition arithmetic, gather order,
GP ownership and the row-split gradient all-reduce.  The compute callables are oracle-backed test doubles;
the product callables (engine.predict / engine.grad_partials) are exercised in test_gpu_multi.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gpmpc_b200 import sharding as S
from oracle import gp_oracle as o


def test_shard_bounds_cover_everything():
    for n in (0, 1, 5, 8192, 8193, 4194304):
        for world in (1, 2, 3, 8):
            blocks = [S.shard_bounds(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [h - l for l, h in blocks]
            assert max(sizes) - min(sizes) <= 1


def test_weighted_bounds_follow_measured_speeds():
    # equal speeds -> equal tile-aligned blocks; a slow rank gets proportionally fewer queries; blocks always tile [0, n)
    assert S.weighted_bounds(1024, [1.0, 1.0], 128) == [(0, 512), (512, 1024)]
    b = S.weighted_bounds(1048576, [1.34, 1.17, 1.30, 1.25, 1.2, 1.33, 1.28, 1.22], 128)
    assert b[0][0] == 0 and b[-1][1] == 1048576 and all(b[i][1] == b[i + 1][0] for i in range(7))
    assert all(l % 128 == 0 for l, _ in b)
    sizes = [h - l for l, h in b]
    w = [1.34, 1.17, 1.30, 1.25, 1.2, 1.33, 1.28, 1.22]
    t = [sz / wi for sz, wi in zip(sizes, w)]
    assert max(t) / min(t) < 1.002                      # equal finishing times to within a tile
    for n in (0, 1, 100, 129):                          # degenerate sizes: still a partition, empty blocks allowed
        bb = S.weighted_bounds(n, [3.0, 1.0, 2.0], 128)
        assert bb[0][0] == 0 and bb[-1][1] == n and all(bb[i][1] == bb[i + 1][0] and bb[i][0] <= bb[i][1] for i in range(2))
    assert S.measured_weights(100, 2.0) == [50.0]      # single process: no process group needed


def test_gp_ownership_round_robin():
    assert [S.gp_owner(p, 8) for p in range(8)] == list(range(8))
    assert S.owned_gps(3, 2, 0) == [0, 2] and S.owned_gps(3, 2, 1) == [1]
    assert S.owned_gps(3, 8, 5) == []


def test_single_process_paths_need_no_process_group():
    Z = torch.arange(12.0).reshape(6, 2)
    lo, hi, outs = S.predict_sharded(lambda z: (z.sum(1)[None],), Z, gather=True)
    assert (lo, hi) == (0, 6) and torch.equal(outs[0][0], Z.sum(1))
    part = S.allreduce_row_partials(lambda a, b: torch.tensor([float(b - a)], dtype=torch.float64), 10)
    assert float(part) == 10.0
    lo, hi, outs = S.rollouts_sharded(lambda a, b: (a + b[:, :1],), [Z, Z * 2.0], gather=True)
    assert (lo, hi) == (0, 6) and torch.equal(outs[0], Z + 2.0 * Z[:, :1])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q_total):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        p = o.synthetic_problem(60, 3, q_total, p=3, seed=5)
        X, Y, Z = p["X"], p["Y"], p["Z"]
        states = [o.fit_state(X, Y[:, i], p["ell"], 1.0, 1e-2) for i in range(3)]

        def predict_fn(z):   # same output convention as engine.predict: (B,q), (B,q), (B,q,d)
            res = [o.predict(z, X, p["ell"], 1.0, s["L"], s["alpha"]) for s in states]
            return tuple(torch.stack([r[k] for r in res]) for k in range(3))

        full = predict_fn(Z)
        # --- query sharding, no gather: my block only
        lo, hi, outs = S.predict_sharded(predict_fn, Z)
        assert (lo, hi) == S.shard_bounds(q_total, world, rank)
        for a, b in zip(outs, full):
            assert torch.equal(a, b[:, lo:hi])
        # --- with gather: everything, original order, on every rank (ragged shards when q_total is odd)
        lo, hi, outs = S.predict_sharded(predict_fn, Z, gather=True)
        assert (lo, hi) == (0, q_total)
        for a, b in zip(outs, full):
            assert torch.equal(a, b)
        # --- speed-weighted blocks (rank 1 three times as fast): same results, same order, one small all-gather outside
        w = S.measured_weights(10 if rank == 0 else 30, 1.0, clamp=10.0)
        assert w == [10.0, 30.0]
        assert S.measured_weights(10 if rank == 0 else 30, 1.0) == [25.5, 30.0]      # default: within 15 % of the median
        wb = S.weighted_bounds(q_total, w, 1)
        lo, hi, outs = S.predict_sharded(predict_fn, Z, bounds=wb)
        assert (lo, hi) == wb[rank] and wb[0][1] == round(q_total * 0.25)
        for a, b in zip(outs, full):
            assert torch.equal(a, b[:, lo:hi])
        lo, hi, outs = S.predict_sharded(predict_fn, Z, gather=True, bounds=wb)
        for a, b in zip(outs, full):
            assert torch.equal(a, b)
        # --- row-split gradient trace: sum of per-rank row blocks == full trace
        n, d = X.shape
        rn, rs, rl = torch.tensor([-2.0]), torch.tensor(0.3), torch.full((1, d), 0.5)
        noise, sf2, ell = o.constrained(rn, rs, rl)
        st = o.fit_state(X, Y[:, 0], ell.reshape(-1), sf2, noise)
        Kf = st["K"] - float(noise) * torch.eye(n, dtype=torch.float64)
        G = torch.outer(st["alpha"], st["alpha"]) - torch.cholesky_inverse(st["L"])
        Xs = X / ell.reshape(-1)

        def partials(r0, r1):   # same layout as engine.grad_partials: [tr G, sum G K_f, sum G K_f dx_j^2 ...]
            GK = (G * Kf)[r0:r1]
            out = [G.diagonal()[r0:r1].sum(), GK.sum()]
            for j in range(d):
                df = Xs[r0:r1, j : j + 1] - Xs[:, j : j + 1].T
                out.append((GK * df * df).sum())
            return torch.stack(out)

        red = S.allreduce_row_partials(partials, n)
        assert torch.allclose(red, partials(0, n), rtol=1e-12, atol=1e-12)
        # --- the fit loop's reducer: balanced row blocks (rank 0's block is EMPTY at n = 60: cuts are 128-aligned), the
        #     product callable replaced by the oracle double, padded like engine.grad_partials (pad_dim(d) + 2 entries)
        from gpmpc_b200 import engine as E

        pad = E._lib.load().gpk_pad_dim(d) + 2

        class _Fit:
            pass

        fake = _Fit()
        fake.n, fake.d, fake.device = n, d, torch.device("cpu")
        rsr = S.rowsplit_grad_reduce()
        assert rsr.rows(n) == S.balanced_row_bounds(n, world, rank)
        assert rsr.rows(n) == ((0, 0) if rank == 0 else (0, n))
        orig = E.grad_partials
        E.grad_partials = lambda fit, ws, r0, r1: torch.cat([partials(r0, r1), torch.zeros(pad - d - 2, dtype=torch.float64)])
        try:
            red2 = rsr(fake, None)
        finally:
            E.grad_partials = orig
        assert red2.shape == (pad,) and torch.allclose(red2[: d + 2], partials(0, n), rtol=1e-12, atol=1e-12)
        # --- output-GP sharding: each rank "fits" its own GPs, everybody gets all results
        calls = []

        def fit_one(pidx):
            calls.append(pidx)
            return torch.tensor([pidx, 10.0 * pidx, rank], dtype=torch.float64)

        allp = S.fit_gps_sharded(fit_one, 3, 3)
        assert calls == S.owned_gps(3, world, rank)
        assert torch.equal(allp[:, 0], torch.tensor([0.0, 1.0, 2.0], dtype=torch.float64))
        assert torch.equal(allp[:, 2], torch.tensor([float(S.gp_owner(i, world)) for i in range(3)], dtype=torch.float64))
        # --- rollout sharding of the MPC-side rows: tightening recursion per rollout, ragged shards, gather order
        import numpy as np
        g = np.load(os.path.join(os.path.dirname(__file__), "golden", "tighten_T25.npz"))
        xp, up = torch.tensor(g["x_prev"]), torch.tensor(g["u_prev"])
        xp, up = torch.cat([xp, xp[:1] * 0.5]), torch.cat([up, up[:1] * 0.5])       # 5 rollouts over 2 ranks: 3 + 2

        def tighten_fn(x_loc, u_loc):
            sc, ic = [], []
            for r in range(x_loc.shape[0]):
                z = np.hstack((x_loc[r].numpy()[:, :-1].T, u_loc[r].numpy().T))
                var_y = np.stack([np.full(z.shape[0], 1e-3 * (k + 1)) + z[:, k] ** 2 * 1e-3 for k in range(3)])
                a, b = o.propagate_constraint_limits(var_y, z, g["noise"], g["dfdx"], g["dfdu"], g["lqr_gain"], g["Bd"],
                                                     float(g["dt"]), float(g["inverse_cdf"]))
                sc.append(torch.tensor(a)), ic.append(torch.tensor(b))
            return torch.stack(sc), torch.stack(ic)

        full_sc, full_ic = tighten_fn(xp, up)
        lo, hi, (sc, ic) = S.rollouts_sharded(tighten_fn, [xp, up])
        assert (lo, hi) == S.shard_bounds(5, world, rank) and torch.equal(sc, full_sc[lo:hi])
        lo, hi, (sc, ic) = S.rollouts_sharded(tighten_fn, [xp, up], gather=True)
        assert (lo, hi) == (0, 5) and torch.equal(sc, full_sc) and torch.equal(ic, full_ic)
        # --- state broadcast
        t = torch.full((4,), float(rank))
        S.broadcast_tensors([t], src=1)
        assert torch.equal(t, torch.ones(4))
        # --- replication of a fitted GPBatch from rank 0 (host logic only: CPU tensors stand in for the device state)
        from gpmpc_b200 import engine as E

        def fake_batch(seed):
            gg = torch.Generator().manual_seed(seed)
            Bn, nn, dd, npad, S_ = 2, 60, 3, 128, 5
            fits = []
            for k in range(Bn):
                fits.append(E.GPFit(X=torch.rand(nn, dd, generator=gg, dtype=torch.float64), y=torch.rand(nn, generator=gg, dtype=torch.float64),
                                    n=nn, d=dd, npad=npad, center=torch.rand(dd, generator=gg, dtype=torch.float64),
                                    ell=torch.rand(dd, generator=gg, dtype=torch.float64) + 0.5, sf2=1.0 + k, noise=1e-2 * (k + 1),
                                    hyp=torch.rand(72, generator=gg, dtype=torch.float64), W=torch.rand(npad, npad, generator=gg, dtype=torch.float64),
                                    alpha=torch.rand(npad, generator=gg, dtype=torch.float64),
                                    wimg=torch.randint(-128, 127, (S_ * npad * npad,), generator=gg, dtype=torch.int8),
                                    wscale=torch.rand(npad, generator=gg, dtype=torch.float64), digits=S_, quad=3.0 + k, logdet=-7.0 * (k + 1)))
            return E.GPBatch.from_fits(fits)

        want = fake_batch(99)
        for with_W in (False, True):
            got, nbytes = S.replicate_batch(fake_batch(99) if rank == 0 else None, src=0, device=torch.device("cpu"), with_W=with_W)
            assert nbytes > 0 and len(got.fits) == 2 and got.digits == 5 and (got.n, got.d, got.npad) == (60, 3, 128)
            for name in ("X", "hyp", "alpha", "wimg", "wscale"):
                assert torch.equal(getattr(got, name), getattr(want, name)), name
            assert (got.W is not None and torch.equal(got.W, want.W)) if (with_W or rank == 0) else got.W is None
            for fg, fw in zip(got.fits, want.fits):
                assert torch.equal(fg.y, fw.y) and torch.equal(fg.ell, fw.ell) and torch.equal(fg.center, fw.center)
                assert (fg.sf2, fg.noise, fg.quad, fg.logdet) == (fw.sf2, fw.noise, fw.quad, fw.logdet)
                assert abs(fg.mll_per_n() - fw.mll_per_n()) == 0.0
        # --- timed reducer keeps working without CUDA (no events on CPU tensors)
        rst = S.rowsplit_grad_reduce(time_allreduce=True)
        assert rst.allreduce_ms() == []
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("q_total", [64, 37])
def test_world_size_2_gloo(q_total):
    mp.spawn(_worker, args=(2, _free_port(), q_total), nprocs=2, join=True)


def test_balanced_row_bounds_partition():
    """Row blocks of the row-split K^-1: disjoint, cover [0, n), 128-aligned inner cuts, equal work under the cost
    model of each mode (all columns: n * integral of (n - i); lower triangle only: r1 * that integral)."""
    from gpmpc_b200 import sharding as S

    for lower_only in (False, True):
        for n in (0, 45, 128, 500, 2048, 32768):
            for world in (1, 2, 3, 8):
                blocks = [S.balanced_row_bounds(n, world, r, lower_only=lower_only) for r in range(world)]
                assert blocks[0][0] == 0 and blocks[-1][1] == n
                for (a0, a1), (b0, b1) in zip(blocks, blocks[1:]):
                    assert a1 == b0 and a0 <= a1
                assert all(lo % 128 == 0 for lo, _ in blocks)
                if n >= 128 * 8 * world:
                    work = [(hi if lower_only else n) * (hi - lo) * (n - 0.5 * (lo + hi)) for lo, hi in blocks]
                    assert max(work) < 1.35 * (sum(work) / world)
                    if not lower_only:          # later rows are cheaper: later blocks are longer
                        sizes = [hi - lo for lo, hi in blocks]
                        assert sizes == sorted(sizes)
