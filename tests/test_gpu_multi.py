"""2-GPU NCCL test of the sharding layer over the real CUDA path (skipped with fewer than 2 GPUs)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import gp_oracle as o

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from gpmpc_b200 import engine as E, sharding as S
        from gpmpc_b200.gp import GaussianProcess, fit_gp

        p = o.synthetic_problem(500, 16, 1001, p=3, seed=11)
        X, Y, Z = p["X"].to(dev), p["Y"].to(dev), p["Z"].to(dev)
        gb = E.fit_batch(X, Y, p["ell"], 1.0, 1e-2)            # replicated fitted state
        full = E.predict(gb, Z)
        lo, hi, outs = S.predict_sharded(lambda z: E.predict(gb, z), Z, gather=True)
        assert (lo, hi) == (0, 1001)
        for a, b in zip(outs, full):
            assert o.rel_err(a.cpu(), b.cpu()) < 1e-12
        # state broadcast: rank 1 overwrites its W / alpha with rank 0's
        f0 = gb.fits[0]
        if rank == 1:
            f0.W.zero_()
            f0.alpha.zero_()
        S.broadcast_tensors([f0.W, f0.alpha], src=0)
        st = o.fit_state(p["X"], p["Y"][:, 0], p["ell"], 1.0, 1e-2)
        assert o.rel_err(f0.alpha[: f0.n].cpu(), st["alpha"]) < 1e-8
        # row-split MLL gradient == single-GPU gradient
        fit, ws = E.factorize(X, Y[:, 0], p["ell"], 1.0, 1e-2, want_kinv=True)
        ref = E.grad_partials(fit, ws).clone()
        red = S.rowsplit_grad_reduce()(fit, ws)
        assert o.rel_err(red.cpu(), ref.cpu()) < 1e-11
        # the same with each rank forming ONLY its rows of K^-1 (what fit_gp does with this reducer)
        rs = S.rowsplit_grad_reduce()
        rows = rs.rows(fit.n)
        fit2, ws2 = E.factorize(X, Y[:, 0], p["ell"], 1.0, 1e-2, want_kinv=rows)
        k0, k1 = E.kinv_row_hull(rows[0], rows[1], fit2.npad)
        ws2.Kinv[:k0].fill_(float("nan"))                  # rows of other ranks must never be read
        ws2.Kinv[k1:].fill_(float("nan"))
        ws2.Kinv[k0:k1, k1:].fill_(float("nan"))           # nor the columns right of the block (symmetric trace)
        assert o.rel_err(rs(fit2, ws2).cpu(), ref.cpu()) < 1e-11
        # fit_gp with the row-split gradient follows the same trajectory as the oracle loop
        Xs, ys = p["X"][:45, :3].contiguous(), p["Y"][:45, 0].contiguous()
        gp = GaussianProcess(Xs, ys)
        fit_gp(gp, n_train=25, lr=0.05, device=f"cuda:{rank}", grad_reduce=S.rowsplit_grad_reduce())
        refl = o.fit_hyperparameters(Xs, ys, n_train=25, lr=0.05)["losses"]
        assert o.rel_err(torch.tensor(gp.fit_losses), torch.tensor(refl)) < 1e-5
        # output-GP sharding: each rank fits the GPs it owns, hyper-parameters gathered everywhere
        def fit_one(pi):
            g = GaussianProcess(Xs, p["Y"][:45, pi].contiguous())
            fit_gp(g, n_train=5, lr=0.05, device=f"cuda:{rank}", compute_covariances=False)
            ell, sf2, noise = g.hyperparameters()
            return torch.tensor([noise, sf2, float(ell[0])], dtype=torch.float64)

        allp = S.fit_gps_sharded(fit_one, 3, 3, device=dev)
        assert allp.shape == (3, 3) and bool((allp > 0).all())
        # rollout sharding of the MPC-side rows: tightening and stage dynamics per rank, gathered == single GPU
        import numpy as np
        from gpmpc_b200.mpc import ConstraintTightening, StageDynamics
        g = np.load(os.path.join(os.path.dirname(__file__), "golden", "tighten_T25.npz"))
        gps = [(torch.tensor(g[f"gp{k}_X"]), torch.tensor(g[f"gp{k}_y"]), torch.tensor(g[f"gp{k}_ell"]),
                float(g[f"gp{k}_sf2"]), float(g["noise"][k])) for k in range(3)]
        ct = ConstraintTightening(gps, [[0], [1, 2, 3], [4, 5, 6]], g["dfdx"], g["dfdu"], g["lqr_gain"], g["Bd"],
                                  float(g["dt"]), float(g["inverse_cdf"]))
        xp = torch.tensor(g["x_prev"]).repeat(3, 1, 1)[:11].to(dev)            # 11 rollouts: 6 + 5
        up = torch.tensor(g["u_prev"]).repeat(3, 1, 1)[:11].to(dev)
        full_sc, full_ic = ct.propagate_constraint_limits(xp, up)
        lo, hi, (sc, ic) = S.rollouts_sharded(ct.propagate_constraint_limits, [xp, up], gather=True)
        assert (lo, hi) == (0, 11) and o.rel_err(sc.cpu(), full_sc.cpu()) < 1e-12 and o.rel_err(ic.cpu(), full_ic.cpu()) < 1e-12
        assert o.rel_err(sc[:4].cpu(), g["state_constraint"]) < 1e-7
        sd = StageDynamics(gps, [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653], 0.02)
        xq, uq = xp[:, :, 0].contiguous(), up[:, :, 0].contiguous()
        full = sd.disc_dyn(xq, uq)
        lo, hi, outs = S.rollouts_sharded(sd.disc_dyn, [xq, uq], gather=True)
        for a, b in zip(outs, full):
            assert torch.equal(a, b)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_nccl_sharding():
    mp.spawn(_worker, args=(2, _free_port()), nprocs=2, join=True)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_second_device_in_one_process():
    """One process, two devices: tensors on cuda:1 while cuda:0 is current.  Launches, function attributes (dynamic
    shared memory), the fit's side stream and the int8 GEMM scratch are all per device."""
    from gpmpc_b200 import engine as E

    assert torch.cuda.current_device() == 0
    p = o.synthetic_problem(2304, 16, 300, p=1, seed=21)
    outs = []
    for idx in (0, 1):
        dev = torch.device("cuda", idx)
        fit, ws = E.factorize(p["X"].to(dev), p["Y"][:, 0].to(dev), p["ell"], p["sf2"], p["noise"], want_kinv=True)
        part = E.grad_partials(fit, ws).cpu()
        mean, var, jac = E.predict(fit, p["Z"].to(dev))
        outs.append((fit.mll_per_n(), part, mean.cpu(), var.cpu(), jac.cpu()))
        assert mean.device == dev and torch.cuda.current_device() == 0
    assert outs[0][0] == outs[1][0]
    for a, b in zip(outs[0][1:], outs[1][1:]):
        assert torch.equal(a, b)
    E.release_scratch()
