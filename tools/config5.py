"""Config 5 (BASELINE.json configs[4]): large-batch linearised-MPC sweep, 65k initial states x horizon 50, one GPU's
shard (1/8 of the states).  Per sweep: (1) the three GPs' mean / variance / Jacobian at every stage input -- the GP work
the config names, at N=2048, d=16 like config 2; (2) the MPC-side consumers on the reference's own model (12 states,
GPs with d = 1, 3, 3): RK4 stage dynamics + sensitivities for every stage, and the chance-constraint tightening
recursion per rollout from the stage variances."""
import os, sys, torch
import torch.distributed as dist
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E
from gpmpc_b200.mpc import ConstraintTightening, StageDynamics
from oracle import gp_oracle as o

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:      # one process per GPU, every rank its own 8192 rollouts (weak scaling, no collective in the data path)
    dist.init_process_group("nccl", device_id=dev)
R = int(sys.argv[1]) if len(sys.argv) > 1 else 8192       # rollouts on this GPU (65536 / 8)
T, N = 50, 2048
g = torch.Generator().manual_seed(5 + rank)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)      # max over ranks
    return float(t)


def say(msg):
    if rank == 0:
        print(msg + (f"  [x{world} GPUs, max over ranks: aggregate = {world} x the per-GPU rate]" if world > 1 else ""), flush=True)


# (1) GP mean / variance / Jacobian per stage, d = 16 (config-2 GPs)
p = o.synthetic_problem(N, 16, 8, p=3, seed=1337)
gb = E.fit_batch(p["X"].to(dev), p["Y"].to(dev), p["ell"], p["sf2"], p["noise"])
Z = ((torch.rand(R * T, 16, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
ms = timed(lambda: E.predict(gb, Z))
say(f"GP stage predictions: {R} rollouts x {T} stages x 3 GPs (N={N}, d=16): {ms:.1f} ms = {3 * R * T / ms * 1e3:.3e} preds/s/GPU")

# (2) MPC-side rows on the reference's model
gps = []
for k, d in enumerate((1, 3, 3)):
    X = torch.rand(N, d, generator=g, dtype=torch.float64) - 0.5
    y = 0.3 * torch.sin(4.0 * X.sum(dim=1)) + 0.01 * torch.randn(N, generator=g, dtype=torch.float64)
    gps.append((X, y, torch.full((d,), 0.3 + 0.1 * k, dtype=torch.float64), 0.5, 1e-2))
sd = StageDynamics(gps, [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653], 0.02)
x = ((torch.rand(R * T, 12, generator=g, dtype=torch.float64) - 0.5) * 0.8).to(dev)
u = ((torch.rand(R * T, 4, generator=g, dtype=torch.float64) - 0.3) * 0.6).to(dev)
ms = timed(lambda: sd.disc_dyn(x, u, work_bytes=8 << 30))
say(f"RK4 stage dynamics + A, B: {R * T} stages: {ms:.1f} ms = {R * T / ms * 1e3:.3e} stage evaluations/s/GPU")
A = torch.eye(12, dtype=torch.float64) + 0.02 * torch.randn(12, 12, generator=g, dtype=torch.float64)
B = 0.1 * torch.randn(12, 4, generator=g, dtype=torch.float64)
K = -0.3 * torch.randn(4, 12, generator=g, dtype=torch.float64)
Bd = torch.eye(12, dtype=torch.float64)[:, [1, 3, 5, 9, 10]]
ct = ConstraintTightening(gps, [[0], [1, 2, 3], [4, 5, 6]], A, B, K, Bd, 0.02, 2.0)
xp = ((torch.rand(R, 12, T + 1, generator=g, dtype=torch.float64) - 0.5) * 0.8).to(dev)
up = ((torch.rand(R, 4, T, generator=g, dtype=torch.float64) - 0.3) * 0.6).to(dev)
ms = timed(lambda: ct.propagate_constraint_limits(xp, up))
say(f"constraint tightening (3-GP variances at {R * T} stage inputs + recursion): {ms:.1f} ms = {R / ms * 1e3:.3e} rollouts/s/GPU")
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
