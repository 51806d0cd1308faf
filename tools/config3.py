"""Config 3: exact GP, N=16384 train points, d=16, batched rollout queries with Jacobians, query-sharded
over the GPUs of one box (one process per GPU; plain `python` = 1 GPU).  Every rank predicts its Q/G block (state
replicated: every rank fits the same GP), no collective in the data path; time = max over ranks.
Usage: [torchrun --nproc-per-node G ...] tools/config3.py [Q_total=262144]"""
import os, sys, time, torch
import torch.distributed as dist
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E, sharding as S

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
N, D, P = 16384, 16, 1
Q = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
lo, hi = S.shard_bounds(Q, world, rank)
g = torch.Generator().manual_seed(1337)
X = torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1
Y = torch.sin(X @ torch.randn(D, P, generator=g, dtype=torch.float64)) + 0.05 * torch.randn(N, P, generator=g, dtype=torch.float64)
g2 = torch.Generator().manual_seed(1337 + rank)
Z = ((torch.rand(hi - lo, D, generator=g2, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
E.fit_batch(X.to(dev), Y.to(dev), torch.full((D,), 2.0, dtype=torch.float64), 1.0, 1e-2)      # warm-up (allocations)
torch.cuda.synchronize()
t0 = time.perf_counter()
gb = E.fit_batch(X.to(dev), Y.to(dev), torch.full((D,), 2.0, dtype=torch.float64), 1.0, 1e-2)
torch.cuda.synchronize()
if rank == 0:
    print(f"fit state (potrf + trtri + alpha + digit image) N={N}: {time.perf_counter() - t0:.3f} s, digits={gb.digits}", flush=True)
out = E.predict(gb, Z[: min(65536, hi - lo)])
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
REPS = 2
e0.record()
for _ in range(REPS):
    out = E.predict(gb, Z)
e1.record()
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / REPS], device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
ms = float(t)
if rank == 0:
    print(f"predict Q={Q} on {world} GPU(s) ({hi - lo} per GPU): {ms:.1f} ms (max over ranks) -> {P * Q / ms * 1e3:.3e} preds/s aggregate; "
          f"4M queries would take {4194304 / Q * ms / 1e3:.2f} s", flush=True)
    print(f"variance contraction fp64-equivalent rate >= {N * N * (hi - lo) / (ms * 1e-3) / 1e12:.1f} TFLOP/s per GPU (whole step)")
    sub = Z[:2048]
    m1, v1, j1 = E.predict(gb, sub)
    m2, v2, j2 = E.predict(gb, sub, var_path="f64")
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
    print("i8 vs DMMA path: mean", rel(m1, m2), "var", rel(v1, v2), "jac", rel(j1, j2))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
