"""Config 4: hyper-parameter fit at large N, one process per GPU (torchrun or plain python for 1 GPU).
  (a) output GPs split across the GPUs: every rank runs MLL + gradient iterations of its own GP (no collective);
      aggregate iterations/s = sum over ranks.
  (b) ONE GP, gradient split by rows: every rank factorises (Cholesky / inverse factor are replicated), forms only
      its rows of K^-1 and of the trace; an all-reduce of d+2 doubles completes the gradient.
Usage: [torchrun --nproc-per-node G ...] tools/config4.py [N=32768] [iters=3]"""
import os, sys, time, torch
import torch.distributed as dist
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E, sharding as S

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
ITERS = int(sys.argv[2]) if len(sys.argv) > 2 else 3
D = 16
ell = torch.full((D,), 2.0, dtype=torch.float64)


def data(seed):
    g = torch.Generator().manual_seed(seed)
    X = (torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    y = (torch.sin(X @ torch.randn(D, generator=g, dtype=torch.float64).to(dev))
         + 0.05 * torch.randn(N, generator=g, dtype=torch.float64).to(dev))
    return X, y


def timed(fn):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(ITERS):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / ITERS], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t), out


# (a) one GP per rank (different targets per rank)
X, y = data(1337 + rank)
def own_iter():
    fit, ws = E.factorize(X, y, ell, 1.0, 1e-2, want_kinv=True, want_image=False)
    return E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
ms, r = timed(own_iter)
if rank == 0:
    print(f"(a) N={N}, {world} GP(s) on {world} GPU(s), {E.FIT_GEMM_DIGITS}-digit int8 GEMMs: {ms:.1f} ms per iteration (max over ranks) "
          f"-> {world / ms * 1e3:.2f} MLL+grad iterations/s aggregate, {N**3 / ms * 1e-9:.1f} TFLOP/s of N^3 per GPU; loss {r[0]:.6f}",
          flush=True)

# (b) one GP (rank 0's data everywhere), K^-1 rows + trace split over the ranks
X, y = data(1337)
rs = S.rowsplit_grad_reduce()
def split_iter():
    fit, ws = E.factorize(X, y, ell, 1.0, 1e-2, want_kinv=rs.rows(N), want_image=False)
    return E.loss_and_grad_from_partials(fit, rs(fit, ws))
ms_b, rb = timed(split_iter)
ref = own_iter() if rank == 0 else None          # rank 0: the unsplit gradient of the same GP
if rank == 0:
    gs = torch.tensor([rb[1], rb[2]] + list(rb[3])); gr = torch.tensor([ref[1], ref[2]] + list(ref[3]))
    print(f"(b) N={N}, one GP, K^-1 rows + gradient trace split over {world} GPU(s) (rows of rank 0: {rs.rows(N)}): "
          f"{ms_b:.1f} ms per iteration -> {1e3 / ms_b:.2f} iterations/s ({ms / ms_b:.2f}x the unsplit iteration); "
          f"gradient vs unsplit: {float((gs - gr).abs().max() / gr.abs().max()):.1e}", flush=True)
if world > 1:
    dist.destroy_process_group()
