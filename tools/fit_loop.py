"""Times the body of fit_gp's loop (loss + gradient on the GPU, Adam step on the host) without the early stop:
python tools/fit_loop.py [N] [iters]     (reference regime: N <= 45, d = 3)"""
import sys, time, torch
sys.path.insert(0, '.')
from gpmpc_b200 import GaussianProcess
from gpmpc_b200 import engine as E
from gpmpc_b200.gp import _loss_and_set_grads
N = int(sys.argv[1]) if len(sys.argv) > 1 else 45
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 100
g = torch.Generator().manual_seed(0)
X = torch.rand(N, 3, generator=g, dtype=torch.float64) - 0.5
y = torch.sin(3 * X.sum(1)) + 0.01 * torch.randn(N, generator=g, dtype=torch.float64)
gp = GaussianProcess(X, y).train()
dev = torch.device("cuda", 0)
optim = torch.optim.Adam(gp.parameters(), lr=0.01)
W = torch.empty(E.round_up(N, 128), E.round_up(N, 128), dtype=torch.float64, device=dev)


def step():
    optim.zero_grad()
    loss, _ = _loss_and_set_grads(gp, dev, W)
    optim.step()
    gp._invalidate()
    return loss


for _ in range(5):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(iters):
    loss = step()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"N={N}: {dt / iters * 1e3:.3f} ms per fit_gp iteration ({iters / dt:.0f} it/s), loss {loss:.6f}")
