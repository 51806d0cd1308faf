"""One MLL + gradient iteration (fit_gp's inner step) at N training points: timing and (under ncu) launch list."""
import sys, time, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E
dev = torch.device('cuda', 0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
D = 16
g = torch.Generator().manual_seed(1337)
X = (torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
y = torch.sin(X @ torch.randn(D, generator=g, dtype=torch.float64).to(dev))
ell = torch.full((D,), 2.0, dtype=torch.float64)
def it():
    fit, ws = E.factorize(X, y, ell, 1.0, 1e-2, want_kinv=True, want_image=False)
    return E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
for _ in range(2): it()
torch.cuda.synchronize()
torch.cuda.profiler.start()
t0 = time.perf_counter()
n = int(sys.argv[2]) if len(sys.argv) > 2 else 5
for _ in range(n): r = it()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
dt = (time.perf_counter() - t0) / n
print(f"N={N}: {dt*1e3:.2f} ms per MLL+grad iteration ({1/dt:.1f} iters/s); N^3 flops -> {N**3/dt/1e12:.2f} TFLOP/s; loss {r[0]:.6f}")
