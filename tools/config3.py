"""Config 3: exact GP, N=16384 train points, d=16, batched rollout queries with Jacobians.
Times a bounded query sample on one GPU (the 4M queries are independent: time scales linearly in Q) and checks
a subset against the DMMA fp64 path."""
import sys, time, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E

dev = torch.device('cuda', 0)
N, D, P = 16384, 16, 1
Q = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
g = torch.Generator().manual_seed(1337)
X = torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1
Y = torch.sin(X @ torch.randn(D, P, generator=g, dtype=torch.float64)) + 0.05 * torch.randn(N, P, generator=g, dtype=torch.float64)
Z = ((torch.rand(Q, D, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
t0 = time.perf_counter()
gb = E.fit_batch(X.to(dev), Y.to(dev), torch.full((D,), 2.0, dtype=torch.float64), 1.0, 1e-2)
torch.cuda.synchronize()
print(f"fit state (potrf + trtri + alpha + digit image) N={N}: {time.perf_counter() - t0:.3f} s, digits={gb.digits}")
for _ in range(2):
    out = E.predict(gb, Z)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    out = E.predict(gb, Z)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(f"predict Q={Q}: {ms:.2f} ms -> {P * Q / ms * 1e3:.3e} preds/s/GPU; 4M queries would take {4194304 / Q * ms / 1e3:.2f} s on 1 GPU")
print(f"variance contraction fp64-equivalent rate >= {N * N * Q / (ms * 1e-3) / 1e12:.1f} TFLOP/s (whole step)")
sub = Z[:2048]
m1, v1, j1 = E.predict(gb, sub)
m2, v2, j2 = E.predict(gb, sub, var_path="f64")
rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
print("i8 vs DMMA path: mean", rel(m1, m2), "var", rel(v1, v2), "jac", rel(j1, j2))
