"""Mean + Jacobian only (engine.predict(..., want_var=False)) at config 2: float32 SIMT kernel and float64 DMMA front kernel."""
import sys, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E
dev = torch.device('cuda', 0)
N, D, Q, P = 2048, 16, 8192, 3
g = torch.Generator().manual_seed(1337)
X = (torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
Y = (torch.sin(X.cpu() @ torch.randn(D, P, generator=g, dtype=torch.float64)) + 0.05 * torch.randn(N, P, generator=g, dtype=torch.float64)).to(dev)
Z = ((torch.rand(Q, D, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
gb = E.fit_batch(X, Y, torch.full((D,), 2.0, dtype=torch.float64), 1.0, 1e-2)
ref = E.predict(gb, Z, want_var=False)
for name, zz in (("f32", Z.float()), ("f64", Z)):
    for _ in range(3):
        out = E.predict(gb, zz, want_var=False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20):
        out = E.predict(gb, zz, want_var=False)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    rel = lambda a, b: float((a.double() - b).abs().max() / b.abs().max())
    print(f"{name}: {ms:.4f} ms -> {P*Q/ms*1e3:.3e} preds/s, {P*Q*N/ms*1e3:.3e} pairs/s; vs f64: mean {rel(out[0], ref[0]):.1e} jac {rel(out[2], ref[2]):.1e}")
