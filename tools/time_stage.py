"""Times StageDynamics.disc_dyn (4 x [3-GP mean+Jacobian launch + stage kernel]): python tools/time_stage.py [Q] [N]"""
import sys
import torch
sys.path.insert(0, ".")
from gpmpc_b200.mpc import StageDynamics
from gpmpc_b200 import _lib

Q = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
gen = torch.Generator().manual_seed(0)
gps = []
for k, d in enumerate((1, 3, 3)):
    X = torch.rand(N, d, generator=gen, dtype=torch.float64) - 0.5
    y = 0.3 * torch.sin(4.0 * X.sum(dim=1)) + 0.01 * torch.randn(N, generator=gen, dtype=torch.float64)
    gps.append((X, y, torch.full((d,), 0.3 + 0.1 * k, dtype=torch.float64), 0.5, 1e-2))
prior = [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653, 9.81]
sd = StageDynamics(gps, prior, 0.02)
x = ((torch.rand(Q, 12, generator=gen, dtype=torch.float64) - 0.5) * 0.8).cuda()
u = ((torch.rand(Q, 4, generator=gen, dtype=torch.float64) - 0.3) * 0.6).cuda()
for _ in range(2):
    sd.disc_dyn(x, u, work_bytes=8 << 30)
lib = _lib.load()
lib.gpk_reset_launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
reps = 5
for _ in range(reps):
    sd.disc_dyn(x, u, work_bytes=8 << 30)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
pairs = 9 * Q * N   # three GPs at the first sub-step, two at the other three
print(f"Q={Q} N={N}: {ms:.3f} ms/call  {Q / ms * 1e3:.3e} stage evals/s  {pairs / ms / 1e6:.1f} G kernel pairs/s  "
      f"launches/call {lib.gpk_launch_count() / reps:.0f}")
