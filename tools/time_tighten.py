"""Times gpk_tighten_constraints (one CTA per rollout) on synthetic rollouts: python tools/time_tighten.py [R] [T]"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from gpmpc_b200 import _lib
from gpmpc_b200._lib import call, ptr

R = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
T = int(sys.argv[2]) if len(sys.argv) > 2 else 25
nx, nu = 12, 4
dev = torch.device("cuda", 0)
g = torch.Generator(device="cpu").manual_seed(0)
A = (torch.eye(nx, dtype=torch.float64) + 0.02 * torch.randn(nx, nx, generator=g, dtype=torch.float64)).to(dev)
B = (0.1 * torch.randn(nx, nu, generator=g, dtype=torch.float64)).to(dev)
K = (-0.3 * torch.randn(nu, nx, generator=g, dtype=torch.float64)).to(dev)
Bd = torch.eye(nx, dtype=torch.float64)[:, [1, 3, 5, 9, 10]].contiguous().to(dev)
z = (torch.rand(R * T, nx + nu, generator=g, dtype=torch.float64) - 0.5).to(dev)
var = (torch.rand(3, R * T, generator=g, dtype=torch.float64) * 1e-2).to(dev)
hyp = torch.zeros(3, _lib.load().gpk_hyp_stride(), dtype=torch.float64, device=dev)
hyp[:, 65] = 1e-3
sc = torch.empty(R, 2 * nx, T + 1, dtype=torch.float64, device=dev)
ic = torch.empty(R, 2 * nu, T, dtype=torch.float64, device=dev)
st = torch.cuda.current_stream().cuda_stream


def run():
    call("gpk_tighten_constraints", ptr(var), var.stride(0), 0, ptr(z), z.stride(0), 6, 7, ptr(hyp), ptr(A), ptr(B),
         ptr(K), ptr(Bd), nx, nu, T, R, 0.02, 2.0, ptr(sc), ptr(ic), st)


for _ in range(3):
    run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(20):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
flops = R * T * (2 * 2 * nx**3 + 2 * nu * nx * nx)
byts = R * T * (5 * 8) + sc.numel() * 8 + ic.numel() * 8
print(f"R={R} T={T}: {ms:.3f} ms  {R / ms * 1e3:.3e} rollouts/s  {flops / ms / 1e9:.2f} TFLOP/s fp64  {byts / ms / 1e6:.1f} GB/s")
