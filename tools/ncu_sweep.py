"""One launch (or a few) of every kernel of the library on shapes of the benchmark configs, for an
`ncu --set full` sweep (profiles/kernels_r01.md is generated from its raw CSV by tools/ncu_table.py).
The profiled region is bracketed by cudaProfilerStart/Stop: run ncu with --profile-from-start off."""
import sys
import torch
sys.path.insert(0, ".")
from gpmpc_b200 import engine as E, _lib
from gpmpc_b200._lib import call, ptr
from gpmpc_b200.mpc import StageDynamics

dev = torch.device("cuda", 0)
st = _lib.stream_ptr(dev)
lib = _lib.load()
g = torch.Generator().manual_seed(1337)
N, D, Q = 2048, 16, 8192
X = (torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
Y = (torch.sin(X.cpu() @ torch.randn(D, 3, generator=g, dtype=torch.float64)) + 0.05 * torch.randn(N, 3, generator=g, dtype=torch.float64)).to(dev)
Z = ((torch.rand(Q, D, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
ell = torch.full((D,), 2.0, dtype=torch.float64)

# warm-up of everything (module load, scratch, caches) outside the profiled region
E.FIT_GEMM_DIGITS = 0
gb = E.fit_batch(X, Y, ell, 1.0, 1e-2)
E.predict(gb, Z); E.predict(gb, Z.float()); E.predict(gb, Z, want_var=False); E.predict(gb, Z.float(), want_var=False)
E.predict(gb, Z[:2048], var_path="f64")
fit, ws = E.factorize(X, Y[:, 0], ell, 1.0, 1e-2, want_kinv=True, want_image=False)
E.grad_partials(fit, ws)
E.factorize(X[:384], Y[:384, 0], ell, 1.0, 1e-2, want_kinv=True)
fit, ws = E.factorize(X, Y[:, 0], ell, 1.0, 1e-2, want_kinv=True, want_image=False)
M = 4096
A = torch.randn(M, M, dtype=torch.float64, device=dev); B = torch.randn(M, M, dtype=torch.float64, device=dev)
C = torch.zeros(M, M, dtype=torch.float64, device=dev)
wb = lib.gpk_oz_dgemm_nt_work_bytes(M, M, M, 0, 6)
work = torch.empty(wb, dtype=torch.uint8, device=dev)
def gemms():
    call("gpk_dgemm_nt", ptr(A), M, ptr(B), M, ptr(C), M, M, M, M, 1.0, 0.0, 0, 0, st)
    call("gpk_oz_dgemm_nt", ptr(A), M, ptr(B), M, ptr(C), M, M, M, M, 1.0, 0.0, 0, 0, 6, ptr(work), wb, st)
gemms()
gen = torch.Generator().manual_seed(0)
gps = []
for k, d in enumerate((1, 3, 3)):
    Xs = torch.rand(N, d, generator=gen, dtype=torch.float64) - 0.5
    ys = 0.3 * torch.sin(4.0 * Xs.sum(dim=1)) + 0.01 * torch.randn(N, generator=gen, dtype=torch.float64)
    gps.append((Xs, ys, torch.full((d,), 0.3 + 0.1 * k, dtype=torch.float64), 0.5, 1e-2))
sd = StageDynamics(gps, [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653, 9.81], 0.02)
xs = ((torch.rand(65536, 12, generator=gen, dtype=torch.float64) - 0.5) * 0.8).to(dev)
us = ((torch.rand(65536, 4, generator=gen, dtype=torch.float64) - 0.3) * 0.6).to(dev)
sd.disc_dyn(xs, us, work_bytes=8 << 30)
torch.cuda.synchronize()

torch.cuda.profiler.start()
E.predict(gb, Z)                                   # oz_front_kernel<double,MJ>, oz_mj_finalize, oz_var_kernel<5>, var_finalize
E.predict(gb, Z.float())                           # float32 queries: oz_front_kernel<float>, oz_var_kernel<4>
E.predict(gb, Z, want_var=False)                   # DMMA front kernel without the digit image
E.predict(gb, Z.float(), want_var=False)           # SIMT mean_jac_kernel<float>
E.predict(gb, Z[:2048], var_path="f64")            # cross_kernel + var_colsq_kernel (DMMA variance path)
# fit kernels at N = 384 (3 leaves: every kernel of the recursion once or a few times, not 190 launches) ...
fit_s, ws_s = E.factorize(X[:384], Y[:384, 0], ell, 1.0, 1e-2, want_kinv=True)   # cross, potf2, leaf_inv, dgemm_nt32/16x128, transposes, trmv, slice_w
# ... and the N^2 kernels of the iteration at N = 2048
E.dense_K(fit)                                     # cross_kernel (K assembly)
call("gpk_kinv", ptr(ws.U), fit.npad, fit.npad, ptr(ws.need_kinv()), fit.npad, st)   # dgemm_nt (upper-triangular A, lower tiles) + mirror
E.grad_partials(fit, ws)                           # grad_trace_kernel
gemms()                                            # dgemm_nt_kernel (128x128 tiles) and the int8 digit GEMM at 4096^3
sd.disc_dyn(xs, us, work_bytes=8 << 30)            # stage_step_kernel + mean_jac_kernel<double> (d <= 8)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("sweep done")
