"""Config 1: the reference's own shapes -- P=3 GPs, d=(1,3,3), N in {15,30,45}, Adam lr 1e-3 <= 500
iterations from raw = 0, then Q=25 variance queries + mean/Jacobian at 25 stages.  Times the drop-in (GPU) and the
CPU oracle loop (torch float64, the reference's arithmetic restated) on the same data."""
import sys, time, torch
sys.path.insert(0, '.')
from oracle import gp_oracle as o

dims = (1, 3, 3)
g = torch.Generator().manual_seed(1)
use_gpu = torch.cuda.is_available()
if use_gpu:
    from gpmpc_b200.gp import GaussianProcess, fit_gp
for N in (15, 30, 45):
    X = torch.rand(N, 7, generator=g, dtype=torch.float64) - 0.5
    Y = 0.3 * torch.sin(3 * X[:, :3]) + 0.01 * torch.randn(N, 3, generator=g, dtype=torch.float64)
    idx = ([0], [1, 2, 3], [4, 5, 6])
    t0 = time.perf_counter()
    iters_cpu = 0
    for p in range(3):
        r = o.fit_hyperparameters(X[:, idx[p]].contiguous(), Y[:, p].contiguous(), n_train=500, lr=1e-3)
        iters_cpu += len(r["losses"])
    t_cpu = time.perf_counter() - t0
    msg = f"N={N}: CPU oracle loop {t_cpu*1e3:.0f} ms for {iters_cpu} Adam iterations ({t_cpu/iters_cpu*1e3:.3f} ms/iter)"
    if use_gpu:
        for rep in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            iters = 0
            for p in range(3):
                gp = GaussianProcess(X[:, idx[p]].contiguous(), Y[:, p].contiguous())
                fit_gp(gp, n_train=500, lr=1e-3, device="cuda:0")
                iters += len(gp.fit_losses)
            torch.cuda.synchronize()
            t_gpu = time.perf_counter() - t0
        msg += f" | drop-in on GPU {t_gpu*1e3:.0f} ms for {iters} iterations ({t_gpu/iters*1e3:.3f} ms/iter)"
        z = torch.rand(25, 16, generator=g, dtype=torch.float64) - 0.5
        gp.eval()
        t0 = time.perf_counter()
        for _ in range(20):
            cov = gp.likelihood(gp(z[:, idx[2]].to("cuda:0"))).covariance_matrix
            d = torch.diag(cov).cpu()
        msg += f"; eval-mode covariance diag (Q=25) {(time.perf_counter()-t0)/20*1e3:.3f} ms/call"
    print(msg, flush=True)
