"""MLL + gradient iteration with the large GEMMs on the DMMA pipe (digits 0) and on the tcgen05 int8 pipe
(5/6/7 digits): time per iteration and deviation of loss / gradient / alpha / W from the DMMA result."""
import sys, time, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E
dev = torch.device('cuda', 0)
D = 16
noise = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-2
E.FIT_GEMM_MIN_WORK_G = int(sys.argv[3]) if len(sys.argv) > 3 else 0
DIGITS = [int(a) for a in sys.argv[4].split(',')] if len(sys.argv) > 4 else [0, 5, 6, 7]
E.FIT_OVERLAP_INVERSE = bool(int(sys.argv[5])) if len(sys.argv) > 5 else E.FIT_OVERLAP_INVERSE
E.FIT_GEMM_DIGITS_KINV = int(sys.argv[6]) if len(sys.argv) > 6 else E.FIT_GEMM_DIGITS_KINV
for N in [int(a) for a in sys.argv[1].split(',')]:
    g = torch.Generator().manual_seed(1337)
    X = (torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    y = (torch.sin(X @ torch.randn(D, generator=g, dtype=torch.float64).to(dev))
         + 0.05 * torch.randn(N, generator=g, dtype=torch.float64).to(dev))
    ell = torch.full((D,), 2.0, dtype=torch.float64)
    ref = None
    for digits in DIGITS:
        E.FIT_GEMM_DIGITS = digits
        E.FIT_GEMM_MIN_NPAD = 1024
        E.release_scratch()
        def it():
            fit, ws = E.factorize(X, y, ell, 1.0, noise, want_kinv=True, want_image=False)
            return fit, E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
        for _ in range(2): fit, r = it()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = 3 if N >= 16384 else 10
        for _ in range(n): fit, r = it()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / n
        vec = torch.tensor([r[0], r[1], r[2]] + list(r[3]), dtype=torch.float64)
        cur = dict(vec=vec, alpha=fit.alpha.clone(), W=fit.W.clone() if N <= 16384 else None)
        msg = ""
        if ref is None:
            ref = cur
        else:
            dl = abs(vec[0] - ref["vec"][0]) / abs(ref["vec"][0])
            dg = (vec[1:] - ref["vec"][1:]).abs().max() / ref["vec"][1:].abs().max()
            da = (cur["alpha"] - ref["alpha"]).abs().max() / ref["alpha"].abs().max()
            dw = ((cur["W"] - ref["W"]).abs().max() / ref["W"].abs().max()).item() if cur["W"] is not None else float('nan')
            msg = f" | vs DMMA: loss {dl:.1e} grad {dg:.1e} alpha {da:.1e} W {dw:.1e}"
        print(f"N={N} noise={noise:g} digits={digits}: {dt*1e3:.2f} ms/iter ({N**3/dt/1e12:.1f} TFLOP/s of N^3) loss {r[0]:.9f}{msg}", flush=True)
        del fit
    E.release_scratch()
