"""Accuracy and timing of the tcgen05 int8 digit GEMM (gpk_oz_dgemm_nt) against the DMMA GEMM (gpk_dgemm_nt)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gpmpc_b200 import _lib
from gpmpc_b200._lib import call, ptr

dev = torch.device("cuda", 0)
lib = _lib.load()
st = _lib.stream_ptr(dev)
torch.manual_seed(0)


def run(M, N, K, a_tri, c_lower, same, digits, alpha=-1.0, beta=1.0, decay=False, reps=3):
    A = torch.randn(M, K, dtype=torch.float64, device=dev)
    if decay:   # rows whose magnitude decays along k, like a Cholesky factor
        A *= torch.logspace(0, -6, K, dtype=torch.float64, device=dev)[None, :]
    if a_tri == 1: A = torch.tril(A)
    if a_tri == 2: A = torch.triu(A)
    B = A if same else torch.randn(N, K, dtype=torch.float64, device=dev)
    C0 = torch.randn(M, N, dtype=torch.float64, device=dev)
    ref = C0.clone()
    call("gpk_dgemm_nt", ptr(A), K, ptr(B), K, ptr(ref), N, M, N, K, alpha, beta, a_tri, c_lower, st)
    wb = lib.gpk_oz_dgemm_nt_work_bytes(M, N, K, int(same), digits)
    work = torch.empty(wb, dtype=torch.uint8, device=dev)
    out = C0.clone()
    call("gpk_oz_dgemm_nt", ptr(A), K, ptr(B), K, ptr(out), N, M, N, K, alpha, beta, a_tri, c_lower, digits, ptr(work), wb, st)
    torch.cuda.synchronize()
    if c_lower:
        # the DMMA kernel skips upper tiles at ITS tile size (32/64/128), this one at 128: compare the true lower triangle
        mask = torch.ones(M // 128, N // 128, device=dev).tril().bool().repeat_interleave(128, 0).repeat_interleave(128, 1)
        err = (out - ref).tril().abs().max().item()
        untouched = ((out - C0).abs() * (~mask)).max().item()
    else:
        err = (out - ref).abs().max().item(); untouched = 0.0
    exact = beta * C0 + alpha * (A @ B.T) if max(M, N, K) <= 4096 else None
    scale = ref.abs().max().item()
    ts = []
    for fn in ("gpk_oz_dgemm_nt", "gpk_dgemm_nt"):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(reps):
            e0.record()
            if fn == "gpk_oz_dgemm_nt":
                call(fn, ptr(A), K, ptr(B), K, ptr(out), N, M, N, K, alpha, beta, a_tri, c_lower, digits, ptr(work), wb, st)
            else:
                call(fn, ptr(A), K, ptr(B), K, ptr(ref), N, M, N, K, alpha, beta, a_tri, c_lower, st)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        ts.append(best)
    flops = 2.0 * M * N * K * (0.5 if a_tri else 1.0) * (0.5 if c_lower else 1.0)
    print(f"M={M} N={N} K={K} tri={a_tri} low={c_lower} same={int(same)} S={digits} decay={int(decay)}: "
          f"|oz-dmma|/max={err/scale:.2e} untouched={untouched:.1e} oz {ts[0]:.3f} ms ({flops/ts[0]*1e-9:.1f} TF eq) "
          f"dmma {ts[1]:.3f} ms ({flops/ts[1]*1e-9:.1f} TF)", flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "one":      # one shape, e.g. for ncu: one M N K tri low same S
        M, N, K, tri, low, same, S = [int(a) for a in sys.argv[2:9]]
        run(M, N, K, tri, low, bool(same), S, reps=2)
        sys.exit(0)
    big = len(sys.argv) > 1 and sys.argv[1] == "big"
    for S in (5, 6, 7):
        run(256, 256, 1024, 0, 0, False, S)
        run(512, 384, 4096, 0, 0, False, S, alpha=1.0, beta=0.0)
        run(1024, 1024, 1024, 1, 0, False, S)
        run(1024, 1024, 1024, 2, 0, False, S, beta=0.0)
        run(1024, 1024, 2048, 0, 1, True, S)
        run(2048, 2048, 2048, 2, 1, True, S, alpha=1.0, beta=0.0)
        run(2048, 2048, 4096, 0, 0, False, S, decay=True)
    if big:
        for S in (5, 6, 7):
            run(4096, 4096, 4096, 0, 1, True, S)
            run(8192, 8192, 8192, 0, 0, False, S)
            run(8192, 8192, 8192, 2, 1, True, S, alpha=1.0, beta=0.0)
            run(16384, 16384, 16384, 0, 1, True, S)
