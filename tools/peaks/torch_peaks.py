"""cuBLAS GEMM peaks through torch (fp64 DGEMM, fp32 SGEMM, TF32) on the current GPU. Prints one JSON line."""
import json, torch
def gemm_tflops(dtype, n, tf32=False, reps=5):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype); b = torch.randn(n, n, device="cuda", dtype=dtype)
    torch.matmul(a, b); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12
out = {"dgemm_4096": gemm_tflops(torch.float64, 4096), "dgemm_8192": gemm_tflops(torch.float64, 8192),
       "sgemm_8192": gemm_tflops(torch.float32, 8192), "tf32_8192": gemm_tflops(torch.float32, 8192, tf32=True)}
# sustained DGEMM for ~2 s
import time
a = torch.randn(8192, 8192, device="cuda", dtype=torch.float64); b = torch.randn_like(a)
torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); n = 0
for _ in range(60): torch.matmul(a, b); n += 1
e1.record(); torch.cuda.synchronize()
out["dgemm_8192_sustained"] = 2.0 * 8192 ** 3 * n / (e0.elapsed_time(e1) * 1e-3) / 1e12
print(json.dumps(out))
