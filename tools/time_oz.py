"""Times the pieces of the int8 variance path separately (CUDA events, L2 flushed between launches)."""
import sys, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E, _lib
from gpmpc_b200._lib import call, ptr

dev = torch.device('cuda', 0)
N, D, Q, P = 2048, 16, 8192, 3
S = int(sys.argv[1]) if len(sys.argv) > 1 else 6
g = torch.Generator().manual_seed(1337)
X = torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1
Y = torch.sin(X @ torch.randn(D, P, generator=g, dtype=torch.float64))
Z = ((torch.rand(Q, D, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
gb = E.fit_batch(X.to(dev), Y.to(dev), torch.full((D,), 2.0, dtype=torch.float64), 1.0, 1e-2, digits=S)
lib = _lib.load()
npad = gb.npad
qpad = E.round_up(Q, 128)
st = _lib.stream_ptr(dev)
eimg = torch.empty(P, S * qpad * npad, dtype=torch.int8, device=dev)
pieces = int(sys.argv[2]) if len(sys.argv) > 2 else lib.gpk_oz_pieces(qpad // 128 * P, npad)
part = torch.empty(P, 16, qpad, dtype=torch.float64, device=dev)
flushbuf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def cross():
    call("gpk_oz_cross", 1, ptr(Z), D, Q, 0, ptr(gb.X), gb.X.stride(1), N, gb.X.stride(0), D, ptr(gb.hyp), ptr(eimg),
         eimg.stride(0), npad, qpad, S, P, st)

def var():
    call("gpk_oz_var_partials", ptr(eimg), eimg.stride(0), ptr(gb.wimg), gb.wimg.stride(0), ptr(gb.wscale),
         gb.wscale.stride(0), ptr(gb.hyp), npad, qpad, S, ptr(part), qpad, 16 * qpad, pieces, P, None, st)

def mj():
    E.predict(gb, Z, want_var=False)

def timeit(fn, n=20, flush=True):
    for _ in range(3): fn()
    evs = []
    for _ in range(n):
        if flush: flushbuf.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); evs.append((e0, e1))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]

cross()
print(f"S={S} pieces={pieces}")
print("cross   ms", timeit(cross))
tv = timeit(var)
ops = 2.0 * S * (S + 1) / 2 * (N * N / 2) * Q * P
print("var     ms", tv, " int8 TOP/s (algorithmic triangular)", ops / tv / 1e9, " fp64-equiv TFLOP/s", N * N * Q * P / tv / 1e9)
print("var warm ms", timeit(var, flush=False))
print("mean_jac ms", timeit(mj))
print("predict ms", timeit(lambda: E.predict(gb, Z)))
