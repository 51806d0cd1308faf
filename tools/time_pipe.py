"""Pipelined prediction step (front end of sub-chunk c+1 co-resident with the contraction of sub-chunk c) against the
sequential one: config-2 and config-3 shapes, CUDA events, L2 flushed between timed calls.
usage: python tools/time_pipe.py [N] [Q] [P]"""
import os, sys, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E, _lib

dev = torch.device('cuda', 0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
P = int(sys.argv[3]) if len(sys.argv) > 3 else 3
D = 16
g = torch.Generator().manual_seed(1337)
X = torch.rand(N, D, generator=g, dtype=torch.float64) * 2 - 1
Y = torch.sin(X @ torch.randn(D, P, generator=g, dtype=torch.float64))
Z = ((torch.rand(Q, D, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
gb = E.fit_batch(X.to(dev), Y.to(dev), torch.full((D,), 2.0, dtype=torch.float64), 1.0, 1e-2, digits=5)
flushbuf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, n=20):
    for _ in range(3): fn()
    evs = []
    for _ in range(n):
        flushbuf.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); evs.append((e0, e1))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]

def run(**env):
    for k in ("GPK_OZ_PIPELINE", "GPK_OZ_PIPE_TSUB", "GPK_OZ_PIPE_L", "GPK_OZ_PIPE_CO", "GPK_OZ_FRONT_CO"):
        os.environ.pop(k, None)
    for k, v in env.items():
        os.environ[k] = str(v)
    out = E.predict(gb, Z)
    torch.cuda.synchronize()
    return out, timeit(lambda: E.predict(gb, Z))

ref, t0 = run(GPK_OZ_PIPELINE=0)
print(f"N={N} Q={Q} P={P}  sequential: median {t0[0]:.4f} ms  min {t0[1]:.4f} ms")
out, tc = run(GPK_OZ_PIPELINE=0, GPK_OZ_FRONT_CO=1)
print(f"sequential, co-variant front alone: median {tc[0]:.4f} ms  min {tc[1]:.4f} ms")
out, t1 = run()
print(f"model plan          : median {t1[0]:.4f} ms  min {t1[1]:.4f} ms")
tiles = (Q + 127) // 128
cands = sorted({t for t in (2, 4, 8, 12, 16, 21, 24, 32, 37, 48, 64, 74) if t < tiles})
for t in cands:
    try:
        out, tt = run(GPK_OZ_PIPELINE=2, GPK_OZ_PIPE_TSUB=t)
    except Exception as e:
        print(f"tsub={t}: {e}")
        continue
    dev_ = max(float((a - b).abs().max() / b.abs().max()) for a, b in zip(out, ref))
    print(f"pipelined tsub={t:3d}: median {tt[0]:.4f} ms  min {tt[1]:.4f} ms   max rel dev vs sequential {dev_:.2e}")
