"""Per-kernel trace of the pipelined prediction step (GPK_OZ_PIPE_TRACE): python tools/trace_pipe.py N Q P tsub"""
import os, sys, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E
N, Q, P, T = (int(x) for x in sys.argv[1:5])
dev = torch.device('cuda', 0)
g = torch.Generator().manual_seed(1337)
X = torch.rand(N, 16, generator=g, dtype=torch.float64) * 2 - 1
Y = torch.sin(X @ torch.randn(16, P, generator=g, dtype=torch.float64))
Z = ((torch.rand(Q, 16, generator=g, dtype=torch.float64) * 2 - 1) * 1.2).to(dev)
gb = E.fit_batch(X.to(dev), Y.to(dev), torch.full((16,), 2.0, dtype=torch.float64), 1.0, 1e-2, digits=5)
os.environ["GPK_OZ_PIPELINE"] = "2"; os.environ["GPK_OZ_PIPE_TSUB"] = str(T)
for _ in range(3): E.predict(gb, Z)
torch.cuda.synchronize()
os.environ["GPK_OZ_PIPE_TRACE"] = "1"
E.predict(gb, Z); E.predict(gb, Z)
