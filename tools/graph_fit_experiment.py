"""Does a CUDA graph of one MLL + gradient iteration (assembly, potrf + trtri on two streams, alpha, K^-1, trace) beat
the stream launches?  Captures the C-ABI calls of engine.factorize / grad_partials with torch.cuda.CUDAGraph."""
import sys, time, torch
sys.path.insert(0, '.')
import bench
from gpmpc_b200 import engine as E, _lib
from gpmpc_b200._lib import call, ptr

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
E.init(dev)
lib = _lib.load()
for n in (512, 1024, 2048, 4096, 8192):
    X, y = bench.fit_data(n, 1337, dev)
    d = X.shape[1]
    ell = torch.full((d,), 2.0, dtype=torch.float64)
    def plain():
        fit, ws = E.factorize(X, y, ell, 1.0, 1e-2, want_kinv=True, want_image=False)
        return E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))
    for _ in range(3): ref = plain()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20): plain()
    torch.cuda.synchronize()
    t_plain = (time.perf_counter() - t0) / 20
    # ---- graph ----
    npad = E.round_up(n, 128)
    ws = E.get_fit_scratch(npad, dev)
    W = torch.empty(npad, npad, dtype=torch.float64, device=dev)
    alpha = torch.empty(npad, dtype=torch.float64, device=dev)
    hyp = E.make_hyp(X.mean(0), ell, 1.0, 1e-2, d, dev)
    ws.y_pad.zero_(); ws.y_pad[:n] = y
    elems = lib.gpk_grad_trace_work_elems(n, d, 0, n)
    work = torch.empty(elems, dtype=torch.float64, device=dev)
    Kinv = ws.need_kinv()
    def enqueue():
        st = _lib.stream_ptr(dev)
        call("gpk_cross_kernel", 1, ptr(X), X.stride(0), n, 0, ptr(X), X.stride(0), n, 0, d, ptr(hyp), ptr(ws.L), npad, 0, npad, npad, 1, 1, st)
        call("gpk_potrf_trtri", ptr(ws.L), npad, npad, ptr(ws.dinv), ptr(ws.info), ptr(W), npad, ptr(ws.U), npad, ptr(ws.T), ws.T.numel(), st)
        call("gpk_solve_alpha", ptr(W), npad, ptr(ws.U), npad, ptr(ws.L), npad, npad, ptr(ws.y_pad), ptr(ws.t_work), ptr(alpha), ptr(ws.scal), st)
        call("gpk_kinv", ptr(ws.U), npad, npad, ptr(Kinv), npad, st)
        call("gpk_mll_grad_reduce_sym", ptr(X), X.stride(0), n, d, ptr(hyp), ptr(Kinv), npad, ptr(alpha), 0, n, ptr(work), elems, ptr(ws.grad_out), st)
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        for _ in range(2): enqueue()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, stream=s):
            enqueue()
    except Exception as e:
        print(f"N={n}: capture failed: {str(e)[:300]}")
        continue
    torch.cuda.synchronize()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        g.replay()
        out = torch.cat([ws.scal, ws.grad_out[:18]]).tolist()      # the one host round trip
    torch.cuda.synchronize()
    t_graph = (time.perf_counter() - t0) / 20
    quad, logdet = out[0], out[1]
    loss = 0.5 * (quad + logdet + n * E.LOG_2PI) / n
    print(f"N={n}: stream launches {t_plain*1e3:.3f} ms/iter ({1/t_plain:.0f} it/s), graph replay {t_graph*1e3:.3f} ms/iter ({1/t_graph:.0f} it/s); "
          f"loss {loss:.12f} vs {ref[0]:.12f}", flush=True)
    E.release_scratch()
