"""bench.py -- GP posterior predictions/s (mean + variance + Jacobian) of the gp-mpc GP hot path on B200.

Workload (BASELINE.json configs[1], SURVEY.md 8d "config 2"): P=3 independent exact GPs sharing N=2048
training points in d=16 (12 states + 4 inputs), Q=8192 horizon queries per step, float64 (the reference's
arithmetic type: float64 data promotes gpytorch's kernel matrix).  One step = one pass of the hot path
over one batch: mean, variance and d-vector Jacobian for Q queries of each of the P GPs = P*Q predictions.
With --gpus N every rank holds the replicated fitted state and its own Q-query shard (weak scaling, no
data-path collective); value = all ranks' predictions / max-over-ranks device time.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Prints ONE JSON line (rank 0).  Synthetic data: seed 1337, X~U(-1,1), y=sin(Xw)+0.05eps, ell=0.5 sqrt(d),
sf2=1, noise=1e-2, Z~U(-1.2,1.2).
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_TRAIN, DIM, N_QUERY, N_GP = 2048, 16, 8192, 3
METRIC = "gp_posterior_preds_per_s"
UNIT = "preds/s"
WORKLOAD = f"config2: P={N_GP} GPs, N={N_TRAIN} train pts, d={DIM}, Q={N_QUERY} queries/step, mean+var+Jacobian"
DMMA_PEAK_TFLOPS = 37.1   # measured on this pool's B200 by tools/peaks (profiles/peaks_micro_r01.json); nominal 40


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the fp32 / MLL side measurements")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
# clocks during the timed region
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smax = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def l2_flusher(device):
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=device)   # > 126 MB L2

    def flush():
        buf.fill_(1)

    return flush


def make_problem(rank: int):
    """Synthetic recipe of SURVEY.md 8d (same numbers as oracle.synthetic_problem, restated here so the
    product arm does not import the oracle)."""
    import math

    g = torch.Generator().manual_seed(1337)
    X = torch.rand(N_TRAIN, DIM, generator=g, dtype=torch.float64) * 2.0 - 1.0
    Wm = torch.randn(DIM, N_GP, generator=g, dtype=torch.float64)
    eps = torch.randn(N_TRAIN, N_GP, generator=g, dtype=torch.float64)
    Y = torch.sin(X @ Wm) + 0.05 * eps
    Z = (torch.rand(N_QUERY, DIM, generator=g, dtype=torch.float64) * 2.0 - 1.0) * 1.2
    if rank > 0:   # every rank predicts its own shard of the global query batch
        g2 = torch.Generator().manual_seed(1337 + rank)
        Z = (torch.rand(N_QUERY, DIM, generator=g2, dtype=torch.float64) * 2.0 - 1.0) * 1.2
    ell = torch.full((DIM,), 0.5 * math.sqrt(DIM), dtype=torch.float64)
    return {"X": X, "Y": Y, "Z": Z, "ell": ell, "sf2": 1.0, "noise": 1e-2}


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle (torch float64 restatement of the gpytorch math) on the host cores.
# gpytorch itself is not installable here, so kind = "port".
# --------------------------------------------------------------------------------------------------
def cpu_predict_rate(problem, n_queries: int, steps: int = 1, warmup: int = 0):
    """Times oracle.predict (mean + variance + Jacobian) for the P GPs on `n_queries` of the workload's
    queries with all host threads.  Fit states (Cholesky, alpha) are prepared outside the timed region,
    like the GPU arm's resident state.  Returns (preds/s, seconds per step)."""
    from oracle import gp_oracle as o

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    X, Y, Z = problem["X"], problem["Y"], problem["Z"][:n_queries]
    states = [o.fit_state(X, Y[:, p], problem["ell"], problem["sf2"], problem["noise"]) for p in range(N_GP)]

    def step():
        for st in states:
            o.predict(Z, X, problem["ell"], problem["sf2"], st["L"], st["alpha"])

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return N_GP * n_queries / dt, dt, cores


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  gpytorch 1.12.0 cannot be installed
    here (no wheel, no network), so this is the oracle port on all host cores; each step is a bounded sample
    of the workload's queries sized so the whole run stays within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    problem = make_problem(0)
    # calibrate on 512 queries, then size the per-step sample for ~2.5 s per step
    rate0, _, cores = cpu_predict_rate(problem, 512, steps=1, warmup=1)
    total_steps = max(1, args.steps + args.warmup)
    budget_s = min(2.5, 150.0 / total_steps)
    sample = int(min(N_QUERY, max(256, rate0 * budget_s / N_GP)))
    rate, dt, cores = cpu_predict_rate(problem, sample, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{sample} of {N_QUERY} queries per step x {N_GP} GPs"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} of {N_QUERY} queries x {N_GP} GPs per step, fit state prepared untimed; "
                                   "gpytorch 1.12.0 not installable, oracle/gp_oracle.py (torch float64) timed instead"},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------------
def timed_steps(fn, steps, flush, device):
    """Sum of per-step CUDA-event durations (ms) on the current stream, L2 flushed before every step."""
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in evs:
        flush()
        e0.record()
        fn()
        e1.record()
    torch.cuda.synchronize(device)
    return sum(e0.elapsed_time(e1) for e0, e1 in evs)


def run_b200(args):
    import torch.distributed as dist

    from gpmpc_b200 import _lib, engine as E

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dev = _lib.require_cuda(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    problem = make_problem(rank)
    X, Y, Z = problem["X"].to(dev), problem["Y"].to(dev), problem["Z"].to(dev)
    t0 = time.perf_counter()
    gb = E.fit_batch(X, Y, problem["ell"], problem["sf2"], problem["noise"])   # replicated fitted state
    torch.cuda.synchronize(dev)
    fit_first_s = time.perf_counter() - t0       # includes one-time module loading and scratch allocation
    t0 = time.perf_counter()
    gb = E.fit_batch(X, Y, problem["ell"], problem["sf2"], problem["noise"])
    torch.cuda.synchronize(dev)
    fit_s = time.perf_counter() - t0
    flush = l2_flusher(dev)
    preds_per_step = N_GP * N_QUERY

    def step():
        return E.predict(gb, Z)

    # ---- value: inputs resident in HBM ----------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(dev.index)
    sampler.start()
    lib.gpk_reset_launch_count()
    torch.cuda.profiler.start()          # ncu --profile-from-start off captures exactly the timed steps
    total_ms = timed_steps(step, args.steps, flush, dev)
    torch.cuda.profiler.stop()
    launches = lib.gpk_launch_count()
    barrier()
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * preds_per_step * args.steps / (total_ms * 1e-3)

    # ---- e2e: host buffers through the public API, H2D + D2H inside the timed region ------------------
    from gpmpc_b200.gp import predict_host

    Zh = problem["Z"].pin_memory()
    out_h = None
    for _ in range(3):
        out_h = predict_host(gb, Zh, out_h)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out_h = predict_host(gb, Zh, out_h)      # returns after the results are on the host
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * preds_per_step * args.steps / float(t.item())
    clocks = sampler.stop()
    h2d = Zh.numel() * 8
    d2h = sum(o_.numel() * o_.element_size() for o_ in out_h)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "per_gpu_queries": N_QUERY, "parallelism": f"query-sharded x{world}, state replicated",
                   "l2": "flushed (256 MB write) before every timed step"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches, "clocks": clocks, "fit_state_s": fit_s, "fit_state_first_call_s": fit_first_s,
    }
    if rank == 0:
        line["roofline"] = measure_roofline(gb, Z, flush, dev, args.steps)
        if not args.no_extras:
            line["extras"] = measure_extras(gb, problem, Z, flush, dev)
        if world == 1 and not args.no_cpu_baseline:
            rate0, _, cores = cpu_predict_rate(problem, 512, steps=1, warmup=1)
            sample = int(min(N_QUERY, max(512, rate0 * 12.0 / N_GP)))   # ~12 s of CPU work
            rate, dt, cores = cpu_predict_rate(problem, sample, steps=1, warmup=0)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{sample} of {N_QUERY} queries x {N_GP} GPs, one pass, fit state prepared "
                                              "untimed; oracle/gp_oracle.py (torch float64, all host threads)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def measure_roofline(gb, Z, flush, dev, steps):
    """Dominant kernel = oz_var_kernel: the variance contraction V = W K_ZX^T on the tcgen05 int8 pipe (exact
    digit products in TMEM, fp64 recombination).  Timed alone with CUDA events on its launch stream, digit
    image of E prepared beforehand, L2 flushed before every launch.

    achieved = int8 tensor-pipe operations the algorithm needs per launch / launch time, where the algorithmic
    count is SURVEY.md 8d's N^2 flops per prediction (triangular contraction) x the S(S+1)/2 digit products an
    S-digit error-free fp64 product consists of x P*Q predictions.  peak = 2 x the measured dense bf16 burst of
    MEASURED_PEAKS.json (the int8 tcgen05 rate is exactly twice the bf16 rate: 8192 vs 4096 MAC/clk/SM,
    tools/peaks/umma_bench); the file holds no int8 figure of its own."""
    from gpmpc_b200 import _lib, engine as E
    from gpmpc_b200._lib import call, ptr

    lib = _lib.load()
    B, n, d, npad, S = len(gb.fits), gb.n, gb.d, gb.npad, gb.digits
    q = Z.shape[0]
    qp = E.round_up(q, lib.gpk_oz_query_tile())
    st = _lib.stream_ptr(dev)
    eimg = torch.empty(B, S * qp * npad, dtype=torch.int8, device=dev)
    pieces = lib.gpk_oz_pieces(qp // lib.gpk_oz_query_tile() * B, npad)
    part = torch.empty(B, 16, qp, dtype=torch.float64, device=dev)
    call("gpk_oz_cross", 1, ptr(Z), Z.stride(0), q, 0, ptr(gb.X), gb.X.stride(1), n, gb.X.stride(0), d, ptr(gb.hyp),
         ptr(eimg), eimg.stride(0), npad, qp, S, B, st)

    def gemm():
        call("gpk_oz_var_partials", ptr(eimg), eimg.stride(0), ptr(gb.wimg), gb.wimg.stride(0), ptr(gb.wscale),
             gb.wscale.stride(0), ptr(gb.hyp), npad, qp, S, ptr(part), qp, 16 * qp, pieces, B, None, st)

    for _ in range(3):
        gemm()
    ms = timed_steps(gemm, steps, flush, dev) / steps
    fp64_flops = float(n) * n * B * q                       # N^2 flops per prediction
    int8_ops = fp64_flops * S * (S + 1) / 2
    achieved = int8_ops / (ms * 1e-3) / 1e12
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    bf16 = peaks.get("bf16_tflops")
    peak = 2.0 * bf16 if bf16 else 2.0 * 1590.0
    return {"kernel": f"oz_var_kernel<{S}> (variance contraction W K_ZX^T, {S}-digit int8 slices on tcgen05, int32 TMEM "
                      "accumulators, fp64 recombination)",
            "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            "unit_note": "int8 tensor operations (multiply + add = 2 ops), i.e. TOP/s on the scale of the bf16 TFLOP/s peak x 2",
            "traffic": ncu_dram_bytes("ncu_oz_var_kernel_S5_r01_v6.csv") if S == 5 else None,
            "traffic_unit": "bytes per launch (dram read + write, ncu --set full, profiles/ncu_oz_var_kernel_S5_r01_v6.csv)",
            "algorithmic_bytes_per_launch": float(B) * S * npad * (qp + npad),
            "ms_per_launch": ms, "algorithmic_int8_ops_per_launch": int8_ops,
            "fp64_equivalent_tflops": fp64_flops / (ms * 1e-3) / 1e12, "fp64_dmma_peak_tflops": DMMA_PEAK_TFLOPS,
            "peak_source": ("2 x MEASURED_PEAKS.json bf16_tflops (burst, of measured)" if bf16 else
                            "2 x 1590 TFLOP/s bf16 (of fallback)") + "; nominal int8 dense 4500 TOP/s"}


def ncu_dram_bytes(name):
    """dram__bytes_read.sum + dram__bytes_write.sum of the committed ncu capture profiles/<name> (same workload), or None."""
    try:
        tot = 0.0
        with open(os.path.join(ROOT, "profiles", name)) as f:
            for line in f:
                k = line.split(",")
                if k[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(k[2]) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[k[1]]
        return tot or None
    except (OSError, KeyError, ValueError, IndexError):
        return None


def measure_extras(gb, problem, Z, flush, dev):
    """Side measurements (not the headline): float32 prediction path, mean+Jacobian kernel alone, MLL+grad."""
    from gpmpc_b200 import engine as E

    out = {}
    Z32 = Z.float()
    for _ in range(3):
        E.predict(gb, Z32)
    ms = timed_steps(lambda: E.predict(gb, Z32), 10, flush, dev) / 10
    out["f32_path_preds_per_s"] = N_GP * N_QUERY / (ms * 1e-3)
    # the same step with the variance contraction on the FP64 DMMA pipe (round-1 first path, kept as var_path="f64")
    for _ in range(3):
        E.predict(gb, Z, var_path="f64")
    ms = timed_steps(lambda: E.predict(gb, Z, var_path="f64"), 10, flush, dev) / 10
    out["dmma_var_path_preds_per_s"] = N_GP * N_QUERY / (ms * 1e-3)
    v_ref = E.predict(gb, Z, var_path="f64")[1]
    out["var_rel_err_vs_dmma_path_digits5"] = float((E.predict(gb, Z)[1] - v_ref).abs().max() / v_ref.abs().max())
    # the same step with 4 (30-bit) or 6 (46-bit) digits: 10 / 21 digit products instead of 15
    for S in (4, 6):
        gbs = E.fit_batch(problem["X"].to(dev), problem["Y"].to(dev), problem["ell"], problem["sf2"], problem["noise"],
                          digits=S)
        for _ in range(3):
            E.predict(gbs, Z)
        ms = timed_steps(lambda: E.predict(gbs, Z), 10, flush, dev) / 10
        out[f"digits{S}_preds_per_s"] = N_GP * N_QUERY / (ms * 1e-3)
        out[f"var_rel_err_vs_dmma_path_digits{S}"] = float((E.predict(gbs, Z)[1] - v_ref).abs().max() / v_ref.abs().max())
        del gbs
    for name, zz in (("f64", Z), ("f32", Z32)):
        for _ in range(3):
            E.predict(gb, zz, want_var=False)
        ms = timed_steps(lambda: E.predict(gb, zz, want_var=False), 10, flush, dev) / 10
        out[f"mean_jac_only_{name}_preds_per_s"] = N_GP * N_QUERY / (ms * 1e-3)
        out[f"mean_jac_only_{name}_pairs_per_s"] = N_GP * N_QUERY * N_TRAIN / (ms * 1e-3)
    # MLL value + full gradient for one GP (fit_gp's inner iteration), N = 2048
    X, y = problem["X"].to(dev), problem["Y"][:, 0].to(dev)

    def mll_iter():
        fit, ws = E.factorize(X, y, problem["ell"], problem["sf2"], problem["noise"], want_kinv=True, want_image=False)
        E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))

    for _ in range(2):
        mll_iter()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(5):
        mll_iter()
    torch.cuda.synchronize(dev)
    out["mll_grad_iters_per_s_n2048"] = 5 / (time.perf_counter() - t0)
    # the same iteration at N = 16384 (half of config 4's N, to keep the default run short): the large GEMMs of the
    # Cholesky / triangular inverse / K^-1 run on the tcgen05 int8 pipe (csrc/ozgemm.cu, E.FIT_GEMM_DIGITS digits),
    # compared with the FP64 DMMA pipe alone
    gb16 = torch.Generator().manual_seed(1337)
    Xb = (torch.rand(16384, DIM, generator=gb16, dtype=torch.float64) * 2.0 - 1.0).to(dev)
    yb = (torch.sin(Xb @ torch.randn(DIM, generator=gb16, dtype=torch.float64).to(dev))
          + 0.05 * torch.randn(16384, generator=gb16, dtype=torch.float64).to(dev))
    big = problem
    saved = E.FIT_GEMM_DIGITS
    losses = {}
    for tag, digits in (("dmma", 0), ("int8", saved or 6)):
        E.FIT_GEMM_DIGITS = digits
        E.release_scratch()

        def big_iter():
            fit, ws = E.factorize(Xb, yb, big["ell"], big["sf2"], big["noise"], want_kinv=True, want_image=False)
            return E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))[0]

        big_iter()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(3):
            losses[tag] = big_iter()
        torch.cuda.synchronize(dev)
        dt = (time.perf_counter() - t0) / 3
        out[f"mll_grad_iters_per_s_n16384_{tag}"] = 1.0 / dt
        out[f"mll_grad_n16384_{tag}_tflops_of_n3"] = 16384.0 ** 3 / dt * 1e-12
    out["mll_grad_n16384_int8_digits"] = int(saved or 6)
    out["mll_n16384_rel_dev_int8_vs_dmma"] = abs(losses["int8"] - losses["dmma"]) / abs(losses["dmma"])
    E.FIT_GEMM_DIGITS = saved
    E.release_scratch()
    del Xb, yb
    out.update(measure_mpc_rows(gb, problem, flush, dev))
    return out


def measure_mpc_rows(gb, problem, flush, dev):
    """The MPC-side rows of SURVEY.md 8(f) on the bench GPs' training set size: batched RK4 stage dynamics with
    sensitivities (4 x [3-GP mean+Jacobian + stage kernel]) and the chance-constraint recursion, for 2048 rollouts of
    32 stages (65536 stage evaluations)."""
    from gpmpc_b200 import _lib
    from gpmpc_b200._lib import call, ptr
    from gpmpc_b200.mpc import StageDynamics

    out = {}
    R, T, nx, nu = 2048, 32, 12, 4
    g = torch.Generator().manual_seed(7)
    X, Y = problem["X"], problem["Y"]
    gps = [(X[:, :dd].contiguous(), Y[:, k].contiguous(), torch.full((dd,), 0.6, dtype=torch.float64), 1.0, 1e-2)
           for k, dd in enumerate((1, 3, 3))]
    sd = StageDynamics(gps, [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653, 9.81], 0.02)
    x = ((torch.rand(R * T, nx, generator=g, dtype=torch.float64) - 0.5) * 0.8).to(dev)
    u = ((torch.rand(R * T, nu, generator=g, dtype=torch.float64) - 0.3) * 0.6).to(dev)
    for _ in range(3):
        sd.disc_dyn(x, u, work_bytes=8 << 30)
    ms = timed_steps(lambda: sd.disc_dyn(x, u, work_bytes=8 << 30), 5, flush, dev) / 5
    out["stage_rk4_evals_per_s_n2048"] = R * T / (ms * 1e-3)
    out["stage_rk4_kernel_pairs_per_s"] = 9 * R * T * X.shape[0] / (ms * 1e-3)   # 3 + 2 + 2 + 2 GP evaluations
    # tightening recursion on its own, 65536 rollouts x 25 stages (variances resident)
    R2, T2 = 65536, 25
    A = (torch.eye(nx, dtype=torch.float64) + 0.02 * torch.randn(nx, nx, generator=g, dtype=torch.float64)).to(dev)
    B = (0.1 * torch.randn(nx, nu, generator=g, dtype=torch.float64)).to(dev)
    K = (-0.3 * torch.randn(nu, nx, generator=g, dtype=torch.float64)).to(dev)
    Bd = torch.eye(nx, dtype=torch.float64)[:, [1, 3, 5, 9, 10]].contiguous().to(dev)
    z = (torch.rand(R2 * T2, nx + nu, generator=g, dtype=torch.float64) - 0.5).to(dev)
    var = (torch.rand(3, R2 * T2, generator=g, dtype=torch.float64) * 1e-2).to(dev)
    sc = torch.empty(R2, 2 * nx, T2 + 1, dtype=torch.float64, device=dev)
    ic = torch.empty(R2, 2 * nu, T2, dtype=torch.float64, device=dev)
    st = _lib.stream_ptr(dev)

    def tighten():
        call("gpk_tighten_constraints", ptr(var), var.stride(0), 0, ptr(z), z.stride(0), 6, 7, ptr(gb.hyp), ptr(A),
             ptr(B), ptr(K), ptr(Bd), nx, nu, T2, R2, 0.02, 2.0, ptr(sc), ptr(ic), st)

    for _ in range(3):
        tighten()
    ms = timed_steps(tighten, 5, flush, dev) / 5
    out["tighten_rollouts_per_s_T25"] = R2 / (ms * 1e-3)
    out["tighten_fp64_tflops"] = R2 * T2 * (4 * nx**3 + 2 * nu * nx * nx) / (ms * 1e-3) / 1e12
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
