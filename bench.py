"""bench.py -- GP posterior predictions/s (mean + variance + Jacobian) of the gp-mpc GP hot path on B200.

Headline workload (BASELINE.json configs[2], SURVEY.md 8d "config 3"): P=3 independent exact GPs sharing N=16384
training points in d=16 (12 states + 4 inputs), ONE fixed global batch of Q rollout queries per step, float64 (the
reference's arithmetic type).  One step = mean, variance and d-vector Jacobian for all Q queries of each of the P
GPs = P*Q predictions, issued through `gpmpc_b200.sharding.predict_sharded`: the global batch is cut into one
contiguous block per rank (strong scaling), the fitted state is computed ONCE on rank 0 and replicated with
`sharding.replicate_batch` (timed, `state_broadcast_s`), and there is no collective in the prediction itself.
Before the timed region every rank checks that the sharded + gathered result equals its own unsharded prediction.
Q defaults to 1,048,576 = a quarter of config 3's 4,194,304 queries (--queries to change): the full batch costs
~20 s per step on one GPU, and the driver's 25 timed + warm-up steps plus the equally long end-to-end leg would not
fit its per-N time budget; per-query cost does not depend on Q at this size (8192 query tiles per GP).

Other configs in the same JSON line (`extras`), so that they land in the driver-run records:
  config2   (N=1 only) BASELINE.json configs[1]: P=3, N=2048, Q=8192 -- round 1's headline, for comparability
  config4a  P = world GPs of N=32768, one per rank: MLL + gradient iterations/s        (output-GP sharding)
  config4b  one GP of N=32768, K^-1 rows + gradient trace split over the ranks, NCCL all-reduce timed  (N>1)
  config5   65,536 rollouts x horizon 50 over the ranks: 50 sequential stage batches (stage dynamics feed the next
            stage's GP queries) + chance-constraint tightening
  config1   (N=1 only) the reference's own shapes: 3 GPs, N=45, Adam loop of fit_gp, beside the CPU oracle loop
and a first-class `mll` block (MLL + gradient iterations/s, half of BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload config3|config2]

Prints ONE JSON line (rank 0).  Synthetic data: seed 1337, X~U(-1,1), y=sin(Xw)+0.05eps, ell=0.5 sqrt(d),
sf2=1, noise=1e-2, Z~U(-1.2,1.2).
"""

from __future__ import annotations

import argparse
import json
import math
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

DIM, N_GP = 16, 3
WORKLOADS = {
    # name: (N train, default global queries per step, what the name means)
    "config3": (16384, 1 << 20, "BASELINE.json configs[2]"),
    "config2": (2048, 8192, "BASELINE.json configs[1]"),
}
METRIC = "gp_posterior_preds_per_s"
UNIT = "preds/s"
NOMINAL_INT8_TOPS = 4500.0   # B200 dense int8, NVIDIA's figure
FALLBACK_BF16_TFLOPS = 1590.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS))
    ap.add_argument("--queries", type=int, default=0, help="global queries per step (default: the workload's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--equal-shards", action="store_true", help="N > 1: equal query blocks instead of speed-weighted ones")
    ap.add_argument("--no-extras", action="store_true", help="skip the other configs / side measurements")
    return ap.parse_args()


def workload_string(name, n, q):
    return (f"{name} ({WORKLOADS[name][2]}): P={N_GP} GPs, N={n} train pts, d={DIM}, Q={q} queries per step (global, "
            f"split over the ranks), mean+var+Jacobian")


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
        return self

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
        sm, smax, power, reasons = [], None, [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smax = float(r[1])
                power.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm), "sm_mhz_min": min(sm) if sm else None, "power_w_max": max(power) if power else None}


def l2_flusher(device):
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=device)   # > 126 MB L2

    def flush():
        buf.fill_(1)

    return flush


def make_problem(n_train: int, n_query: int):
    """Synthetic recipe of SURVEY.md 8d (same numbers as oracle.synthetic_problem, restated here so the product arm
    does not import the oracle).  Identical on every rank: Z is the GLOBAL query batch."""
    g = torch.Generator().manual_seed(1337)
    X = torch.rand(n_train, DIM, generator=g, dtype=torch.float64) * 2.0 - 1.0
    Wm = torch.randn(DIM, N_GP, generator=g, dtype=torch.float64)
    eps = torch.randn(n_train, N_GP, generator=g, dtype=torch.float64)
    Y = torch.sin(X @ Wm) + 0.05 * eps
    Z = (torch.rand(n_query, DIM, generator=g, dtype=torch.float64) * 2.0 - 1.0) * 1.2
    ell = torch.full((DIM,), 0.5 * math.sqrt(DIM), dtype=torch.float64)
    return {"X": X, "Y": Y, "Z": Z, "ell": ell, "sf2": 1.0, "noise": 1e-2}


def load_peaks():
    peaks = {}
    for fn in ("MEASURED_PEAKS.json", os.path.join("profiles", "peaks_micro_r02.json")):
        try:
            with open(os.path.join(ROOT, fn)) as f:
                peaks.update(json.load(f))
        except (OSError, ValueError):
            pass
    return peaks


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle (torch float64 restatement of the gpytorch math) on the host cores.
# --------------------------------------------------------------------------------------------------
def try_gpytorch():
    """the reference's own GP stack if the box has it (it does not in this image)."""
    try:
        import gpytorch  # noqa: F401
        return "importable (version %s) but the reference arm still times the oracle port" % gpytorch.__version__
    except Exception as e:  # noqa: BLE001
        return f"not importable here ({type(e).__name__}): oracle port timed instead"


class CpuArm:
    """oracle.predict (mean + variance + Jacobian) for the P GPs with all host threads; fit states (Cholesky, alpha)
    are prepared outside the timed region like the GPU arm's resident state.  Two formulations of the kernel matrix
    are timed -- the direct-difference loop the oracle is written in (gpmpc/gp.py:12-14) and the GEMM form gpytorch's
    RBFKernel really uses (oracle.se_kernel_gpytorch) -- and the faster is reported."""

    def __init__(self, problem):
        from oracle import gp_oracle as o

        self.o, self.p = o, problem
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        # untimed preparation: at large N the kernel matrix is assembled in the GEMM form (16 passes over a 2 GB matrix
        # otherwise), and only L and alpha are kept
        saved = o.se_kernel
        if problem["X"].shape[0] >= 8192:
            o.se_kernel = o.se_kernel_gpytorch
        try:
            self.states = []
            for k in range(N_GP):
                st = o.fit_state(problem["X"], problem["Y"][:, k], problem["ell"], problem["sf2"], problem["noise"])
                self.states.append({"L": st["L"], "alpha": st["alpha"]})
                del st
        finally:
            o.se_kernel = saved
        self.kernel = "direct"

    def _step(self, Z, kernel):
        o, p = self.o, self.p
        saved = o.se_kernel
        if kernel == "gemm":
            o.se_kernel = o.se_kernel_gpytorch
        try:
            for st in self.states:
                o.predict(Z, p["X"], p["ell"], p["sf2"], st["L"], st["alpha"])
        finally:
            o.se_kernel = saved

    def calibrate(self, q=256):
        """Pick the faster kernel formulation; returns preds/s of it on q queries."""
        Z = self.p["Z"][:q]
        rates = {}
        for kern in ("direct", "gemm"):
            self._step(Z[:32], kern)
            t0 = time.perf_counter()
            self._step(Z, kern)
            rates[kern] = N_GP * q / (time.perf_counter() - t0)
        self.kernel = max(rates, key=rates.get)
        self.rates = rates
        return rates[self.kernel]

    def rate(self, n_queries, steps=1, warmup=0):
        Z = self.p["Z"][:n_queries]
        for _ in range(warmup):
            self._step(Z, self.kernel)
        t0 = time.perf_counter()
        for _ in range(steps):
            self._step(Z, self.kernel)
        dt = (time.perf_counter() - t0) / steps
        return N_GP * n_queries / dt, dt

    def describe(self, sample, q_total):
        return (f"{sample} of {q_total} queries x {N_GP} GPs per step, fit state prepared untimed; oracle/gp_oracle.py "
                f"(torch float64, {self.cores} host threads), kernel matrix by {self.kernel} form "
                f"(direct {self.rates['direct']:.0f} / gemm {self.rates['gemm']:.0f} preds/s in calibration); gpytorch "
                f"1.12.0 {try_gpytorch()}")


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  gpytorch 1.12.0 cannot be installed here (no
    wheel, no network), so this is the oracle port on all host cores; each step is a bounded sample of the workload's
    queries sized so the whole run stays within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_train, q_default, _ = WORKLOADS[args.workload]
    q_total = args.queries or q_default
    problem = make_problem(n_train, min(q_total, 16384))
    arm = CpuArm(problem)
    rate0 = arm.calibrate(min(256, problem["Z"].shape[0]))
    total_steps = max(1, args.steps + args.warmup)
    budget_s = min(2.5, 150.0 / total_steps)
    sample = int(min(problem["Z"].shape[0], max(64, rate0 * budget_s / N_GP)))
    rate, dt = arm.rate(sample, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, n_train, q_total),
                   "sample": f"{sample} of {q_total} queries per step x {N_GP} GPs"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": arm.cores, "kind": "port",
                         "sample": arm.describe(sample, q_total)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------------
class Ctx:
    """Per-process distributed context."""

    def __init__(self):
        import torch.distributed as dist

        from gpmpc_b200 import _lib

        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dev = _lib.require_cuda(f"cuda:{self.local_rank}")
        torch.cuda.set_device(self.dev)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.lib = _lib.load()
        self.flush = l2_flusher(self.dev)

    def barrier(self):
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return float(v)
        t = torch.tensor([v], dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def gather_objects(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out


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


def rel_dev(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-300))


def elementwise_var_err(v, v_ref, sf2):
    """max |dv| / max(v_ref, 1e-3 sf2): element-wise relative error of the variance with a floor at 1e-3 of the prior
    variance (the variance cancels near training data, so a normwise figure hides the small entries)."""
    return float(((v - v_ref).abs() / v_ref.clamp_min(1e-3 * sf2)).max())


def run_b200(args):
    from gpmpc_b200 import engine as E, sharding as S
    from gpmpc_b200.gp import predict_host

    ctx = Ctx()
    dev, world, rank, dist = ctx.dev, ctx.world, ctx.rank, ctx.dist
    n_train, q_default, _ = WORKLOADS[args.workload]
    q_total = args.queries or q_default
    problem = make_problem(n_train, q_total)
    warmup = max(args.warmup, 3)

    # ---- fitted state: once on rank 0, replicated to every rank ------------------------------------
    fit_first_s = fit_s = None
    gb = None
    if rank == 0:
        X, Y = problem["X"].to(dev), problem["Y"].to(dev)
        t0 = time.perf_counter()
        gb = E.fit_batch(X, Y, problem["ell"], problem["sf2"], problem["noise"])
        torch.cuda.synchronize(dev)
        fit_first_s = time.perf_counter() - t0       # includes one-time module loading and scratch allocation
        del gb
        t0 = time.perf_counter()
        gb = E.fit_batch(X, Y, problem["ell"], problem["sf2"], problem["noise"])
        torch.cuda.synchronize(dev)
        fit_s = time.perf_counter() - t0
        E.release_scratch()                           # the fit scratch (L, U, GEMM images) is not prediction state
    ctx.barrier()
    t0 = time.perf_counter()
    gb, bcast_bytes = S.replicate_batch(gb, src=0, device=dev)
    ctx.barrier()
    bcast_s = time.perf_counter() - t0

    Zd = problem["Z"].to(dev)                         # the global batch, resident in HBM on every rank
    lo, hi = S.shard_bounds(q_total, world, rank)
    preds_per_step = N_GP * q_total

    # ---- self-check: sharded (+ gathered) == unsharded, on every rank ---------------------------------------
    qc = min(q_total, 8192)
    full = E.predict(gb, Zd[:qc])
    _, _, gathered = S.predict_sharded(lambda z: E.predict(gb, z), Zd[:qc], gather=True)
    check = max(rel_dev(a, b) for a, b in zip(gathered, full))
    check = ctx.max_over_ranks(check)
    assert check < 1e-11, f"sharded prediction deviates from the unsharded one: {check:.3e}"
    del full, gathered

    bounds = [S.shard_bounds(q_total, world, r) for r in range(world)]

    def step():
        return S.predict_sharded(lambda z: E.predict(gb, z), Zd, bounds=bounds)

    # ---- value: inputs resident in HBM ----------------------------------------------------------
    for _ in range(warmup):
        step()
    # Speed-weighted blocks: the GPUs of one box hold different clocks under the power cap (the step ends with the
    # slowest), so every rank times TWO more warm-up steps on its equal block (rank-local CUDA events), the speeds are
    # all-gathered once -- outside the step -- and the blocks are re-cut in proportion (sharding.weighted_bounds,
    # 128-query granularity).  The timed steps still contain no collective.
    shard_weights = None
    if world > 1 and not args.equal_shards:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.barrier()
        e0.record()
        step()
        step()
        e1.record()
        torch.cuda.synchronize(dev)
        shard_weights = S.measured_weights(2 * (hi - lo), e0.elapsed_time(e1) * 1e-3)
        bounds = S.weighted_bounds(q_total, shard_weights, 128)
        lo, hi = bounds[rank]
        step()                                         # one warm-up step on the new blocks
    ctx.barrier()
    sampler = ClockSampler(dev.index).start()
    ctx.lib.gpk_reset_launch_count()
    torch.cuda.profiler.start()          # ncu --profile-from-start off captures exactly the timed steps
    total_ms = timed_steps(step, args.steps, ctx.flush, dev)
    torch.cuda.profiler.stop()
    launches = ctx.lib.gpk_launch_count()
    clocks = sampler.stop()
    ctx.barrier()
    total_ms = ctx.max_over_ranks(total_ms)
    value = preds_per_step * args.steps / (total_ms * 1e-3)

    # ---- e2e: host buffers through the public API, H2D + D2H inside the timed region ------------------
    Zh = problem["Z"].pin_memory()
    out_h = [None]

    def e2e_step():
        def fn(z_host):
            out_h[0] = predict_host(gb, z_host, out_h[0])      # returns after the results are on the host
            return out_h[0]
        return S.predict_sharded(fn, Zh, bounds=bounds)

    for _ in range(2):
        e2e_step()
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    e2e_s = ctx.max_over_ranks(time.perf_counter() - t0)
    e2e_value = preds_per_step * args.steps / e2e_s
    h2d = (hi - lo) * DIM * 8
    d2h = sum(o_.numel() * o_.element_size() for o_ in out_h[0])
    h2d_all = sum(ctx.gather_objects(h2d))
    d2h_all = sum(ctx.gather_objects(d2h))
    out_h[0] = None
    del Zh

    all_clocks = ctx.gather_objects(clocks)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, n_train, q_total), "global_queries": q_total,
                   "queries_per_rank": [b[1] - b[0] for b in bounds],
                   "shard_speeds_preds_per_s": shard_weights,
                   "sharding": ("equal contiguous blocks" if shard_weights is None else
                                "contiguous blocks proportional to each rank's measured speed on two warm-up steps "
                                "(sharding.measured_weights + weighted_bounds, 128-query granularity)"),
                   "parallelism": f"sharding.predict_sharded: contiguous query blocks over {world} rank(s), fitted state "
                                  "fitted on rank 0 and replicated (sharding.replicate_batch), no collective in the step",
                   "l2": "flushed (256 MB write) before every timed step; the step's operands (W image "
                         f"{gb.wimg.numel() / 1e9:.2f} GB + E image chunks) exceed the 126 MB L2 anyway"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all},
        "gpu_launches": launches, "clocks": clocks,
        "clocks_all_ranks": {"sm_mhz": [c.get("sm_mhz") for c in all_clocks], "sm_mhz_min": [c.get("sm_mhz_min") for c in all_clocks],
                             "reasons": sorted({r for c in all_clocks for r in c.get("reasons", [])})},
        "fit_state_s": fit_s, "fit_state_first_call_s": fit_first_s,
        "state_broadcast_s": bcast_s if world > 1 else 0.0, "state_broadcast_bytes": bcast_bytes,
        "sharded_vs_unsharded_max_rel_dev": check,
        "variance_digits": {"digits": gb.digits, "digits_f32_queries": gb.digits_f32 or gb.digits,
                            "policy": "engine.select_digits: smallest digit count whose probed |var_i8 - var_f64dmma| x "
                                      "2 <= 1e-6 x noise (element-wise 1e-6 on the predictive variance var_f + noise)",
                            "probed_abs_err_per_gp": [f.var_err for f in gb.fits], "noise": problem["noise"]},
        "comm_nranks_seen": dist.get_world_size() if world > 1 else 1,
    }
    roof = measure_roofline(gb, Zd[lo:hi], ctx, max(3, min(args.steps, 10)))
    # step-level fraction: the whole step's algorithmic int8 operations against the same peak
    S_ = gb.digits
    step_ops = float(gb.n) ** 2 * S_ * (S_ + 1) / 2 * preds_per_step / world       # per GPU
    roof["step_frac"] = step_ops / (total_ms / args.steps * 1e-3) / 1e12 / roof["peak"]
    roof["step_frac_note"] = ("whole prediction step (front end + variance contraction + finalize), algorithmic int8 "
                              "operations of the step / ms_per_step / peak, per GPU")
    line["roofline"] = roof

    extras = {}
    if not args.no_extras:
        del Zd
        gb_keep = gb if args.workload == "config2" else None
        del gb
        E.release_scratch()
        torch.cuda.empty_cache()
        extras.update(measure_fit_configs(ctx))
        extras.update(measure_config5(ctx))
        if world == 1:
            extras.update(measure_single_gpu_extras(ctx, gb_keep, args.workload))
    if rank == 0:
        if extras:
            line["mll"] = extras.pop("mll", None)
            line["extras"] = extras
        if world == 1 and not args.no_cpu_baseline:
            arm = CpuArm(make_problem(n_train, min(q_total, 16384)))
            rate0 = arm.calibrate(256)
            sample = int(min(arm.p["Z"].shape[0], max(256, rate0 * 12.0 / N_GP)))   # ~12 s of CPU work
            rate, dt = arm.rate(sample, steps=1, warmup=0)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": arm.cores, "kind": "port",
                                    "sample": arm.describe(sample, q_total)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def measure_roofline(gb, Z, ctx, steps):
    """Dominant kernel = oz_var_kernel: the variance contraction V = W K_ZX^T on the tcgen05 int8 pipe (exact digit
    products in TMEM, fp64 recombination).  Timed alone with CUDA events on its launch stream, digit image of E
    prepared beforehand, L2 flushed before every launch, on (a chunk of) this rank's query block.

    achieved = int8 tensor-pipe operations the algorithm needs per launch / launch time: SURVEY.md 8d's N^2 flops per
    prediction (triangular contraction) x the S(S+1)/2 digit products of an S-digit error-free fp64 product x the
    predictions of the launch.  peak = 2 x the measured dense bf16 burst of MEASURED_PEAKS.json (tcgen05 int8 issues at
    twice the bf16 rate: 8192 vs 4096 MAC/clk/SM); the int8 rate measured directly by tools/peaks/umma_bench and the
    nominal 4500 TOP/s are reported beside it."""
    from gpmpc_b200 import _lib, engine as E
    from gpmpc_b200._lib import call, ptr

    lib, dev = ctx.lib, ctx.dev
    B, n, d, npad, S = len(gb.fits), gb.n, gb.d, gb.npad, gb.digits
    qt = lib.gpk_oz_query_tile()
    # one work chunk of the prediction step (the step walks the query block in chunks that fit its scratch)
    per_q = S * npad * B
    q = int(min(Z.shape[0], max(qt, (E.VAR_WORK_BYTES // per_q) // qt * qt)))
    Z = Z[:q].contiguous()
    qp = E.round_up(q, qt)
    st = _lib.stream_ptr(dev)
    eimg = torch.empty(B, S * qp * npad, dtype=torch.int8, device=dev)
    pieces = lib.gpk_oz_pieces(qp // qt * B, npad)
    part = torch.empty(B, 16, qp, dtype=torch.float64, device=dev)
    call("gpk_oz_cross", 1, ptr(Z), Z.stride(0), q, 0, ptr(gb.X), gb.X.stride(1), n, gb.X.stride(0), d, ptr(gb.hyp),
         ptr(eimg), eimg.stride(0), npad, qp, S, B, st)

    def gemm():
        call("gpk_oz_var_partials", ptr(eimg), eimg.stride(0), ptr(gb.wimg), gb.wimg.stride(0), ptr(gb.wscale),
             gb.wscale.stride(0), ptr(gb.hyp), npad, qp, S, ptr(part), qp, 16 * qp, pieces, B, None, st)

    for _ in range(2):
        gemm()
    ms = timed_steps(gemm, steps, ctx.flush, dev) / steps
    fp64_flops = float(n) * n * B * q                       # N^2 flops per prediction
    int8_ops = fp64_flops * S * (S + 1) / 2
    achieved = int8_ops / (ms * 1e-3) / 1e12
    peaks = load_peaks()
    # MEASURED_PEAKS.json holds a burst figure (best of 10 short GEMMs) and a sustained one (4 s back to back, clocks
    # settled under the power cap).  A launch that is itself long (tens of ms, repeated for seconds: config 3) runs in
    # the sustained regime; a sub-millisecond launch timed alone between L2 flushes (config 2) in the burst regime.
    sustained = ms >= 10.0 and peaks.get("bf16_tflops_sustained")
    bf16 = peaks.get("bf16_tflops_sustained") if sustained else peaks.get("bf16_tflops")
    peak = 2.0 * (bf16 or FALLBACK_BF16_TFLOPS)
    dmma_peak = peaks.get("dmma_tflops_bps8", 37.1)
    prof = f"ncu_oz_var_kernel_S{S}_r02_n{n}.csv"
    if not os.path.exists(os.path.join(ROOT, "profiles", prof)) and n == 2048 and S == 5:
        prof = "ncu_oz_var_kernel_S5_r01_v6.csv"        # same kernel, same shape, captured in round 1
    out = {"kernel": f"oz_var_kernel<{S}> (variance contraction W K_ZX^T, {S}-digit int8 slices on tcgen05, int32 TMEM "
                     "accumulators, fp64 recombination)",
           "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
           "unit_note": "int8 tensor operations (multiply + add = 2 ops), i.e. TOP/s on the scale of the bf16 TFLOP/s peak x 2",
           "frac_of_nominal": achieved / NOMINAL_INT8_TOPS, "nominal_peak": NOMINAL_INT8_TOPS,
           "traffic": ncu_dram_bytes(prof),
           "traffic_unit": f"bytes per launch (dram read + write, ncu --set full, profiles/{prof})",
           "algorithmic_bytes_per_launch": float(B) * S * npad * (qp + npad / 2),
           "queries_per_launch": q, "ms_per_launch": ms, "algorithmic_int8_ops_per_launch": int8_ops,
           "fp64_equivalent_tflops": fp64_flops / (ms * 1e-3) / 1e12, "fp64_dmma_peak_tflops": dmma_peak,
           "peak_source": (f"2 x MEASURED_PEAKS.json {'bf16_tflops_sustained (sustained' if sustained else 'bf16_tflops (burst'}"
                           f", of measured): {ms:.2f} ms per launch, {steps + 2} launches back to back" if bf16 else
                           f"2 x {FALLBACK_BF16_TFLOPS} TFLOP/s bf16 (of fallback)")}
    # what bounds the kernel besides the tensor pipe (profiles/pipeline_r02.md): operand fetches from shared memory
    g = sum(range(1, S + 1))                              # digit groups multiplied per k-step: columns = 64 g
    a_reads = S + (1 if 64 * S > 256 else 0)              # the widest instruction is split in two: its A tile is read twice
    smem_bytes_per_kstep = a_reads * 4096 + 64 * g * 32
    out["shared_memory_bound"] = {
        "bytes_per_clk_at_full_tensor_rate": smem_bytes_per_kstep / (64 * g / 2.0), "sm_bytes_per_clk": 128,
        "note": "per 32-byte k-step a CTA's tcgen05.mma read the E digit tiles (4 KB each) and the concatenated W digit columns "
                "(32 B each) from shared memory in the N/2 clocks the tensor pipe needs: the kernel tops out near "
                "128 / that figure of tensor-pipe-active (ncu: 89.7 % at config 3), and a co-resident kernel that needs "
                "shared memory starves (measured)"}
    if peaks.get("bf16_tflops") and peaks.get("bf16_tflops_sustained"):
        out["frac_of_burst_peak"] = achieved / (2.0 * peaks["bf16_tflops"])
        out["frac_of_sustained_peak"] = achieved / (2.0 * peaks["bf16_tflops_sustained"])
    if peaks.get("int8_tops_burst"):
        out["frac_of_measured_int8"] = achieved / peaks["int8_tops_burst"]
        out["measured_int8_tops"] = {"burst": peaks["int8_tops_burst"], "sustained": peaks.get("int8_tops_sustained"),
                                     "source": "profiles/peaks_micro_r02.json (tools/peaks/umma_bench --json)"}
    return out


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


# --------------------------------------------------------------------------------------------------
# config 4 (hyper-parameter fit at N = 32768) and the MLL block -- every rank participates
# --------------------------------------------------------------------------------------------------
def fit_data(n, seed, dev):
    g = torch.Generator().manual_seed(seed)
    X = (torch.rand(n, DIM, generator=g, dtype=torch.float64) * 2.0 - 1.0).to(dev)
    y = (torch.sin(X @ torch.randn(DIM, generator=g, dtype=torch.float64).to(dev))
         + 0.05 * torch.randn(n, generator=g, dtype=torch.float64).to(dev))
    return X, y


def timed_iters(fn, iters, ctx):
    """ms per call of fn (one warm-up, then `iters` calls between two CUDA events), max over ranks."""
    fn()
    ctx.barrier()
    sampler = ClockSampler(ctx.dev.index).start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = None
    for _ in range(iters):
        out = fn()
    e1.record()
    torch.cuda.synchronize(ctx.dev)
    clocks = sampler.stop()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1) / iters)
    return ms, out, clocks


def measure_fit_configs(ctx, n_big: int = 32768):
    from gpmpc_b200 import engine as E, sharding as S

    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    ell = torch.full((DIM,), 0.5 * math.sqrt(DIM), dtype=torch.float64)
    peaks = load_peaks()
    dmma_peak = peaks.get("dmma_tflops_bps8", 37.1)
    out = {}

    # MLL block, small sizes first (latency-bound: measured before the N = 32768 runs heat the chip into its power
    # cap; the clocks of each point are recorded)
    mll = {"unit": "MLL+gradient iterations/s (one GP, one GPU)", "points": []}
    time.sleep(1.0)
    for n in (2048, 16384):
        Xn, yn = fit_data(n, 1337, dev)

        def it():
            fit, ws = E.factorize(Xn, yn, ell, 1.0, 1e-2, want_kinv=True, want_image=False)
            return E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))

        ms_plain, res, clk = timed_iters(it, 20 if n <= 4096 else 3, ctx)
        # the same iteration replayed as a CUDA graph (what fit_gp does from its third iteration on)
        gm = E.GraphedMLL(Xn, yn)
        for _ in range(E.GRAPH_FIT_WARMUP + 1):
            gm(ell, 1.0, 1e-2)
        ms, res_g, clk = timed_iters(lambda: gm(ell, 1.0, 1e-2), 20 if n <= 4096 else 3, ctx)
        mll["points"].append({"n": n, "value": 1e3 / ms, "ms_per_iter": ms, "tflops_of_n3": float(n) ** 3 / ms * 1e-9,
                              "frac_of_fp64_dmma_peak": float(n) ** 3 / ms * 1e-9 / dmma_peak, "loss": res_g[0],
                              "path": "CUDA graph replay (engine.GraphedMLL)", "stream_launch_ms_per_iter": ms_plain,
                              "loss_stream_launch": res[0], "sm_mhz": clk.get("sm_mhz")})
        del Xn, yn, gm
        E.release_scratch()

    # (a) one GP per rank (different targets per rank): no collective
    X, y = fit_data(n_big, 1337 + rank, dev)

    def own_iter():
        fit, ws = E.factorize(X, y, ell, 1.0, 1e-2, want_kinv=True, want_image=False)
        return E.loss_and_grad_from_partials(fit, E.grad_partials(fit, ws))

    ms_a, res_a, clk_a = timed_iters(own_iter, 2, ctx)
    clk_all = ctx.gather_objects({"sm_mhz": clk_a.get("sm_mhz"), "reasons": clk_a.get("reasons")})
    out["config4a"] = {"what": f"MLL + gradient iteration, N={n_big}, d={DIM}, one GP per rank x {world} rank(s), no collective",
                       "ms_per_iter_max_over_ranks": ms_a, "iters_per_s_aggregate": world / ms_a * 1e3,
                       "tflops_of_n3_per_gpu": float(n_big) ** 3 / ms_a * 1e-9, "fit_gemm_digits": int(E.FIT_GEMM_DIGITS),
                       "kinv_digits": int(E.FIT_GEMM_DIGITS_KINV), "loss_rank0": res_a[0], "clocks_per_rank": clk_all}

    # (b) one GP (rank 0's data everywhere), K^-1 rows + trace split over the ranks, all-reduce of d+2 doubles
    if world > 1:
        X0, y0 = fit_data(n_big, 1337, dev)
        rs = S.rowsplit_grad_reduce(time_allreduce=True)

        def split_iter():
            fit, ws = E.factorize(X0, y0, ell, 1.0, 1e-2, want_kinv=rs.rows(n_big), want_image=False)
            return E.loss_and_grad_from_partials(fit, rs(fit, ws))

        ms_b, res_b, clk_b = timed_iters(split_iter, 2, ctx)
        ar = rs.allreduce_ms()[1:]                                 # drop the warm-up call
        ar_ms = ctx.max_over_ranks(sum(ar) / max(1, len(ar)))
        # the collective alone, ranks aligned by a barrier: latency of an all_reduce of d+2 doubles
        buf = torch.zeros(DIM + 2, dtype=torch.float64, device=dev)
        for _ in range(5):
            ctx.dist.all_reduce(buf)
        ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            ctx.dist.all_reduce(buf)
        e1.record()
        torch.cuda.synchronize(dev)
        ar_us = ctx.max_over_ranks(e0.elapsed_time(e1) / 50 * 1e3)
        # unsplit gradient of the same GP on this rank, for the parity figure
        X, y = X0, y0
        ref = own_iter()
        gs = torch.tensor([res_b[1], res_b[2]] + list(res_b[3]))
        gr = torch.tensor([ref[1], ref[2]] + list(ref[3]))
        dev_grad = ctx.max_over_ranks(float((gs - gr).abs().max() / gr.abs().max()))
        out["config4b"] = {"what": f"one GP, N={n_big}: every rank factorises (replicated), K^-1 rows + gradient trace split over "
                                   f"{world} ranks (sharding.RowSplitGradReduce), NCCL all_reduce of {DIM + 2} doubles",
                           "ms_per_iter_max_over_ranks": ms_b, "iters_per_s": 1e3 / ms_b,
                           "speedup_vs_unsplit": ms_a / ms_b,
                           "ceiling_note": "potrf + trtri (2/3 of the N^3 work) are replicated on every rank: <= 1.5x at any rank count",
                           "allreduce_ms_in_iter_max_over_ranks": ar_ms,
                           "allreduce_ms_note": "CUDA events around dist.all_reduce on the compute stream: wait for the slowest rank's "
                                                "trace + the collective",
                           "allreduce_us_isolated": ar_us, "rows_rank0": list(rs.rows(n_big)),
                           "grad_rel_dev_vs_unsplit": dev_grad, "clocks_rank0": {"sm_mhz": clk_b.get("sm_mhz"), "reasons": clk_b.get("reasons")}}
        del X0, y0
    del X, y
    E.release_scratch()
    torch.cuda.empty_cache()

    # the config-4 size point of the MLL block comes from (a)
    mll["points"].append({"n": n_big, "value": 1e3 / ms_a, "ms_per_iter": ms_a, "tflops_of_n3": float(n_big) ** 3 / ms_a * 1e-9,
                          "frac_of_fp64_dmma_peak": float(n_big) ** 3 / ms_a * 1e-9 / dmma_peak, "loss": res_a[0]})
    mll["value"] = mll["points"][-1]["value"]
    mll["n"] = n_big
    mll["roofline"] = {"bound": "tensor", "unit": "TFLOP/s", "peak": dmma_peak, "achieved": mll["points"][-1]["tflops_of_n3"],
                       "frac": mll["points"][-1]["frac_of_fp64_dmma_peak"],
                       "note": "N^3 flops per iteration (SURVEY.md 8d) against the FP64 DMMA pipe the iteration was specified on; "
                               "above 1 because the large GEMMs run as exact int8 digit products on tcgen05 "
                               f"({int(E.FIT_GEMM_DIGITS)} digits, K^-1 {int(E.FIT_GEMM_DIGITS_KINV)})",
                       "peak_source": "profiles/peaks_micro_r02.json dmma_tflops_bps8" if "dmma_tflops_bps8" in peaks
                                      else "37.1 TFLOP/s (round-1 measurement, profiles/peaks_micro_r01.json)"}
    if rank == 0 and world == 1:
        # CPU oracle beside it (closed-form loss + gradient, all host threads) at the size the reference can reach
        from oracle import gp_oracle as o

        torch.set_num_threads(os.cpu_count() or 1)
        g = torch.Generator().manual_seed(1337)
        Xc = torch.rand(2048, DIM, generator=g, dtype=torch.float64) * 2.0 - 1.0
        yc = torch.sin(Xc @ torch.randn(DIM, generator=g, dtype=torch.float64)) + 0.05 * torch.randn(2048, generator=g, dtype=torch.float64)
        rn, rs_, rl = o.inv_softplus(torch.tensor([1e-2 - 1e-6])), o.inv_softplus(torch.tensor(1.0)), o.inv_softplus(ell.reshape(1, -1))
        o.loss_and_grad(Xc[:256], yc[:256], rn, rs_, rl)
        t0 = time.perf_counter()
        o.loss_and_grad(Xc, yc, rn, rs_, rl)
        mll["cpu_oracle"] = {"n": 2048, "value": 1.0 / (time.perf_counter() - t0), "cores": os.cpu_count(), "kind": "port",
                             "what": "oracle.loss_and_grad (torch float64: Cholesky, cholesky_inverse, d trace passes), one call"}
    out["mll"] = mll
    return out


# --------------------------------------------------------------------------------------------------
# config 5: 65,536 rollouts x horizon 50 over the ranks
# --------------------------------------------------------------------------------------------------
def measure_config5(ctx, rollouts: int = 65536, horizon: int = 50, n_train: int = 2048):
    """Large-batch linearised-MPC sweep (BASELINE.json configs[4]).  Every rank owns rollouts/world initial states
    (sharding.rollouts_sharded) and walks the horizon: per stage t the three d=16 GPs are evaluated at z_t = (x_t, u_t)
    (mean, variance, Jacobian: what the linearisation and the tightening consume), and the RK4 stage dynamics with
    sensitivities on the reference's 12-state model (GPs with d = 1, 3, 3) produce x_{t+1}: stage t+1 depends on
    stage t, so the sweep is 50 sequential batches of rollouts/world queries.  The chance-constraint tightening
    recursion over the visited states closes the sweep."""
    from gpmpc_b200 import engine as E, sharding as S
    from gpmpc_b200.mpc import ConstraintTightening, StageDynamics

    dev, world = ctx.dev, ctx.world
    p = make_problem(n_train, 8)
    gb = E.fit_batch(p["X"].to(dev), p["Y"].to(dev), p["ell"], p["sf2"], p["noise"])
    g = torch.Generator().manual_seed(5)
    gps = []
    for k, d in enumerate((1, 3, 3)):
        Xk = torch.rand(n_train, d, generator=g, dtype=torch.float64) - 0.5
        yk = 0.3 * torch.sin(4.0 * Xk.sum(dim=1)) + 0.01 * torch.randn(n_train, generator=g, dtype=torch.float64)
        gps.append((Xk, yk, torch.full((d,), 0.3 + 0.1 * k, dtype=torch.float64), 0.5, 1e-2))
    sd = StageDynamics(gps, [12.1432, 1.8118, -72.08, -7.5755, 39.8653, -72.08, -7.5755, 39.8653], 0.02)
    A = torch.eye(12, dtype=torch.float64) + 0.02 * torch.randn(12, 12, generator=g, dtype=torch.float64)
    B = 0.1 * torch.randn(12, 4, generator=g, dtype=torch.float64)
    K = -0.3 * torch.randn(4, 12, generator=g, dtype=torch.float64)
    Bd = torch.eye(12, dtype=torch.float64)[:, [1, 3, 5, 9, 10]]
    ct = ConstraintTightening(gps, [[0], [1, 2, 3], [4, 5, 6]], A, B, K, Bd, 0.02, 2.0)
    x0_all = ((torch.rand(rollouts, 12, generator=g, dtype=torch.float64) - 0.5) * 0.8).to(dev)
    u_all = ((torch.rand(rollouts, 4, horizon, generator=g, dtype=torch.float64) - 0.3) * 0.6).to(dev)
    ev = {k: [] for k in ("gp", "dyn", "tighten")}

    def sweep_local(x0, u):
        """x0 (R,12), u (R,4,T) -> per-rollout summaries (R, 2): trace of the final A and the last stage's GP-0 variance."""
        R = x0.shape[0]
        xs = torch.empty(R, 12, horizon + 1, dtype=torch.float64, device=dev)
        xs[:, :, 0] = x0
        x = x0
        last_var = None
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(2 * horizon + 2)]
        for t in range(horizon):
            ut = u[:, :, t].contiguous()
            marks[2 * t].record()
            z = torch.cat([x, ut], dim=1)
            mean, var, jac = E.predict(gb, z)                      # 3 GPs x R queries: mean, variance, Jacobian
            marks[2 * t + 1].record()
            x, Ax, Bu = sd.disc_dyn(x, ut, work_bytes=2 << 30)     # RK4 + sensitivities (4 x 3-GP mean+Jacobian inside)
            xs[:, :, t + 1] = x
            last_var = var[0]
        marks[2 * horizon].record()
        sc, ic = ct.propagate_constraint_limits(xs, u)
        marks[2 * horizon + 1].record()
        ev["marks"] = marks
        return (torch.stack([Ax.diagonal(dim1=1, dim2=2).sum(1), last_var, sc[:, 0, -1]], dim=1),)

    def sweep():
        return S.rollouts_sharded(sweep_local, [x0_all, u_all])

    ms, res, clk = timed_iters(sweep, 2, ctx)
    lo, hi, (summary,) = res
    marks = ev["marks"]
    gp_ms = sum(marks[2 * t].elapsed_time(marks[2 * t + 1]) for t in range(horizon))
    dyn_ms = sum(marks[2 * t + 1].elapsed_time(marks[2 * t + 2]) for t in range(horizon))
    tight_ms = marks[2 * horizon].elapsed_time(marks[2 * horizon + 1])
    finite = bool(torch.isfinite(summary).all())
    finite = ctx.max_over_ranks(0.0 if finite else 1.0) == 0.0
    clk_all = ctx.gather_objects({"sm_mhz": clk.get("sm_mhz"), "reasons": clk.get("reasons")})
    out = {"config5": {
        "what": f"{rollouts} rollouts x horizon {horizon} over {world} rank(s) ({hi - lo} rollouts per rank): {horizon} sequential "
                f"stage batches [3-GP mean+var+Jacobian (N={n_train}, d=16) + RK4 stage dynamics with sensitivities "
                f"(GPs d=1,3,3, N={n_train})] + constraint tightening; sharding.rollouts_sharded, no collective",
        "ms_per_sweep_max_over_ranks": ms, "stage_evals_per_s_aggregate": rollouts * horizon / (ms * 1e-3),
        "gp_preds_per_s_aggregate": N_GP * rollouts * horizon / (ms * 1e-3),
        "rank0_breakdown_ms": {"gp_predictions": gp_ms, "stage_dynamics": dyn_ms, "tightening": tight_ms},
        "all_finite": finite, "clocks_per_rank": clk_all}}
    del gb, sd, ct
    E.release_scratch()
    torch.cuda.empty_cache()
    return out


# --------------------------------------------------------------------------------------------------
# N = 1 only: config 2 (round 1's headline), the mean+Jacobian kernel alone, float32 path, config 1, MPC rows
# --------------------------------------------------------------------------------------------------
def measure_single_gpu_extras(ctx, gb, workload):
    from gpmpc_b200 import engine as E
    from gpmpc_b200.gp import predict_host

    dev, flush = ctx.dev, ctx.flush
    out = {}
    n_train, n_query, _ = WORKLOADS["config2"]
    problem = make_problem(n_train, n_query)
    if gb is None:
        gb = E.fit_batch(problem["X"].to(dev), problem["Y"].to(dev), problem["ell"], problem["sf2"], problem["noise"])
    Z = problem["Z"].to(dev)
    npreds = N_GP * n_query
    c2 = {"what": workload_string("config2", n_train, n_query)}
    for _ in range(3):
        E.predict(gb, Z)
    steps = 20
    ms = timed_steps(lambda: E.predict(gb, Z), steps, flush, dev) / steps
    c2["value"], c2["ms_per_step"] = npreds / (ms * 1e-3), ms
    Zh = problem["Z"].pin_memory()
    oh = None
    for _ in range(3):
        oh = predict_host(gb, Zh, oh)
    t0 = time.perf_counter()
    for _ in range(steps):
        oh = predict_host(gb, Zh, oh)
    c2["e2e_value"] = npreds * steps / (time.perf_counter() - t0)
    roof = measure_roofline(gb, Z, ctx, 10)
    S_ = gb.digits
    roof["step_frac"] = float(n_train) ** 2 * S_ * (S_ + 1) / 2 * npreds / (ms * 1e-3) / 1e12 / roof["peak"]
    c2["roofline"] = {k: roof[k] for k in ("kernel", "achieved", "peak", "frac", "frac_of_nominal", "step_frac", "ms_per_launch",
                                           "fp64_equivalent_tflops")}
    # accuracy of the int8 variance path against the FP64 DMMA path, normwise and element-wise, per digit count
    v_ref = E.predict(gb, Z, var_path="f64")[1]
    sf2 = problem["sf2"]
    v5 = E.predict(gb, Z)[1]
    c2["var_rel_err_vs_dmma_path"] = {"digits": S_, "normwise": rel_dev(v5, v_ref), "elementwise_floor_1e-3_sf2": elementwise_var_err(v5, v_ref, sf2)}
    Z32 = Z.float()
    for _ in range(3):
        E.predict(gb, Z32)
    ms32 = timed_steps(lambda: E.predict(gb, Z32), 10, flush, dev) / 10
    c2["f32_path_preds_per_s"] = npreds / (ms32 * 1e-3)
    for _ in range(3):
        E.predict(gb, Z, var_path="f64")
    msd = timed_steps(lambda: E.predict(gb, Z, var_path="f64"), 10, flush, dev) / 10
    c2["dmma_var_path_preds_per_s"] = npreds / (msd * 1e-3)
    for S in (4, 6):
        gbs = E.fit_batch(problem["X"].to(dev), problem["Y"].to(dev), problem["ell"], problem["sf2"], problem["noise"], digits=S)
        for _ in range(3):
            E.predict(gbs, Z)
        msS = timed_steps(lambda: E.predict(gbs, Z), 10, flush, dev) / 10
        vS = E.predict(gbs, Z)[1]
        c2[f"digits{S}"] = {"preds_per_s": npreds / (msS * 1e-3), "var_normwise": rel_dev(vS, v_ref),
                            "var_elementwise_floor_1e-3_sf2": elementwise_var_err(vS, v_ref, sf2)}
        del gbs
    out["config2"] = c2
    # north-star kernel (a): mean + Jacobian alone, against the FP32 / FP64 FMA pipes (68 flops per pair at d=16)
    peaks = load_peaks()
    mj = {}
    for name, zz, pk in (("f64", Z, peaks.get("dmma_tflops_bps8", 37.1)), ("f32", Z32, peaks.get("ffma_tflops_bps8", 72.5))):
        for _ in range(3):
            E.predict(gb, zz, want_var=False)
        msj = timed_steps(lambda: E.predict(gb, zz, want_var=False), 10, flush, dev) / 10
        pairs = npreds * n_train / (msj * 1e-3)
        mj[name] = {"preds_per_s": npreds / (msj * 1e-3), "pairs_per_s": pairs, "tflops_at_68_per_pair": pairs * 68e-12,
                    "pipe_peak_tflops": pk, "frac": pairs * 68e-12 / pk}
    # the same at config 5's per-stage batch (Q = 65536): enough CTAs to fill the waves, per-CTA set-up amortised
    pz = make_problem(n_train, 65536)["Z"].to(dev)
    for name, zz, pk in (("f64_q65536", pz, peaks.get("dmma_tflops_bps8", 37.1)), ("f32_q65536", pz.float(), peaks.get("ffma_tflops_bps8", 72.5))):
        for _ in range(3):
            E.predict(gb, zz, want_var=False)
        msj = timed_steps(lambda: E.predict(gb, zz, want_var=False), 10, flush, dev) / 10
        pairs = N_GP * 65536 * n_train / (msj * 1e-3)
        mj[name] = {"preds_per_s": N_GP * 65536 / (msj * 1e-3), "pairs_per_s": pairs, "tflops_at_68_per_pair": pairs * 68e-12,
                    "pipe_peak_tflops": pk, "frac": pairs * 68e-12 / pk}
    del pz
    out["mean_jac_only"] = mj
    out.update(measure_mpc_rows(gb, problem, flush, dev))
    out.update(measure_config1(ctx))
    return out


def measure_config1(ctx):
    """BASELINE.json configs[0] restated as shapes: P=3 GPs, d=(1,3,3), N=45.  (i) the drop-in fit_gp
    exactly as the reference calls it (Adam lr 1e-3 <= 500 iterations from raw = 0, early stop, then K and K^-1) beside
    the CPU oracle's restatement of the same call; (ii) the cost of ONE Adam iteration (loss + gradient + step) on both
    sides, 200 iterations without the early stop; (iii) the Q=25 eval-mode covariance diagonal of gpmpc.py:441-445."""
    from gpmpc_b200 import gp as G
    from oracle import gp_oracle as o

    g = torch.Generator().manual_seed(1)
    N = 45
    X = torch.rand(N, 7, generator=g, dtype=torch.float64) - 0.5
    Y = 0.3 * torch.sin(3 * X[:, :3]) + 0.01 * torch.randn(N, 3, generator=g, dtype=torch.float64)
    idx = ([0], [1, 2, 3], [4, 5, 6])
    Xs = [X[:, i].contiguous() for i in idx]
    torch.set_num_threads(1)       # 45 x 45 matrices: one thread is the fastest CPU setting (16 threads: ~100x slower)

    def cpu_fit():
        n_it = 0
        for p in range(3):
            r = o.fit_hyperparameters(Xs[p], Y[:, p].contiguous(), n_train=500, lr=1e-3)
            noise, sf2, ell = o.constrained(r["raw_noise"].double(), r["raw_sf2"].double(), r["raw_ell"].double())
            K = o.assemble_K(Xs[p], ell.reshape(-1), sf2, noise)
            torch.linalg.inv(K)                                 # gp.K, gp.K_inv  (gpmpc/gp.py:45-46, 69)
            n_it += len(r["losses"])
        return n_it

    cpu_fit()
    t0 = time.perf_counter()
    iters_cpu = cpu_fit()
    t_cpu = time.perf_counter() - t0
    # one CPU iteration: closed-form loss + gradient + Adam step
    rn, rs, rl = (torch.zeros(1, requires_grad=True), torch.zeros((), requires_grad=True), torch.zeros(1, 1, requires_grad=True))
    opt = torch.optim.Adam([rn, rs, rl], lr=1e-3)
    y1 = Y[:, 1].contiguous()
    t0 = time.perf_counter()
    for _ in range(200):
        opt.zero_grad()
        loss, gn, gs, gl = o.loss_and_grad(Xs[1], y1, rn.detach().double(), rs.detach().double(), rl.detach().double())
        rn.grad, rs.grad, rl.grad = gn.float().reshape(1), gs.float().reshape(()), gl.float().reshape(1, 1)
        opt.step()
    cpu_iter_ms = (time.perf_counter() - t0) / 200 * 1e3
    torch.set_num_threads(os.cpu_count() or 1)

    def gpu_fit():
        n_it = 0
        for p in range(3):
            gp = G.GaussianProcess(Xs[p], Y[:, p].contiguous())
            G.fit_gp(gp, n_train=500, lr=1e-3, device="cpu")    # gp_mpc_config.yaml:19 device: "cpu" (tensors stay on the
                                                                # host, the arithmetic runs on the current CUDA device)
            n_it += len(gp.fit_losses)
        return n_it, gp

    gpu_fit()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    iters, gp = gpu_fit()
    torch.cuda.synchronize()
    t_gpu = time.perf_counter() - t0

    def gpu_fit_batched():      # the three GPs in lockstep: one launch per Adam iteration for all of them (gp.fit_gps)
        gps3 = [G.GaussianProcess(Xs[p], Y[:, p].contiguous()) for p in range(3)]
        G.fit_gps(gps3, n_train=500, lr=1e-3, device="cpu")
        return sum(len(q.fit_losses) for q in gps3)

    gpu_fit_batched()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    iters_b = gpu_fit_batched()
    torch.cuda.synchronize()
    t_gpu_b = time.perf_counter() - t0
    gp2 = G.GaussianProcess(Xs[1], y1)
    gp2.train()
    opt = torch.optim.Adam(gp2.parameters(), lr=1e-3)
    for _ in range(5):
        G._loss_and_set_grads(gp2, ctx.dev, None)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(200):
        opt.zero_grad()
        G._loss_and_set_grads(gp2, ctx.dev, None)
        opt.step()
        gp2._invalidate()
    torch.cuda.synchronize()
    gpu_iter_ms = (time.perf_counter() - t0) / 200 * 1e3
    z = torch.rand(25, 3, generator=g, dtype=torch.float64) - 0.5
    gp.eval()
    torch.diag(gp.likelihood(gp(z)).covariance_matrix).cpu()      # first call builds the prediction state
    t0 = time.perf_counter()
    for _ in range(20):
        torch.diag(gp.likelihood(gp(z)).covariance_matrix).cpu()
    t_eval = (time.perf_counter() - t0) / 20
    return {"config1": {"what": "reference shapes: 3 GPs d=(1,3,3), N=45; fit_gp as the reference calls it (Adam lr=1e-3 <= 500 "
                                "iterations, early stop, then K and K^-1), one Adam iteration alone, and the eval-mode covariance "
                                "diagonal at Q=25 (gpmpc/gpmpc.py:441-445)",
                        "gpu_fit_3gps_ms": t_gpu * 1e3, "gpu_iters": iters, "gpu_fit_gps_batched_ms": t_gpu_b * 1e3,
                        "gpu_batched_iters": iters_b, "cpu_oracle_fit_3gps_ms": t_cpu * 1e3,
                        "cpu_iters": iters_cpu, "cpu_threads": 1,
                        "gpu_ms_per_adam_iter": gpu_iter_ms, "cpu_oracle_ms_per_adam_iter": cpu_iter_ms,
                        "gpu_iter_path": "gpk_mll_small: one fused launch + one host round trip per iteration",
                        "gpu_eval_cov_diag_ms": t_eval * 1e3}}


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
