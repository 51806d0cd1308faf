"""Config-2 step time against the number of k-splits of the prediction front end (GPK_OZ_FRONT_SPLITS override)."""
import os, sys, torch
sys.path.insert(0, '.')
import bench
from gpmpc_b200 import engine as E
dev = torch.device("cuda", 0)
p = bench.make_problem(2048, 8192)
gb = E.fit_batch(p["X"].to(dev), p["Y"].to(dev), p["ell"], p["sf2"], p["noise"], digits=5)
Z = p["Z"].to(dev)
flush = bench.l2_flusher(dev)
for ns in ["auto", 1, 2, 3, 4, 6, 8, 12, 16]:
    if ns == "auto": os.environ.pop("GPK_OZ_FRONT_SPLITS", None)
    else: os.environ["GPK_OZ_FRONT_SPLITS"] = str(ns)
    for mode, kw in (("step", {}), ("mean_jac_f64", {"want_var": False})):
        for _ in range(3): E.predict(gb, Z, **kw)
        ms = bench.timed_steps(lambda: E.predict(gb, Z, **kw), 20, flush, dev) / 20
        print(f"splits={ns} {mode}: {ms:.4f} ms", flush=True)
