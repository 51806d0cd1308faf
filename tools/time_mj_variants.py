"""A/B timing of mean+Jacobian-only prediction (config-2 shape) over library builds in build/variants/ (GPMPC_B200_LIB)."""
import glob, os, subprocess, sys
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    sys.path.insert(0, '.')
    import bench
    from gpmpc_b200 import engine as E
    dev = torch.device("cuda", 0)
    p = bench.make_problem(2048, 65536)
    gb = E.fit_batch(p["X"].to(dev), p["Y"].to(dev), p["ell"], p["sf2"], p["noise"], digits=5)
    flush = bench.l2_flusher(dev)
    for Q in (8192, 65536):
        for name, Z in (("f32", p["Z"][:Q].to(dev).float()), ("f64", p["Z"][:Q].to(dev))):
            for kw in ({"want_var": False}, {"want_var": False, "mj_path": "simt"}):
                if name == "f32" and "mj_path" in kw: continue
                for _ in range(3): E.predict(gb, Z, **kw)
                ms = bench.timed_steps(lambda: E.predict(gb, Z, **kw), 20, flush, dev) / 20
                print(f"  Q={Q} {name} {kw.get('mj_path','auto')}: {ms:.4f} ms = {3*Q*2048/ms*1e-6:.1f} G pairs/s", flush=True)
else:
    for lib in sorted(glob.glob("build/variants/lib_*.so")):
        print(lib, flush=True)
        subprocess.run([sys.executable, __file__, "child"], env={**os.environ, "GPMPC_B200_LIB": os.path.abspath(lib)})
