"""Measured error of the int8-digit variance path against the FP64 DMMA path over conditioning (noise), input
dimension and training-set size, next to the quantities an error model can use at fit time (max |W|, N, sf2):
the calibration data behind engine.choose_digits.  Writes gpurun_out/digit_error_sweep.json.
Usage (GPU box): python tools/digit_error_sweep.py"""
import json, math, os, sys, torch
sys.path.insert(0, '.')
from gpmpc_b200 import engine as E

dev = torch.device("cuda", 0)
rows = []
for d, ell0 in ((1, 0.5), (3, 0.8), (16, 2.0)):
    for n in (256, 1024, 2048, 8192) + ((16384,) if d == 16 else ()):
        for noise in (1e-2, 1e-4, 1e-5, 1e-6):
            for sf2 in (1.0, 4.0):
                if sf2 != 1.0 and (n != 1024):
                    continue
                g = torch.Generator().manual_seed(1000 * d + n)
                X = torch.rand(n, d, generator=g, dtype=torch.float64) * 2 - 1
                y = math.sqrt(sf2) * torch.sin(X @ torch.randn(d, generator=g, dtype=torch.float64)) + math.sqrt(noise) * torch.randn(n, generator=g, dtype=torch.float64)
                Zr = (torch.rand(512, d, generator=g, dtype=torch.float64) * 2 - 1) * 1.2
                Z = torch.cat([X[:: max(1, n // 512)][:512], Zr]).to(dev)
                ell = torch.full((d,), ell0, dtype=torch.float64)
                try:
                    fit, ws = E.factorize(X.to(dev), y.to(dev), ell, sf2, noise, digits=6)
                except E.NotPSDError as e:
                    rows.append({"d": d, "n": n, "noise": noise, "sf2": sf2, "error": str(e)[:80]})
                    continue
                wmax = float(fit.W.abs().max())
                v_ref = E.predict(fit, Z, var_path="f64")[1][0]
                row = {"d": d, "n": n, "noise": noise, "sf2": sf2, "jitter": fit.jitter, "wmax": wmax,
                       "var_min": float(v_ref.min()), "var_max": float(v_ref.max())}
                for S in (4, 5, 6):
                    fs, _ = E.factorize(X.to(dev), y.to(dev), ell, sf2, noise, digits=S)
                    v = E.predict(fs, Z)[1][0]
                    err = (v - v_ref).abs()
                    row[f"abs{S}"] = float(err.max())
                    row[f"abs{S}_train"] = float(err[:512].max())
                    row[f"model_c{S}"] = float(err.max()) / (sf2 ** 1.5 * wmax * math.sqrt(n) * 2.0 ** (-8 * S))
                    if S == 5:
                        v32 = E.predict(fs, Z.float())[1][0].double()
                        row["abs_f32path"] = float((v32 - v_ref).abs().max())
                rows.append(row)
                print(json.dumps(row), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/digit_error_sweep.json", "w"), indent=1)
cs = {S: sorted(r[f"model_c{S}"] for r in rows if f"model_c{S}" in r) for S in (4, 5, 6)}
for S in (4, 5, 6):
    c = cs[S]
    print(f"S={S}: model constant C = err / (sf2^1.5 wmax sqrt(N) 2^-8S): min {c[0]:.3g} median {c[len(c)//2]:.3g} max {c[-1]:.3g}")
