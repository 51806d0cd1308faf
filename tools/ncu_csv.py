"""ncu raw page of a .ncu-rep -> 'metric,unit,value' CSV of ONE kernel launch (the longest launch whose name matches).
Usage: python tools/ncu_csv.py prof.ncu-rep <kernel name regex> > profiles/ncu_<kernel>_r02.csv   (needs ncu; no GPU)"""
import csv, io, re, subprocess, sys

rep, pat = sys.argv[1], re.compile(sys.argv[2])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
kn, du = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
cand = [r for r in data if pat.search(r[kn])]
assert cand, "no launch matches"
best = max(cand, key=lambda r: float(r[du].replace(",", "")))
print("metric,unit,value")
print(f"kernel,,{best[kn]}")
print(f"launches_matching,,{len(cand)}")
for h, u, v in zip(hdr, units, best):
    if "__" in h and v != "":
        print(f"{h},{u},{v.replace(',', '')}")
