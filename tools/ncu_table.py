"""ncu raw-page CSV (ncu -i X.ncu-rep --page raw --csv) -> markdown table: per kernel name the LONGEST launch."""
import csv, re, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
col = {h: i for i, h in enumerate(hdr)}

def get(r, name, default=""):
    i = col.get(name)
    if i is None:      # some metrics only appear with a section prefix ("FBSP.TriageCompute.dram__throughput...")
        i = next((j for h, j in col.items() if h.endswith("." + name)), None)
    return r[i] if i is not None and i < len(r) and r[i] != "" else default

def f(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return float("nan")

best = {}
count = {}
for r in data:
    name = re.sub(r"\(.*", "", get(r, "Kernel Name")).replace("void ", "").replace("gpk::", "")
    dur = f(get(r, "gpu__time_duration.sum"))
    count[name] = count.get(name, 0) + 1
    if name not in best or dur > best[name][0]:
        best[name] = (dur, r)
unit = rows[1][col["gpu__time_duration.sum"]]
print(f"| kernel | launches | longest ({unit}) | grid x block | regs | tensor pipe % | FP64 pipe % | FMA pipe % | ALU pipe % | issue active % | DRAM GB/s | DRAM % of peak | L2 hit % | achieved occupancy % |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for name, (dur, r) in sorted(best.items(), key=lambda kv: -kv[1][0]):
    rd, wr = f(get(r, "dram__bytes_read.sum.per_second")), f(get(r, "dram__bytes_write.sum.per_second"))
    bw_unit = rows[1][col["dram__bytes_read.sum.per_second"]] if "dram__bytes_read.sum.per_second" in col else ""
    scale = {"Tbyte/s": 1e3, "Gbyte/s": 1.0, "Mbyte/s": 1e-3, "Kbyte/s": 1e-6, "byte/s": 1e-9}.get(bw_unit, 1.0)
    print("| `%s` | %d | %.4g | %s x %s | %s | %.1f | %.1f | %.1f | %.1f | %.1f | %.0f | %.1f | %.1f | %.1f |" % (
        name[:60], count[name], dur, get(r, "launch__grid_size"), get(r, "launch__block_size"),
        get(r, "launch__registers_per_thread"),
        f(get(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "nan")),
        f(get(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "nan")),
        f(get(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "nan")),
        f(get(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "nan")),
        f(get(r, "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "nan")),
        (rd + wr) * scale,
        f(get(r, "dram__throughput.avg.pct_of_peak_sustained_elapsed", "nan")),
        f(get(r, "lts__t_sector_hit_rate.pct", "nan")),
        f(get(r, "sm__warps_active.avg.pct_of_peak_sustained_active", "nan"))))
