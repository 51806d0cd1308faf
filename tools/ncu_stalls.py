"""Print the warp-stall breakdown (cycles per issued instruction by reason) of the first kernel in an ncu raw CSV."""
import csv, sys
r = list(csv.reader(open(sys.argv[1])))
h, v = r[0], r[2]
out = []
for i, n in enumerate(h):
    if "issue_stalled" in n and n.endswith("per_warp_active.pct") is False and "ratio" in n:
        try:
            out.append((float(v[i].replace(",", "")), n))
        except ValueError:
            pass
for val, n in sorted(out, reverse=True)[:12]:
    print(f"{val:8.3f}  {n}")
for key in ("smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fma.sum",
            "sm__inst_executed_pipe_alu.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__cycles_active.avg"):
    for i, n in enumerate(h):
        if n == key:
            print(key, v[i])
