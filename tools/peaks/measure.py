"""Runs the pipe micro-benchmarks on the GPU box and writes one merged JSON (default gpurun_out/peaks_micro.json):
tools/peaks/peaks (FP64 DFMA / DMMA, FP32 FFMA, MUFU ex2) and tools/peaks/umma_bench --json (tcgen05 int8 dense rate).
Build both first (nvcc cross-compiles in the build container; the binaries travel with the gpurun snapshot):
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/peaks/peaks tools/peaks/peaks.cu
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -Igpmpc_b200/csrc -o tools/peaks/umma_bench tools/peaks/umma_bench.cu
"""
import json
import os
import subprocess
import sys

here = os.path.dirname(os.path.abspath(__file__))
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "..", "..", "gpurun_out", "peaks_micro.json")
merged = {}
for cmd in ([os.path.join(here, "peaks")], [os.path.join(here, "umma_bench"), "--json"]):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not line:
        print("failed:", cmd, r.stdout[-500:], r.stderr[-500:])
        continue
    merged.update(json.loads(line[-1]))
os.makedirs(os.path.dirname(out), exist_ok=True)
with open(out, "w") as f:
    json.dump(merged, f, indent=1)
print(json.dumps(merged))
