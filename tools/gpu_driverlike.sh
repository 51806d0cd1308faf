#!/bin/bash
# what the driver runs at round end on one GPU: reference arm, then our arm, with its flags; wall-clock of each
mkdir -p gpurun_out
( time python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/ref_n1.json 2> gpurun_out/ref_n1.err ) 2> gpurun_out/ref_n1.time; tail -3 gpurun_out/ref_n1.time
( time python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_n1_full.json 2> gpurun_out/bench_n1_full.err ) 2> gpurun_out/bench_n1_full.time; tail -3 gpurun_out/bench_n1_full.time; tail -3 gpurun_out/bench_n1_full.err
python - <<PY
import json
r=json.loads([x for x in open('gpurun_out/ref_n1.json') if x.startswith('{')][-1]); print('ref', r['value'], r['ms_per_step'], r['cpu_baseline']['sample'][:80])
l=json.loads([x for x in open('gpurun_out/bench_n1_full.json') if x.startswith('{')][-1])
print(l['value'], l['ms_per_step'], l['e2e']['value'], l['roofline']['frac'], l['roofline']['step_frac'], l['roofline']['peak_source'][:60], l['clocks'])
print('ratio e2e', l['e2e']['value']/r['value'])
PY
python __graft_entry__.py smoke 2>&1 | tail -2
