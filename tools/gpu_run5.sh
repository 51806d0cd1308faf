#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_n1.err
python - <<PY
import json
l=json.loads([x for x in open('gpurun_out/bench_n1.json') if x.startswith('{')][-1])
print(l['value'], l['ms_per_step'], l['roofline']['frac'], l['roofline']['traffic'])
print(json.dumps(l['mll']['points'])); print(l['extras']['config1']); print(l['extras']['mean_jac_only']); print(l['extras']['config2']['value'], l['extras']['config2']['ms_per_step'])
PY
