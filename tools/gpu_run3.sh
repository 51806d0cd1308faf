#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q --durations=8 ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
python tools/front_splits_sweep.py 2>&1 | tee gpurun_out/front_splits.log
timeout 600 python bench.py --workload config2 --steps 20 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_c2.err
python -c "
import json; l=json.load(open('gpurun_out/bench_c2.json')); print(l['value'], l['ms_per_step'], l['e2e']['value']); print(l['extras']['config1']); print(l['mll']['points'][0])"
