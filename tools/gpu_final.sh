#!/bin/bash
# end-of-round evidence with the final code: full GPU test suite, launch list of one N = 2048 MLL + gradient iteration,
# full captures of the two new fit kernels
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_fit_n2048_r02_v4.csv python tools/fit_iter.py 2048 1 > gpurun_out/ncu_l.log 2>&1; echo "launch list rc=$?"
python tools/agg_launches.py gpurun_out/launches_fit_n2048_r02_v4.csv 2>/dev/null | head -8
timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:potf2_mma -s 4 -c 1 \
    -o gpurun_out/prof_potf2_mma_r02 -f python tools/fit_iter.py 2048 1 > gpurun_out/ncu_p.log 2>&1; echo "capture rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:trsm128 -s 8 -c 1 \
    -o gpurun_out/prof_trsm128_r02 -f python tools/fit_iter.py 2048 1 > gpurun_out/ncu_t.log 2>&1; echo "capture rc=$?"
timeout 200 python tools/graph_fit_experiment.py 2>&1 | tail -5
