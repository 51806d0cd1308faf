#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_fit_n16384_r02.csv python tools/fit_iter.py 16384 1 > gpurun_out/ncu_l16.log 2>&1; echo "launch list rc=$?"
python tools/agg_launches.py gpurun_out/launches_fit_n16384_r02.csv
