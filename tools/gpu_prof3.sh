#!/bin/bash
# round-2 (late): launch list of one MLL + gradient iteration at N = 2048 with the DMMA leaf, and a full capture of the leaf
mkdir -p gpurun_out
timeout 120 python tools/fit_iter.py 2048 3 2>&1 | tail -1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_fit_n2048_r02_v2.csv python tools/fit_iter.py 2048 1 > gpurun_out/ncu_l.log 2>&1; echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:potf2_mma -s 4 -c 1 \
    -o gpurun_out/prof_potf2_mma_r02 -f python tools/fit_iter.py 2048 1 > gpurun_out/ncu_p.log 2>&1; echo "capture rc=$?"
python tools/agg_launches.py gpurun_out/launches_fit_n2048_r02_v2.csv
