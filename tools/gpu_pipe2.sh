#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/fit_iter.py 2048 3 2>&1 | tail -1
timeout 200 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:potf2_kernel -s 3 -c 1 \
    -o gpurun_out/prof_potf2_r02 -f python tools/fit_iter.py 2048 1 > gpurun_out/ncu_potf2.log 2>&1; echo "rc=$?"
ls -la gpurun_out/prof_potf2_r02.ncu-rep
