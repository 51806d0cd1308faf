#!/bin/bash
# round-2 ncu evidence for the headline workload (config 3, N=16384): launch list of the timed steps + one full capture
# of the dominant kernel and of the front kernel.  bench.py brackets its timed steps with cudaProfilerStart/Stop.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --queries 34816 --no-extras --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain_run.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain_run.log; exit 86; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_r02_config3.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:oz_var_kernel -c 1 \
    -o gpurun_out/prof_var_r02 -f $CMD > gpurun_out/ncu_var.log 2>&1; echo "var capture rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:oz_front_kernel -c 1 \
    -o gpurun_out/prof_front_r02 -f $CMD > gpurun_out/ncu_front.log 2>&1; echo "front capture rc=$?"
# config 2 step as well (round-1 headline): launch list
CMD2="python bench.py --workload config2 --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_r02_config2.csv $CMD2 > gpurun_out/ncu_launches2.log 2>&1; echo "launch list 2 rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_r02_*.csv
