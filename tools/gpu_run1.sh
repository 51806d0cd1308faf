#!/bin/bash
# round-2 GPU call 1: parity tests, first bench line of the new workload, peaks, digit-error calibration
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tools/peaks/measure.py gpurun_out/peaks_micro_r02.json > gpurun_out/peaks.log 2>&1; tail -2 gpurun_out/peaks.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_n1.json; tail -5 gpurun_out/bench_n1.err
timeout 600 python tools/digit_error_sweep.py > gpurun_out/digit_sweep.log 2>&1; tail -4 gpurun_out/digit_sweep.log
