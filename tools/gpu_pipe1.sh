#!/bin/bash
mkdir -p gpurun_out
( time timeout 240 python -m pytest tests/test_gpu_ozaki.py -m gpu -q -x ) > gpurun_out/pytest_pipe.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_pipe.log
timeout 120 python tools/time_pipe.py 2048 8192 3 > gpurun_out/pipe_c2.log 2>&1; echo "rc=$?"; cat gpurun_out/pipe_c2.log
timeout 240 python tools/time_pipe.py 16384 28416 3 > gpurun_out/pipe_c3.log 2>&1; echo "rc=$?"; cat gpurun_out/pipe_c3.log
