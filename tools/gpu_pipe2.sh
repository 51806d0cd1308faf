#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/time_pipe.py 2048 8192 3 2>&1 | head -2
timeout 120 python tools/time_mj.py 2>&1 | tail -3
timeout 200 python -m pytest tests/test_gpu_ozaki.py tests/test_gpu_kernels.py -m gpu -q -x 2>&1 | tail -2
