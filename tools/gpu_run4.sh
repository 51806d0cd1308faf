#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
rm -f build/variants/*.so; mkdir -p build/variants; cp gpmpc_b200/libgpmpc_b200.so build/variants/lib_current.so
python tools/time_mj_variants.py 2>&1 | tee gpurun_out/mj_variants3.log
python tools/front_splits_sweep.py 2>&1 | head -2
