#!/bin/bash
timeout 100 python tools/trace_pipe.py 2048 8192 3 32 2>&1 | tail -4
timeout 100 python tools/trace_pipe.py 16384 18944 3 37 2>&1 | tail -6
