#!/bin/bash
# round 2, call 24: the driver's round-end form of the GPU suite, one process, final tree
mkdir -p gpurun_out
( time timeout 285 python -m pytest tests -x -q -m gpu --durations=8 ) > gpurun_out/r2c24_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -16 gpurun_out/r2c24_pytest_gpu.log
