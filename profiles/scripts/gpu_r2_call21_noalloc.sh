#!/bin/bash
# round 2, call 21: rt_solve_3d_device_static (the solve without device allocations).
# (1) its test alone in a fresh process, (2) after the 44 parity tests in one process, (3) the rest of the GPU suite
mkdir -p gpurun_out
( time timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/r2c21_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r2c21_alone.log
( time timeout 260 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q --durations=5 ) > gpurun_out/r2c21_after.log 2>&1; echo "after rc=$?"; tail -12 gpurun_out/r2c21_after.log
( time timeout 200 python -m pytest tests/test_gpu_depth.py -m gpu -x -q ) > gpurun_out/r2c21_depth.log 2>&1; echo "depth rc=$?"; tail -4 gpurun_out/r2c21_depth.log
