#!/bin/bash
# round 2, call 23: rt_solve_3d_device_static against rt_solve_3d_device on short launch lists
mkdir -p gpurun_out
timeout 150 python tools/bench_static.py 15 > gpurun_out/r2c23_static_vs_device.json 2> gpurun_out/r2c23_static.err; echo "rc=$?"; cat gpurun_out/r2c23_static_vs_device.json | head -60; tail -3 gpurun_out/r2c23_static.err
