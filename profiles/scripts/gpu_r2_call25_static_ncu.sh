#!/bin/bash
# round 2, call 25: launch list (durations only) of the allocation-free solve's test; nothing printed under ncu is a bench value
mkdir -p gpurun_out
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2c25_static_launches.csv \
  python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2c25_static_ncu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r2c25_static_ncu.log; wc -l gpurun_out/r2c25_static_launches.csv
