#!/bin/bash
# round 2, call 22: last sanity of what the driver runs at round end, on the rebuilt libraries: smoke() and the default bench
mkdir -p gpurun_out
( time timeout 90 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > gpurun_out/r2c22_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/r2c22_smoke.log
( time timeout 230 python bench.py ) > gpurun_out/r2c22_bench.log 2>&1; echo "bench rc=$?"; tail -5 gpurun_out/r2c22_bench.log
