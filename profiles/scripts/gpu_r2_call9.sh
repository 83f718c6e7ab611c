#!/bin/bash
# round 2, call 9: FP64 pipe probes (what real instruction streams sustain), f4 / grazing tests
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "analytic_partials or grazing or primitives" > gpurun_out/r2c9_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c9_pytest.log
python tools/pipe_probe.py > gpurun_out/r2c9_pipe_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.per_cycle_active,smsp__inst_executed.sum,sm__cycles_elapsed.avg.per_second \
    --clock-control none -k regex:rt_pipe_kernel --csv --log-file gpurun_out/r2c9_pipe_ncu.csv python tools/pipe_probe.py > gpurun_out/r2c9_pipe_ncu.log 2>&1
cat gpurun_out/r2c9_pipe_plain.log
