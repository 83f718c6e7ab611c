#!/bin/bash
# round 2, call 1: parity suite, bench line after the controller deadband, chain timeline, DFMA peak under ncu
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > gpurun_out/r2c1_smi.log 2>&1
nproc >> gpurun_out/r2c1_smi.log
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c1_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2c1_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2c1_bench.json 2> gpurun_out/r2c1_bench.err; echo "bench rc=$?"
RT_DEBUG=1 python bench.py --steps 1 --warmup 3 --cpu-sample 2000 > gpurun_out/r2c1_dbg.json 2> gpurun_out/r2c1_dbg.err
python tools/peak_probe.py > gpurun_out/r2c1_peak_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.avg,sm__cycles_elapsed.avg.per_second,sm__cycles_active.avg,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,sm__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__issue_active.avg.per_cycle_active \
    --clock-control none -k regex:dfma_peak --csv --log-file gpurun_out/r2c1_peak_ncu.csv python tools/peak_probe.py > gpurun_out/r2c1_peak_ncu.log 2>&1
cat gpurun_out/r2c1_peak_plain.log
tail -c 1500 gpurun_out/r2c1_bench.json
