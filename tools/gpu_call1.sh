#!/bin/bash
# One GPU call: parity tests, bench, debug timeline, ncu launch list + one full capture of the hot kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_cfg3.json
RT_DEBUG=1 python tools/dev_gpu_check.py cfg3 cfg2s > gpurun_out/dev_check.log 2>&1
tail -40 gpurun_out/dev_check.log
python bench.py --steps 1 --warmup 1 --cpu-sample 2000 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --cpu-sample 2000 > gpurun_out/ncu_l.log 2>&1
python bench.py --steps 1 --warmup 1 --cpu-sample 2000 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_integrate -s 2 -c 2 -o gpurun_out/prof_integrate -f python bench.py --steps 1 --warmup 1 --cpu-sample 2000 > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out
