#!/bin/bash
# round-1 profile v3: bench line, ncu full captures (steady full-list launch, die-off launch, cooperative launch), launch list
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --cpu-sample 4000"
$CMD > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err || { tail -5 gpurun_out/bench_v3.err; exit 1; }
python tools/prof_slices.py > gpurun_out/plain_slices.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0E -s 2 -c 1 -o gpurun_out/prof3_steady -f python tools/prof_slices.py > gpurun_out/ncu3_a.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0E -s 8 -c 1 -o gpurun_out/prof3_dieoff -f python tools/prof_slices.py > gpurun_out/ncu3_b.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb1E -s 6 -c 1 -o gpurun_out/prof3_coop -f python tools/prof_slices.py > gpurun_out/ncu3_c.log 2>&1
CMD2="python bench.py --steps 1 --warmup 1 --cpu-sample 2000"
$CMD2 > gpurun_out/plain_v3b.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_v3.csv $CMD2 > gpurun_out/ncu3_l.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -c 600 gpurun_out/bench_v3.json
