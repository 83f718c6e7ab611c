#!/bin/bash
# ncu full captures: one bulk launch and one cooperative launch of the 3D integrate kernel (developer tool)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --cpu-sample 2000"
$CMD > gpurun_out/plain_p2.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0E -s 8 -c 2 -o gpurun_out/prof_bulk -f $CMD > gpurun_out/ncu_bulk.log 2>&1
$CMD > gpurun_out/plain_p3.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb1E -s 6 -c 1 -o gpurun_out/prof_coop -f $CMD > gpurun_out/ncu_coop.log 2>&1
$CMD > gpurun_out/plain_p4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_v2.csv $CMD > gpurun_out/ncu_l2.log 2>&1
ls -la gpurun_out/*.ncu-rep
