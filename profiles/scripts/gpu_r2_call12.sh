#!/bin/bash
# round 2, call 12: ncu --set full captures of the final kernel: a full-list launch and the die-off launch
mkdir -p gpurun_out
python tools/prof_slices.py > gpurun_out/r2c12_plain_slices.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0ELb0E -s 2 -c 1 -o gpurun_out/r2c12_steady -f python tools/prof_slices.py > gpurun_out/r2c12_ncu_a.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0ELb0E -s 9 -c 1 -o gpurun_out/r2c12_dieoff -f python tools/prof_slices.py > gpurun_out/r2c12_ncu_b.log 2>&1
ls -la gpurun_out/*.ncu-rep; cat gpurun_out/r2c12_plain_slices.log
