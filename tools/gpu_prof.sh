#!/bin/bash
# usage: gpu_prof.sh <lib.so> <tag> : timeline + ncu full capture of two integrate launches (developer tool)
export RT_LIB=/root/repo/ray-trace_b200/$1
tag=$2
mkdir -p gpurun_out
RT_DEBUG=1 python tools/dev_gpu_check.py cfg3 > gpurun_out/timeline_$tag.log 2>&1
grep -E "rep1|DFMA" gpurun_out/timeline_$tag.log; grep "\[rt\]" gpurun_out/timeline_$tag.log | tail -24 | head -12
python bench.py --steps 1 --warmup 1 --cpu-sample 2000 > gpurun_out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rt_integrate -s 2 -c 2 -o gpurun_out/prof_$tag -f python bench.py --steps 1 --warmup 1 --cpu-sample 2000 > gpurun_out/ncu_$tag.log 2>&1
ls -la gpurun_out/prof_$tag.ncu-rep
