#!/bin/bash
# round 2, call 14: long slices for the priority chains' die-off, with and without the one-bulk-block-beside placement
mkdir -p gpurun_out
bash tools/gpu_env_matrix.sh "-" "RT_PRIO_KEEP_ATT=128" "RT_PRIO_KEEP_ATT=128 RT_PRIO_SMEM_KB=130" "RT_PRIO_KEEP_ATT=128 RT_PRIO_SMEM_KB=130 RT_COOP_POLICY=1 RT_COOP_SUM=5000" "RT_PRIO_KEEP_ATT=128 RT_PRIO_SMEM_KB=130 RT_PRIO_SMEM_RAYS=14000" "RT_PRIO_KEEP_ATT=256 RT_PRIO_SMEM_KB=130" "RT_PRIO_KEEP_ATT=128 RT_PRIO_SMEM_KB=105" 2>&1 | tee gpurun_out/r2c14_matrix.log
RT_LIB=/root/repo/ray-trace_b200/libraytrace_b200_dev.so RT_DEBUG=1 RT_PRIO_KEEP_ATT=128 RT_PRIO_SMEM_KB=130 python bench.py --steps 1 --warmup 2 --cpu-sample 2000 > gpurun_out/r2c14_dbg.json 2> gpurun_out/r2c14_dbg.err
