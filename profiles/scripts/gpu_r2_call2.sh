#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_env_matrix.sh "-" "RT_PRIO_SMEM_KB=130" "RT_PRIO_SMEM_KB=150" "RT_PRIO_SMEM_KB=180" "RT_PRIO_SMEM_KB=130 RT_PRIO_SMEM_RAYS=60000" "RT_PRIO_SMEM_KB=130 RT_PRIO_SMEM_RAYS=30000" "RT_PRIO_SMEM_KB=130 RT_DIEOFF_ATT=2" "RT_DIEOFF_ATT=2" "RT_PRIO_SMEM_KB=130 RT_COOP_SUM=60000" "RT_PRIO_SMEM_KB=130 RT_COOP_SUM=8000" "RT_PRIO_SMEM_KB=100" 2>&1 | tee gpurun_out/r2c2_matrix.log
RT_DEBUG=1 RT_PRIO_SMEM_KB=130 python bench.py --steps 1 --warmup 2 --cpu-sample 2000 > gpurun_out/r2c2_dbg.json 2> gpurun_out/r2c2_dbg.err
