#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_env_matrix.sh "-" "RT_COOP_WIN_LO=192 RT_COOP_WIN_HI=448" "RT_COOP_WIN_LO=192 RT_COOP_WIN_HI=576" "RT_COOP_WIN_LO=192 RT_COOP_WIN_HI=320" "RT_COOP_WIN_LO=448 RT_COOP_WIN_HI=704" "RT_COOP_WIN_LO=576 RT_COOP_WIN_HI=832" "RT_COOP_WIN_LO=128 RT_COOP_WIN_HI=384 RT_COOP_ATT=16" "RT_COOP_WIN_LO=320 RT_COOP_WIN_HI=576 RT_COOP_ATT=64" 2>&1 | tee gpurun_out/r2c3_matrix.log
RT_DEBUG=1 RT_COOP_WIN_LO=192 RT_COOP_WIN_HI=448 python bench.py --steps 1 --warmup 2 --cpu-sample 2000 > gpurun_out/r2c3_dbg.json 2> gpurun_out/r2c3_dbg.err
