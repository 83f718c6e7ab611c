#!/bin/bash
# round 2, call 17: when the early result copy of the host-pointer entry point should start (rays still running at the snapshot = n / RT_EARLY_DIV)
mkdir -p gpurun_out
bash tools/gpu_env_matrix.sh "-" "RT_EARLY_DIV=32" "RT_EARLY_DIV=64" "RT_EARLY_DIV=128" "RT_EARLY_DIV=256" "RT_EARLY_DIV=2048" "RT_EARLY_MIN=100000000" 2>&1 | tee gpurun_out/r2c17_matrix.log
