#!/bin/bash
# chain-1 cooperative mode under contention, with the bulk kernel's residency limited by shared memory
for cfg in "1 32 0" "0 64 0" "0 64 80" "0 128 80" "0 64 115" "1 32 80"; do
  set -- $cfg
  echo "=== policy=$1 coop_att=$2 bulk_smem_kb=$3"
  RT_COOP_POLICY=$1 RT_COOP_ATT=$2 RT_BULK_SMEM_KB=$3 RT_DEBUG=1 python tools/dev_gpu_check.py cfg3 2>&1 | grep -E "^\[rt\] chain|integrate|ms" | awk '/chain 1/{l1=$0} /chain 0/{l0=$0} !/chain/{print} END{print l0; print l1}'
done
