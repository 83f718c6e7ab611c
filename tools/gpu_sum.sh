#!/bin/bash
for cfg in "2048 128" "2048 64" "2048 256" "12000 128" "25000 128" "25000 64" "50000 128"; do
  set -- $cfg
  echo "=== coop_sum=$1 wait_att=$2"
  RT_COOP_SUM=$1 RT_WAIT_ATT=$2 RT_DEBUG=1 python tools/dev_gpu_check.py cfg3 2>&1 | grep -E "^\[rt\] chain|rep" | awk '/chain 1/{l1=$0} /chain 0/{l0=$0} /rep/{print} END{print l0; print l1}'
done
