#!/bin/bash
# developer tool: bench.py (cfg3) under a list of environment settings, one line of results each
# usage: tools/gpu_env_matrix.sh "A=1 B=2" "A=3" ...   ("-" = defaults)
mkdir -p gpurun_out
export RT_LIB=/root/repo/ray-trace_b200/libraytrace_b200_dev.so   # the environment switches exist in the developer build only
i=0
for setting in "$@"; do
  i=$((i+1))
  if [ "$setting" = "-" ]; then envs=""; else envs="$setting"; fi
  env $envs python bench.py --steps 3 --warmup 2 --cpu-sample 2000 > gpurun_out/matrix_$i.json 2> gpurun_out/matrix_$i.err
  rc=$?
  python - "$setting" "$rc" gpurun_out/matrix_$i.json <<'PY'
import json, sys
setting, rc, path = sys.argv[1:4]
try:
    d = json.loads(open(path).read().strip().splitlines()[-1])
    print("%-60s ms/step %.2f  frac %.4f  e2e_ms %.2f  steady %.3f  launches %d" % (setting, d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["ms_per_step"], d["roofline"]["steady_state"]["frac"], d["roofline"]["launches"]))
except Exception as e:
    print("%-60s FAILED rc=%s %s" % (setting, rc, e))
PY
done
