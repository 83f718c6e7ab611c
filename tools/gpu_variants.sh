#!/bin/bash
# Run the parity tests on one variant build and the bench on each variant (developer tool).
mkdir -p gpurun_out
first=1
for lib in "$@"; do
  export RT_LIB=/root/repo/ray-trace_b200/$lib
  if [ $first = 1 ]; then
    python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$lib.log 2>&1; echo "pytest($lib) rc=$?"; tail -3 gpurun_out/pytest_$lib.log
    first=0
  fi
  python bench.py --steps 2 --warmup 1 --cpu-sample 2000 > gpurun_out/bench_$lib.json 2> gpurun_out/bench_$lib.err
  echo "== $lib rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$lib.json").read().strip().splitlines()[-1])
    print("ms/step %.2f  value %.3e  frac %.3f  launches %d avg_launch_ms %.3f" % (d["ms_per_step"], d["value"], d["roofline"]["frac"], d["roofline"]["launches"], d["roofline"]["avg_launch_ms"]))
except Exception as e:
    print("failed", e); print(open("gpurun_out/bench_$lib.err").read()[-2000:])
PY
done
