#!/bin/bash
# round 2, call 8: full parity suite and cfg3 line after the vote-skipped L^-4.5 / acos correction and the shared 1/omega^2
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c8_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2c8_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2c8_bench_cfg3.json 2> gpurun_out/r2c8_bench_cfg3.err; echo "cfg3 rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2c8_bench_cfg3.json").read().strip().splitlines()[-1])
print("ms/step %.2f value %.4e frac %.4f steady %.4f e2e %.2f"%(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["roofline"]["steady_state"]["frac"], d["e2e"]["ms_per_step"]))
PY
