#!/bin/bash
# round 2, call 5: parity suite after the product / developer split, records for cfg2 and cfg5 (one GPU)
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c5_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2c5_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c5_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2c5_smoke.log
python bench.py --workload cfg2 --steps 2 --warmup 1 > gpurun_out/r2c5_bench_cfg2.json 2> gpurun_out/r2c5_bench_cfg2.err; echo "cfg2 rc=$?"
python bench.py --workload cfg5 --steps 3 --warmup 3 > gpurun_out/r2c5_bench_cfg5_1gpu.json 2> gpurun_out/r2c5_bench_cfg5_1gpu.err; echo "cfg5 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2c5_bench_cfg2.json","gpurun_out/r2c5_bench_cfg5_1gpu.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "ms/step %.2f value %.4e frac %.4f e2e %.2f"%(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["e2e"]["ms_per_step"]))
    except Exception as e: print(f, "failed", e)
PY
