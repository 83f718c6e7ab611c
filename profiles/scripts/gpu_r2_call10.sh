#!/bin/bash
# round 2, call 10: final parity suite, the driver's bench command (both arms), ncu captures of the round-2 kernel
# (full-list launch and die-off launch; the 3-lanes-per-ray launch and the launch list are in call 11: 64 MiB limit)
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c10_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "^E |FAILED|passed|failed" gpurun_out/r2c10_pytest.log | head -12
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2c10_bench_cfg3.json 2> gpurun_out/r2c10_bench_cfg3.err; echo "cfg3 rc=$?"
python bench.py --impl reference --gpus 1 --steps 5 --warmup 1 > gpurun_out/r2c10_bench_reference.json 2> gpurun_out/r2c10_bench_reference.err; echo "ref rc=$?"
python tools/prof_slices.py > gpurun_out/r2c10_plain_slices.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0ELb0E -s 2 -c 1 -o gpurun_out/r2c10_steady -f python tools/prof_slices.py > gpurun_out/r2c10_ncu_a.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb0ELb0E -s 9 -c 1 -o gpurun_out/r2c10_dieoff -f python tools/prof_slices.py > gpurun_out/r2c10_ncu_b.log 2>&1
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2c10_bench_cfg3.json").read().strip().splitlines()[-1])
print("ms/step %.2f value %.4e frac %.4f peak %.2f steady %.4f e2e %.2f"%(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["roofline"]["peak"], d["roofline"]["steady_state"]["frac"], d["e2e"]["ms_per_step"]), d["clocks"])
PY
