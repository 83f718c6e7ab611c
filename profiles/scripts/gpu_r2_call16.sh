#!/bin/bash
# round 2, call 16: final sanity on the committed tree: smoke, parity suite, the driver's two bench commands
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/r2c16_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2c16_smoke.log
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c16_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "^E |FAILED|passed|failed" gpurun_out/r2c16_pytest.log | head
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2c16_reference.json 2> gpurun_out/r2c16_reference.err; echo "reference rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2c16_bench.json 2> gpurun_out/r2c16_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2c16_bench.json').read().strip().splitlines()[-1]); r=json.loads(open('gpurun_out/r2c16_reference.json').read().strip().splitlines()[-1])
print('cfg3: ms/step %.2f value %.4e frac %.4f peak %.2f e2e %.4e; reference %.4e on %d cores; e2e ratio %.1f'%(d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline']['peak'], d['e2e']['value'], r['value'], r['cpu_baseline']['cores'], d['e2e']['value']/r['value']))"
