#!/bin/bash
# round 2, call 18: the new early-copy default on the product library: its test and the driver's bench line (e2e is compared bit for bit inside bench.py)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "early_result or multi_gpu_entry" > gpurun_out/r2c18_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2c18_pytest.log
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2c18_bench.json 2> gpurun_out/r2c18_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2c18_bench.err
python -c "
import json
d=json.loads(open('gpurun_out/r2c18_bench.json').read().strip().splitlines()[-1])
print('cfg3: ms/step %.2f value %.4e frac %.4f e2e ms %.2f e2e value %.4e'%(d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['e2e']['value']))"
