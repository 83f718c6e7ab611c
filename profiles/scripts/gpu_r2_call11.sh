#!/bin/bash
# round 2, call 11: the cfg4 record (one full step: 10^6 rays x 10^6 attempts) and a look at the switch-over thresholds
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c11_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "^E |FAILED|passed|failed" gpurun_out/r2c11_pytest.log | head -12
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2c11_bench_cfg3.json 2> gpurun_out/r2c11_bench_cfg3.err; echo "cfg3 rc=$?"
python bench.py --impl reference --gpus 1 --steps 5 --warmup 1 > gpurun_out/r2c11_bench_reference.json 2> gpurun_out/r2c11_bench_reference.err; echo "ref rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2c11_bench_cfg3.json').read().strip().splitlines()[-1]); print('FINAL cfg3: ms/step %.2f value %.4e frac %.4f peak %.2f steady %.4f e2e %.2f'%(d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline']['peak'], d['roofline']['steady_state']['frac'], d['e2e']['ms_per_step']))"
for th in 6000 16000 24000; do
  python bench.py --steps 3 --warmup 2 --cpu-sample 2000 --coop-threshold $th > gpurun_out/r2c11_th$th.json 2>/dev/null
  python -c "
import json; d=json.loads(open('gpurun_out/r2c11_th$th.json').read().strip().splitlines()[-1]); print('coop_threshold $th: ms/step %.2f frac %.4f'%(d['ms_per_step'], d['roofline']['frac']))"
done
bash tools/gpu_env_matrix.sh "-" "RT_COOP_SUM=10000" "RT_COOP_SUM=40000" "RT_COOP_SUM=60000" 2>&1 | tee gpurun_out/r2c11_matrix.log
unset RT_LIB
( time RT_DEEP=1 python -m pytest tests/test_gpu_depth.py -q -k cfg4_options ) > gpurun_out/r2c11_deep_cfg4.log 2>&1; echo "deep cfg4 rc=$?"; tail -4 gpurun_out/r2c11_deep_cfg4.log
python tools/prof_slices.py > gpurun_out/r2c11_plain_slices.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:ILi1ELb1ELb0E -s 6 -c 1 -o gpurun_out/r2c11_coop -f python tools/prof_slices.py > gpurun_out/r2c11_ncu_c.log 2>&1
CMD2="python bench.py --steps 1 --warmup 1 --cpu-sample 2000"
$CMD2 > gpurun_out/r2c11_plain_b.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2c11_launches.csv $CMD2 > gpurun_out/r2c11_ncu_l.log 2>&1
python bench.py --workload cfg4 --steps 1 --warmup 0 --skip-e2e > gpurun_out/r2c11_bench_cfg4.json 2> gpurun_out/r2c11_bench_cfg4.err; echo "cfg4 rc=$?"; tail -2 gpurun_out/r2c11_bench_cfg4.err
tail -c 1500 gpurun_out/r2c11_bench_cfg4.json
