#!/bin/bash
# round 2, call 4: parity suite (staged trajectories, multi entry point), cfg3 line, save workload, store efficiency under ncu
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2c4_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2c4_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2c4_bench_cfg3.json 2> gpurun_out/r2c4_bench_cfg3.err; echo "cfg3 rc=$?"
python bench.py --workload cfg2save --steps 3 --warmup 3 > gpurun_out/r2c4_bench_cfg2save.json 2> gpurun_out/r2c4_bench_cfg2save.err; echo "cfg2save rc=$?"; tail -3 gpurun_out/r2c4_bench_cfg2save.err
python tools/prof_save.py > gpurun_out/r2c4_prof_save_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_write.sum,dram__bytes_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum,l1tex__t_requests_pipe_lsu_mem_global_op_st.sum,smsp__inst_executed_op_global_st.sum,smsp__sass_average_data_bytes_per_sector_mem_global_op_st.pct,lts__t_sectors_op_write.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none --kernel-name-base mangled -k regex:rt_integrate_kernelILi0ELb0ELb1E -s 4 -c 6 --csv --log-file gpurun_out/r2c4_save_ncu.csv python tools/prof_save.py > gpurun_out/r2c4_save_ncu.log 2>&1
cat gpurun_out/r2c4_prof_save_plain.log
python - <<'PY'
import json
for f in ("gpurun_out/r2c4_bench_cfg3.json","gpurun_out/r2c4_bench_cfg2save.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "ms/step %.2f frac %.4f e2e %.2f"%(d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["ms_per_step"]), d.get("trajectory"))
    except Exception as e: print(f, "failed", e)
PY
