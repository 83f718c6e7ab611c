#!/bin/bash
# round 2, call 6 (8 GPUs): the multi-GPU entry point on real devices, weak scaling (cfg3) and strong scaling (cfg5, fixed 10^7 list)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
python -m pytest tests -m gpu -x -q -k "multi_gpu_entry_point or analytic_partials" > gpurun_out/r2c6_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c6_pytest.log
run() { # N workload steps warmup tag
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500+$1)) bench.py --gpus $1 --workload $2 --steps $3 --warmup $4 > gpurun_out/r2c6_$5.json 2> gpurun_out/r2c6_$5.err; echo "$5 rc=$?"
}
run 8 cfg3 5 3 cfg3_8gpu
run 2 cfg5 3 2 cfg5_2gpu
run 4 cfg5 3 2 cfg5_4gpu
run 8 cfg5 3 2 cfg5_8gpu
python tools/bench_abi_multi.py --gpus 8 --workload cfg3 > gpurun_out/r2c6_abi_cfg3_8gpu.json 2> gpurun_out/r2c6_abi_cfg3_8gpu.err; echo "abi cfg3 rc=$?"
python tools/bench_abi_multi.py --gpus 1 --workload cfg3 > gpurun_out/r2c6_abi_cfg3_1gpu.json 2> gpurun_out/r2c6_abi_cfg3_1gpu.err
python tools/bench_abi_multi.py --gpus 8 --workload cfg5 --steps 3 --warmup 2 > gpurun_out/r2c6_abi_cfg5_8gpu.json 2> gpurun_out/r2c6_abi_cfg5_8gpu.err; echo "abi cfg5 rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2c6_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if "roofline" in d: print(f, "N", d["n_gpus"], "ms/step %.2f value %.4e frac %.4f e2e_ms %.2f"%(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["e2e"]["ms_per_step"]), [round(b["ms_per_step"],2) for b in (d.get("by_rank") or [])])
        else: print(f, "N", d["n_gpus"], "ms/step %.2f min %.2f rhs/s %.4e"%(d["ms_per_step"], d["ms_min"], d["rhs_evals_per_s"]), [round(x,1) for x in d["device_solve_ms_last_step"]])
    except Exception as e: print(f, "failed", e)
PY
