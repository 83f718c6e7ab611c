#!/bin/bash
# round 2, call 19 (8 GPUs): the strong-scaling curve of cfg5 (fixed 10^7-ray list) with the block-cyclic default, and cfg3 weak at 2 / 4 / 8
mkdir -p gpurun_out
run() { # N workload steps warmup tag
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29800+$1)) bench.py --gpus $1 --workload $2 --steps $3 --warmup $4 > gpurun_out/r2c19_$5.json 2> gpurun_out/r2c19_$5.err; echo "$5 rc=$?"
}
run 2 cfg5 3 2 cfg5_2gpu
run 4 cfg5 3 2 cfg5_4gpu
run 8 cfg5 3 2 cfg5_8gpu
run 2 cfg3 10 3 cfg3_2gpu
run 4 cfg3 10 3 cfg3_4gpu
run 8 cfg3 10 3 cfg3_8gpu
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2c19_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "N", d["n_gpus"], "ms/step %.2f value %.4e frac %.4f e2e_ms %.2f"%(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["e2e"]["ms_per_step"]), d["config"]["partition"], d.get("gather_root_ms"), [round(b["ms_per_step"],2) for b in (d.get("by_rank") or [])])
    except Exception as e: print(f, "failed", e)
PY
