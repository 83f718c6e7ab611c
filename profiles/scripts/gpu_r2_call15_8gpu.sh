#!/bin/bash
# round 2, call 15 (8 GPUs): what the gather costs its root under the index-placed and the block-placed layouts
mkdir -p gpurun_out
run() { # N workload steps warmup tag layout
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700+$1)) bench.py --gpus $1 --workload $2 --steps $3 --warmup $4 --layout "$6" > gpurun_out/r2c15_$5.json 2> gpurun_out/r2c15_$5.err; echo "$5 rc=$?"
}
run 8 cfg3 8 3 cfg3_8gpu_rootlight ""
run 8 cfg3 8 3 cfg3_8gpu_bc4096 "blockcyclic:4096"
run 8 cfg3 8 3 cfg3_8gpu_bc1000 "blockcyclic:1000"
run 8 cfg3 8 3 cfg3_8gpu_rotated "rotated"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2c15_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "ms/step %.2f value %.4e e2e_ms %.2f"%(d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"]), d["config"]["partition"], d.get("gather_root_ms"), [round(b["ms_per_step"],2) for b in (d.get("by_rank") or [])])
    except Exception as e: print(f, "failed", e)
PY
