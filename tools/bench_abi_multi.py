"""Single-process multi-GPU through the C ABI (rt_solve_3d_multi): what a Julia session calling the library once gets.

    python tools/bench_abi_multi.py [--gpus N] [--workload cfg3|cfg5] [--steps K] [--warmup W]

cfg3: 10^6 rays per GPU (the sweep refined along longitude, as bench.py does: weak scaling); cfg5: the fixed 10^7-ray list
(strong scaling).  Host arrays are pinned; the timed region is the whole call (H2D of the launch columns, solve, D2H of
the terminal records into launch order), wall clock around rt_solve_3d_multi.  Prints one JSON line."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import raytrace_b200 as R
from raytrace_b200 import sweeps

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=0)
ap.add_argument("--workload", default="cfg3")
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--block", type=int, default=0)
a = ap.parse_args()
ngpu = a.gpus or torch.cuda.device_count()
full = sweeps.sweep_3d(100, 100 * ngpu, 100) if a.workload == "cfg3" else sweeps.sweep_3d(100, 1000, 100)
n = full.shape[1]
launch = torch.from_numpy(full).pin_memory().numpy()
m = R.MultiRayTracer("ray3d", devices=list(range(ngpu)), block=a.block)
m.set_solver(r_esc=10 * sweeps.RE)
res = R.api.SolveResult(n, 7)
for name in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
    setattr(res, name, torch.from_numpy(getattr(res, name)).pin_memory().numpy())
for _ in range(a.warmup):
    m.solve(launch, res)
t = []
for _ in range(a.steps):
    t0 = time.perf_counter()
    m.solve(launch, res)
    t.append(time.perf_counter() - t0)
nfe = int(res.nfe.astype(np.int64).sum())
per_dev = [m.stats(g).total_ms for g in range(ngpu)]
b, counts = m.partition(n)
ms = 1e3 * float(np.mean(t))
print(json.dumps({"what": "rt_solve_3d_multi, one process, pinned host arrays, wall clock", "workload": a.workload, "n_gpus": ngpu,
                  "rays": n, "block": b, "rays_per_device": counts, "ms_per_step": ms, "ms_min": 1e3 * min(t),
                  "rhs_evals_per_step": nfe, "rhs_evals_per_s": nfe / (ms * 1e-3), "rays_per_s": n / (ms * 1e-3),
                  "device_solve_ms_last_step": per_dev,
                  "status_counts": {R.STATUS_NAMES[int(k)]: int(v) for k, v in zip(*np.unique(res.status, return_counts=True))}}))
m.close()
