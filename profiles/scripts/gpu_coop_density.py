"""Experiment: per-attempt latency of the cooperative kernel as a function of how many long rays share the GPU."""
import sys
import numpy as np
sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps
RE = 6.3712e6
full = sweeps.sweep_3d()
rt = R.RayTracer("ray3d"); rt.set_solver(r_esc=10 * RE)
res = rt.solve(full)
long_idx = np.nonzero(res.naccept + res.nreject > 1000)[0]
print("long rays:", long_idx.size, flush=True)
base = full[:, long_idx]
for mult in (1, 2, 3):
    launch = np.ascontiguousarray(np.tile(base, (1, mult)))
    for coop in (10**9, -1):
        rt.set_solver(r_esc=10 * RE, coop_threshold=coop, single_chain=1)
        rt.solve(launch)
        r2 = rt.solve(launch)
        att = int((r2.naccept + r2.nreject).max())
        print(f"rays={launch.shape[1]:6d} mode={'coop3' if coop > 0 else 'bulk '} total_ms={r2.stats.total_ms:7.2f} us/attempt={1e3 * r2.stats.total_ms / att:6.2f} "
              f"G rhs/s={r2.nfe.astype(np.int64).sum() / r2.stats.total_ms / 1e6:.3f}", flush=True)
