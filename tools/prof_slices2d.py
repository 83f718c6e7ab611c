"""Developer probe: the cfg2 sweep cut off after 256 attempts, one chain, fixed 16-attempt slices (for ncu)."""
import sys

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

rt = R.RayTracer("lat2d")
rt.set_solver(maxiters=256, single_chain=1, slice_attempts=16, r_esc=10 * 6.3712e6)
res = rt.solve(sweeps.sweep_2d())
print(res.stats.integrate_ms, res.stats.slices, int(res.nfe.astype("int64").sum()) / res.stats.integrate_ms / 1e6, "G RHS/s")
