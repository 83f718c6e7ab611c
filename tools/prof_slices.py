"""Developer probe: one single-chain cfg3 solve with fixed 4-attempt slices (run under ncu --metrics for a per-slice picture)."""
import sys

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

rt = R.RayTracer("ray3d")
rt.set_solver(r_esc=10 * 6.3712e6, single_chain=1, slice_attempts=4)
res = rt.solve(sweeps.sweep_3d())
print(res.stats.integrate_ms, res.stats.slices)
