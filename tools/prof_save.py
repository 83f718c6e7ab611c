"""Developer probe: a 2D sweep with save-every-step trajectory output (20 000 rays x up to 2 048 rows, 64-attempt slices),
for ncu: store efficiency of the staged trajectory flush in rt_integrate_kernel<0, false, true>."""
import sys

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

rt = R.RayTracer("lat2d")
rt.set_solver(r_esc=10 * 6.3712e6, maxiters=2046, save_mode=1, max_samples=2048, single_chain=1, slice_attempts=64)
res = rt.solve(sweeps.sweep_2d(200, 100), with_trajectories=False)
print(res.stats.integrate_ms, res.stats.slices, int(res.naccept.sum()))
