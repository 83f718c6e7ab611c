"""Developer probe: small 2D / 3D solves (both lane mappings, trajectories, post-processing) for compute-sanitizer."""
import sys

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

RE = 6.3712e6
rt = R.RayTracer("ray3d")
for coop in (-1, 10 ** 9):
    rt.set_solver(maxiters=120, r_esc=10 * RE, coop_threshold=coop, save_mode=1, max_samples=64)
    res = rt.solve(sweeps.sweep_3d(4, 3, 8), derived=True)
    print("3d", coop, np.unique(res.status, return_counts=True), res.traj.shape)
rt.close()
rt = R.RayTracer("lat2d")
for coop in (-1, 10 ** 9):
    rt.set_solver(maxiters=120, coop_threshold=coop, save_mode=1, max_samples=64)
    res = rt.solve(sweeps.sweep_2d(6, 8), derived=True)
    print("2d", coop, np.unique(res.status, return_counts=True), res.traj.shape)
print(rt.plasmasphere(np.linspace(RE, 5 * RE, 100), np.zeros(100))["ne_total"][:2])
print(rt.meridian_maps(np.linspace(-2 * RE, 2 * RE, 9), np.linspace(-2 * RE, 2 * RE, 7))["L"].shape)
rt.close()
