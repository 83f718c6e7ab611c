"""Developer probe: attempts per ray of cfg3 next to the launch parameters and dr/dt at launch (saved to gpurun_out/)."""
import sys

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

L = sweeps.sweep_3d()
rt = R.RayTracer("ray3d")
rt.set_solver(r_esc=10 * 6.3712e6)
res = rt.solve(L)
u0 = np.ascontiguousarray(L[:7])
du = rt.haselgrove(u0, [L[7]])
att = (res.naccept + res.nreject).astype(np.int32)
np.savez_compressed("gpurun_out/attempts_cfg3.npz", att=att, drdt=du[0], dthdt=du[1], nrefl=res.nreflect, t_end=res.t_end)
print(att.max(), np.percentile(att, [50, 90, 99, 99.9]))
