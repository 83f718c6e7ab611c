"""Developer probe: the cfg4 list (10^6-ray 2D sweep, tspan 5e10, reltol 1e-7) cut off after 2000 attempts per ray."""
import sys

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

rt = R.RayTracer("scratch")
peak = rt.measure_fp64_peak()
rt.close()
rt = R.RayTracer("lat2d")
rt.set_solver(t1=5.0e10, reltol=1e-7, maxiters=2000, r_esc=10 * 6.3712e6)
L = sweeps.sweep_2d_long()
for rep in range(2):
    res = rt.solve(L)
nfe = int(res.nfe.astype(np.int64).sum())
ms = res.stats.integrate_ms
u, c = np.unique(res.status, return_counts=True)
print(f"cfg4 (maxiters 2000): n={L.shape[1]} ms={ms:.1f} rhs/s={nfe / ms / 1e6:.3f} G  frac={nfe * 2307 / (ms * 1e-3) / peak:.3f}",
      {R.STATUS_NAMES[int(a)]: int(b) for a, b in zip(u, c)}, "launches", res.stats.kernel_launches)
