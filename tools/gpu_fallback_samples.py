"""Developer probe: which inputs still leave the branch-free 3D right-hand side (saved to gpurun_out/)."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

rt = R.RayTracer("ray3d")
rt.set_solver(r_esc=10 * 6.3712e6)
lib = R.api._lib()
fb = (C.c_ulonglong * 8)()
lib.rt_debug_fallback_counts(rt._ctx, fb, 1)
res = rt.solve(sweeps.sweep_3d())
lib.rt_debug_fallback_counts(rt._ctx, fb, 0)
n = min(int(fb[0]), 8192)
buf = np.zeros((8192, 8))
lib.rt_debug_fallback_samples(rt._ctx, buf.ctypes.data_as(C.c_void_p), 8192)
np.save("gpurun_out/fallback_samples.npy", buf[:n])
print(list(fb), n)
