"""Developer probe: fallback census of cfg3 with the cooperative mode off / on."""
import ctypes as C
import sys

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

lib = R.api._lib()
L = sweeps.sweep_3d()
for thr in (1, 0):
    rt = R.RayTracer("ray3d")
    rt.set_solver(r_esc=10 * 6.3712e6, coop_threshold=thr)
    fb = (C.c_ulonglong * 8)()
    lib.rt_debug_fallback_counts(rt._ctx, fb, 1)
    res = rt.solve(L)
    lib.rt_debug_fallback_counts(rt._ctx, fb, 0)
    print("coop_threshold", thr, list(fb), res.stats.integrate_ms)
    rt.close()
