"""Developer probe: time the sweeps on the GPU through the public API and print per-status counts."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

RE = 6.3712e6


def run(name, preset, launch, **kw):
    rt = R.RayTracer(preset, dev=True)  # developer build: environment switches and the exact-path census
    rt.set_solver(**kw)
    for rep in range(2):
        t = time.time()
        res = rt.solve(launch)
        wall = time.time() - t
        st = res.stats
        nfe = int(res.nfe.astype(np.int64).sum())
        print(f"{name} rep{rep}: n={launch.shape[1]} wall={wall:.3f}s total_ms={st.total_ms:.1f} integ_ms={st.integrate_ms:.1f} "
              f"slices={st.slices} nfe={nfe:.3e} rhs/s={nfe / (st.total_ms * 1e-3):.3e} rays/s={launch.shape[1] / (st.total_ms * 1e-3):.3e}",
              flush=True)
    import ctypes as C
    fb = (C.c_ulonglong * 8)()
    R.api._dev_lib().rt_debug_fallback_counts(rt._ctx, fb, 1)
    print("   exact-path fallbacks:", list(fb), "of", nfe * 2, "RHS evaluations (2 reps)", flush=True)
    u, c = np.unique(res.status, return_counts=True)
    print("   status:", {R.STATUS_NAMES[int(a)]: int(b) for a, b in zip(u, c)}, "naccept min/med/max",
          res.naccept.min(), int(np.median(res.naccept)), res.naccept.max(), "nreject max", res.nreject.max(), flush=True)
    rt.close()
    return res


if __name__ == "__main__":
    rt = R.RayTracer("lat2d")
    print("DFMA peak TFLOP/s:", rt.measure_fp64_peak() / 1e12, flush=True)
    rt.close()
    which = sys.argv[1:] or ["cfg3", "cfg2s"]
    if "cfg3" in which:
        run("cfg3", "ray3d", sweeps.sweep_3d(), r_esc=10 * RE)
    if "cfg3slices" in which:
        for sl in (2, 4, 8, 16, 32):
            run(f"cfg3 slice={sl}", "ray3d", sweeps.sweep_3d(), r_esc=10 * RE, slice_attempts=sl)
    if "cfg3prof" in which:
        run("cfg3-200k", "ray3d", sweeps.sweep_3d(20, 100, 100), r_esc=10 * RE)
    if "cfg2s" in which:
        run("cfg2-small", "lat2d", sweeps.sweep_2d(100, 50), r_esc=10 * RE)
    if "cfg2" in which:
        run("cfg2", "lat2d", sweeps.sweep_2d(), r_esc=10 * RE)
