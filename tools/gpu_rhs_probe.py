"""Developer probe: branch-free vs exact 3D right-hand side on the inputs that left the fast path (fallback_samples.npy)."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R

s = np.load(sys.argv[1] if len(sys.argv) > 1 else "tools/fallback_samples.npy")
s = np.ascontiguousarray(s[:, :8])
n = len(s)
rt = R.RayTracer("ray3d")
lib = R.api._lib()
df = np.zeros((n, 7)); de = np.zeros((n, 7)); ok = np.zeros(n, dtype=np.int32)
p = lambda a: a.ctypes.data_as(C.c_void_p)
rc = lib.rt_debug_rhs3d(rt._ctx, n, p(s), p(df), p(ok), p(de))
assert rc == 0
nan_e = np.isnan(de).all(1)
same = (df.view(np.uint64) == de.view(np.uint64)).all(1) | (np.isnan(df).all(1) & nan_e)
print("n", n, "ok", ok.sum(), "exact all-NaN", nan_e.sum(), "ok and same", (ok.astype(bool) & same).sum(), "ok but DIFFERENT", (ok.astype(bool) & ~same).sum())
re = 6.3712e6
L = s[:, 0] / (re * np.cos(np.pi / 2 - s[:, 1]) ** 2)
bad = np.where(~ok.astype(bool))[0]
for i in bad[:12]:
    print("not ok: r=%.5g th=%.5f L=%.5g rho=(%.3g,%.3g,%.3g) exact=%s" % (s[i, 0], s[i, 1], L[i], s[i, 3], s[i, 4], s[i, 5], de[i, :2]))
np.save("gpurun_out/probe_ok.npy", ok)
np.save("gpurun_out/probe_de.npy", de)
