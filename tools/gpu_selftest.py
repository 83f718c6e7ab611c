"""GPU probe: bit-equality of the branch-free primitives and FP64 dependent-chain latencies."""
import ctypes as C
import sys

sys.path.insert(0, ".")
from raytrace_b200 import api

L = api._lib()
names = ["div", "div_y", "div_c", "sqrt", "exp", "exp10", "sincos", "acos", "atan", "pow_m4p5"]
cnt = (C.c_ulonglong * 30)()
rc = L.rt_selftest_fastmath(0, C.c_longlong(int(sys.argv[1]) if len(sys.argv) > 1 else 2000), C.c_ulonglong(12345), cnt, 30)
print("selftest rc", rc)
bad = 0
for k, nm in enumerate(names):
    print(f"{nm:9s} compared={cnt[3*k]:12d} mismatches={cnt[3*k+1]:8d} flagged={cnt[3*k+2]:10d}")
    bad += cnt[3 * k + 1]
print("TOTAL MISMATCHES", bad)
res = (C.c_double * 6)()
for w in (1, 2, 4, 8, 16):
    L.rt_probe_latency(0, w, res)
    print(f"warps/SM={w:2d} cycles per op: DFMA {res[0]:.2f} DMUL {res[1]:.2f} DADD {res[2]:.2f} RCP64H {res[3]:.2f} | 2-chain {res[4]:.2f} 4-chain {res[5]:.2f}")
