"""What the FP64 pipe sustains for instruction streams shaped like real code (rt_probe_pipe, developer build):
fractions of the nominal rate at 3, 4, 8 and 12 warps per scheduler.  Run plain, then under
ncu --metrics sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.per_cycle_active -k regex:rt_pipe_kernel."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytrace_b200 as R

L = R.api._dev_lib()
names = ["DFMA reuse", "DFMA 3 regs", "DFMA/DMUL/DADD", "+1 int/op", "+FP32 cmp+sel/op"]
for w in (3, 4, 8, 12):
    res = (C.c_double * 5)()
    assert L.rt_probe_pipe(0, w, C.c_double(1950.0), res) == 0
    print(f"warps/scheduler {w:2d}: " + "  ".join(f"{n} {v:.3f}" for n, v in zip(names, res)), flush=True)
