"""The FP64 roofline denominator by itself (rt_measure_fp64_peak): run plain, then under
ncu --metrics sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.avg.per_second,... -k regex:dfma_peak
(profiles/r2_dfma_peak_ncu.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytrace_b200 as R

rt = R.RayTracer("ray3d")
for k in range(3):
    print(f"DFMA peak {rt.measure_fp64_peak() / 1e12:.3f} TFLOP/s")
rt.close()
