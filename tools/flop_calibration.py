"""Run the batch RHS evaluator on the cfg3 / cfg2 launch points so that ncu can count FP64 instructions per RHS:
ncu --metrics sm__sass_thread_inst_executed_op_dadd_pred_on.sum,sm__sass_thread_inst_executed_op_dmul_pred_on.sum,\
sm__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__thread_inst_executed.sum -k regex:rt_rhs_kernel python tools/flop_calibration.py"""
import sys

import numpy as np

sys.path.insert(0, ".")
import raytrace_b200 as R
from raytrace_b200 import sweeps

L3 = sweeps.sweep_3d(40, 50, 50)  # 1e5 points of the cfg3 grid
rt = R.RayTracer("ray3d")
du = rt.haselgrove(L3[:7], [L3[7]])
print("3D n =", L3.shape[1], "finite:", bool(np.isfinite(du).all()))
rt.close()
L2 = sweeps.sweep_2d(400, 250)
rt = R.RayTracer("lat2d")
du = rt.haselgrove(L2[:4])
print("2D n =", L2.shape[1], "finite:", bool(np.isfinite(du).all()))
rt.close()
