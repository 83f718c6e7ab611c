"""rt_solve_3d_device_static against rt_solve_3d_device on short cfg3 launch lists (inputs resident in HBM): wall time of
the synchronous call and the library's CUDA-event time, median of `reps` calls each, plus a bit-equality check.

    python tools/bench_static.py [reps] > gpurun_out/static_vs_device.json
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytrace_b200 as R  # noqa: E402
from raytrace_b200 import sweeps  # noqa: E402

RE = 6.3712e6


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 15
    dev = torch.device("cuda", 0)
    full = sweeps.sweep_3d()
    nfull = full.shape[1]
    stride = 49999
    while np.gcd(stride, nfull) != 1:
        stride += 2
    rt = R.RayTracer("ray3d", device=0)
    rt.set_solver(r_esc=10 * RE)
    rows = []
    for n in (256, 1024, 4096, 16384):
        launch = torch.from_numpy(np.ascontiguousarray(full[:, (np.arange(n, dtype=np.int64) * stride) % nfull])).to(dev)
        cols = [launch[k].data_ptr() for k in range(9)]
        rec = {}
        outs = {}
        for name, fn in (("device", rt.solve_device), ("static", rt.solve_device_static)):
            d = torch.zeros((9, n), dtype=torch.float64, device=dev)
            i = torch.zeros((5, n), dtype=torch.int32, device=dev)
            ptr = {"u_end": d[:7].data_ptr(), "t_end": d[7].data_ptr(), "group_delay": d[8].data_ptr(), "status": i[0].data_ptr(),
                   "naccept": i[1].data_ptr(), "nreject": i[2].data_ptr(), "nfe": i[3].data_ptr(), "nreflect": i[4].data_ptr()}
            for _ in range(3):
                fn(n, cols, ptr)
            wall, evt, launches = [], [], 0
            for _ in range(reps):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                st = fn(n, cols, ptr)
                wall.append((time.perf_counter() - t0) * 1e3)
                evt.append(st.total_ms)
                launches = st.kernel_launches
            outs[name] = (d.cpu().numpy().view(np.int64), i.cpu().numpy())
            rec[name] = {"wall_ms": float(np.median(wall)), "event_ms": float(np.median(evt)), "kernel_launches": int(launches)}
        rec["n"] = n
        rec["nfe_sum"] = int(outs["device"][1][3].sum())
        rec["bit_equal"] = bool(np.array_equal(outs["device"][0], outs["static"][0]) and np.array_equal(outs["device"][1], outs["static"][1]))
        rows.append(rec)
    print(json.dumps({"workload": "cfg3 rays (strided sample), bench options", "reps": reps, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
