"""rt_solve_3d_device_static: the solve for callers that cannot allocate device memory (no counterpart in the reference,
include/raytrace_b200.h).  Its working set sits in a file-scope device
array of the library; the call must give the allocating entry point's results bit for bit (the tolerance of every
integer / terminal-record comparison in this suite: none), on consecutive calls that share the array.

Collected last on purpose.  Passes alone in a fresh process and after the others.
"""
import os
import re

import numpy as np
import pytest

import raytrace_b200 as R
from raytrace_b200 import sweeps

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RE = 6.3712e6


def _function_body(src, head):
    i = src.index(head)
    j = src.index("{", i)
    depth, k = 1, j + 1
    while depth:
        depth += {"{": 1, "}": -1}.get(src[k], 0)
        k += 1
    return src[j:k]


def test_entry_point_source_has_no_allocation_and_resolves_the_symbol():
    """Static: the scratch is a file-scope aligned device array; its address comes from a checked cudaGetSymbolAddress in
    the same file; neither the entry point nor its set-up helper allocates or frees."""
    src = open(os.path.join(ROOT, "ray-trace_b200", "csrc", "rt_kernels.cu")).read()
    hdr = open(os.path.join(ROOT, "include", "raytrace_b200.h")).read()
    assert re.search(r"^static __device__ __align__\(256\) unsigned char rt_static_scratch\[RT_STATIC_SCRATCH_BYTES\];", src, flags=re.M)
    cap = int(re.search(r"#define RT_STATIC_MAX_RAYS (\d+)", hdr).group(1))
    assert cap == R.api.STATIC_MAX_RAYS and 256 + 100 * cap < 4 << 20  # modest: under 4 MB per device
    ready = _function_body(src, "static int static_scratch_ready(")
    assert "cudaSetDevice(ctx->device)" in ready and "cudaGetSymbolAddress(&p, rt_static_scratch)" in ready
    assert "e != cudaSuccess" in ready
    body = _function_body(src, 'extern "C" int rt_solve_3d_device_static(')
    for text in (ready, body):
        for word in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "ensure_work", "ensure_stage", "std::vector", "cub::", "new "):
            assert word not in text, word
    assert "ctx->static_scratch" in body and "rt_static_scratch" not in body  # host code never passes the stand-in itself
    assert "no counterpart in the reference" in hdr.lower().replace("\n * ", " ")


def _device_result(torch, n, dev, fill):
    d = torch.full((9, n), float(fill), dtype=torch.float64, device=dev)
    i = torch.full((5, n), int(fill), dtype=torch.int32, device=dev)
    views = {"u_end": d[:7], "t_end": d[7], "group_delay": d[8], "status": i[0], "naccept": i[1], "nreject": i[2],
             "nfe": i[3], "nreflect": i[4]}
    return d, i, {k: v.data_ptr() for k, v in views.items()}


def _bits(d, i):
    return d.cpu().numpy().view(np.int64).copy(), i.cpu().numpy().copy()


@pytest.mark.gpu
def test_static_scratch_solve_matches_the_allocating_entry_point_twice_in_a_row():
    import torch

    dev = torch.device("cuda", 0)
    full = sweeps.sweep_3d()
    nfull = full.shape[1]
    n = 12000  # above the 3-lanes-per-ray threshold at first: both kernel variants run
    assert n <= R.api.STATIC_MAX_RAYS
    stride = 49999
    while np.gcd(stride, nfull) != 1:
        stride += 2
    launch = np.ascontiguousarray(full[:, (np.arange(n, dtype=np.int64) * stride) % nfull])
    launch_d = torch.from_numpy(launch).to(dev)
    cols = [launch_d[k].data_ptr() for k in range(9)]
    rt = R.RayTracer("ray3d", device=0)
    rt.set_solver(r_esc=10 * RE)  # the bench's options (RayTrace_3D.jl:395 defaults)
    out = [_device_result(torch, n, dev, fill) for fill in (-7, -9, -11)]
    torch.cuda.synchronize()

    st1 = rt.solve_device_static(n, cols, out[0][2])           # first call: set-up (symbol address) + solve
    free_before = torch.cuda.mem_get_info()[0]
    st2 = rt.solve_device_static(n, cols, out[1][2])           # second call in a row: shares the array with the first
    free_after = torch.cuda.mem_get_info()[0]
    assert free_after >= free_before, (free_before, free_after)  # nothing was allocated on the device by the call
    rt.solve_device(n, cols, out[2][2])                        # the existing entry point
    torch.cuda.synchronize()
    a, b, ref = (_bits(o[0], o[1]) for o in out)
    for got in (a, b):
        assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])  # every terminal record, bit for bit
    status = ref[1][0]
    assert (status == 2).mean() > 0.5 and status.min() >= 1  # real work: rays come down to the ground, none left running
    att = ref[1][1] + ref[1][2]
    assert att.max() > 400 and st1.slices > 8 and st2.slices == st1.slices

    # a shorter list through a SECOND context of the same device, after the array holds the long list's state
    rt2 = R.RayTracer("ray3d", device=0)
    rt2.set_solver(r_esc=10 * RE)
    m = 777
    sub_d = launch_d[:, 5000:5000 + m].contiguous()
    cols_m = [sub_d[k].data_ptr() for k in range(9)]
    s_out, r_out = _device_result(torch, m, dev, -3), _device_result(torch, m, dev, -5)
    rt2.solve_device_static(m, cols_m, s_out[2])
    rt.solve_device(m, cols_m, r_out[2])
    torch.cuda.synchronize()
    s, r = _bits(s_out[0], s_out[1]), _bits(r_out[0], r_out[1])
    assert np.array_equal(s[0], r[0]) and np.array_equal(s[1], r[1])
    assert np.array_equal(s[0], ref[0][:, 5000:5000 + m])  # and the same rays as in the long list

    # limits are refused with a clear message, nothing is launched
    rt.solve_device_static(0, cols, out[0][2])  # empty list: accepted like the other entry points
    big = R.api.STATIC_MAX_RAYS + 1
    with pytest.raises(RuntimeError, match="RT_STATIC_MAX_RAYS = 16384"):
        rt.solve_device_static(big, cols, out[0][2])
    rt.set_solver(save_mode=1, max_samples=64)
    with pytest.raises(RuntimeError, match="save_mode must be 0"):
        rt.solve_device_static(n, cols, out[0][2])
    rt2d = R.RayTracer("lat2d", device=0)
    with pytest.raises(RuntimeError, match="does not match the 3D call"):
        rt2d.solve_device_static(n, cols, out[0][2])
    for t in (rt, rt2, rt2d):
        t.close()
