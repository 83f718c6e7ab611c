"""GPU parity at the depth the benchmark actually runs: BASELINE.json's configs on their own grids and options, the CUDA
path (through the C ABI) against the CPU oracle, bit for bit.

  cfg1  the reference's own ray (RayTrace_lat.jl:333-338) with its full save-every-step trajectory (5 000+ rows)
  cfg2  2D sweep rays at maxiters = 10^5 (93 % of cfg2's rays end in that class; RayTrace_lat.jl:338 options)
  cfg3  10 000 rays of the 10^6-ray 3D grid at the bench's options (RayTrace_3D.jl:395 defaults), incl. the ~1 000-attempt
        magnetospheric paths
  cfg4  long-duration rays at tspan 5e10 (scratch.jl:390), reltol 1e-7, maxiters 10^6 (RT_DEEP=1: 64 rays, ~5 min of
        oracle time on 16 cores; default 16 rays at full depth + 64 rays at 10^5 attempts)
plus ContinuousCallback's interior safety check on rays that graze the surface inside one step.
"""
import os

import numpy as np
import pytest

import oracle_lib as O
import raytrace_b200 as R
from raytrace_b200 import sweeps
from test_gpu_parity import _check_solve, _same

pytestmark = pytest.mark.gpu

RE = 6.3712e6
NCPU = os.cpu_count() or 4


def _pick(full, count, stride=49999):
    n = full.shape[1]
    while np.gcd(stride, n) != 1:
        stride += 2
    return np.ascontiguousarray(full[:, (np.arange(count, dtype=np.int64) * stride) % n])


def test_cfg1_reference_ray_with_full_trajectory():
    launch = sweeps.cfg1()
    rt = R.RayTracer("lat2d")
    rt.set_solver(save_mode=1, max_samples=8192)  # RayTrace_lat.jl:338 options are the lat2d preset's
    res = rt.solve(launch)
    rt.close()
    E = O.env(O.LAT2D)
    ref = O.solve_ray_2d(E, O.opts(save_mode=1, max_samples=8192), launch[:4, 0], 0.0, cap=8192)
    assert res.status[0] == ref.status == 1 and res.t_end[0] == 5.0e9
    assert res.naccept[0] == ref.naccept and res.nreject[0] == ref.nreject and ref.naccept > 5000
    tr = res.trajectory(0)
    assert tr.shape == ref.traj.shape == (ref.naccept + 1, 9)
    assert _same(tr, ref.traj)  # t, r, lam, chi, u4, mu, dmudpsi, dip, psi at every accepted step
    assert _same(res.u_end[:, 0], ref.u_end) and res.group_delay[0] == ref.u_end[3] - 1000.0


def test_cfg3_sample_of_the_bench_grid_at_bench_options():
    launch = _pick(sweeps.sweep_3d(), 10000)
    rt = R.RayTracer("ray3d")
    rt.set_solver(r_esc=10 * RE)  # ray3d preset = RayTrace_3D.jl:395 defaults; r_esc as in bench.py
    res = rt.solve(launch)
    rt.close()
    ref = O.solve_batch(O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, r_esc=10 * RE), launch, True, NCPU)
    _check_solve(res, ref, 1e-7)
    att = res.naccept + res.nreject
    assert att.max() > 800 and (att > 200).mean() > 0.005  # the long magnetospheric paths are in the sample
    assert _same(res.group_delay, ref["u_end"][6])


def test_cfg2_rays_at_the_full_iteration_cap():
    launch = _pick(sweeps.sweep_2d(), 64)
    rt = R.RayTracer("lat2d")
    rt.set_solver(r_esc=10 * RE)  # maxiters 10^5, reltol 1e-7, dtmax 1e6, dtmin 1e-8, tspan 5e9
    res = rt.solve(launch)
    rt.close()
    ref = O.solve_batch(O.env(O.LAT2D), O.opts(r_esc=10 * RE), launch, False, NCPU)
    _check_solve(res, ref, 1e-7)
    assert (res.status == 4).sum() >= 32 and (res.naccept + res.nreject).max() == 100000


def test_cfg4_options_long_duration_rays():
    deep = os.environ.get("RT_DEEP") == "1"
    full = sweeps.sweep_2d_long()
    kw = dict(reltol=1e-7, abstol=1e-6, dtmax=1e6, dtmin=1e-8, t1=5.0e10, r_esc=10 * RE)
    rt = R.RayTracer("lat2d")
    E = O.env(O.LAT2D)
    for count, maxiters in ((64, 1000000),) if deep else ((64, 100000), (16, 1000000)):
        launch = _pick(full, count)
        rt.set_solver(maxiters=maxiters, **kw)
        res = rt.solve(launch)
        ref = O.solve_batch(E, O.opts(maxiters=maxiters, **kw), launch, False, NCPU)
        _check_solve(res, ref, 1e-7)
        assert (res.naccept + res.nreject).max() == maxiters
    rt.close()


def test_event_inside_a_step_grazing_rays():
    """Rays launched almost horizontally just above the surface dip through r = re and come back within one step at
    the 3D default tolerance; DiffEq's interior check (interp_points = 10) terminates them.  The kernel evaluates the
    eight interior points only when a bound on the interpolant says a crossing is possible; the oracle always does."""
    chi = np.deg2rad(np.linspace(60, 120, 241))
    rows = []
    for alt in (2.0e3, 1.0e4):
        for c in chi:
            rows.append([RE + alt, np.pi / 3, 0.3, np.cos(c), np.sin(c) * np.cos(0.1), np.sin(c) * np.sin(0.1), 0.0, 1000.0, 0.0])
    launch = np.ascontiguousarray(np.array(rows).T)
    E3 = O.env(O.RAY3D)
    okw = dict(reltol=1e-3, dtmax=0, dtmin=0, maxiters=3000, r_esc=10 * RE)
    O.interior_events(reset=True)
    ref = O.solve_batch(E3, O.opts(**okw), launch, True, 1)
    assert O.interior_events(reset=True) >= 2  # (4 to 13 of the 482 launches, depending on last-bit details)
    rt = R.RayTracer("ray3d")
    for coop in (-1, 10 ** 9):
        rt.set_solver(maxiters=3000, r_esc=10 * RE, coop_threshold=coop)
        _check_solve(rt.solve(launch), ref, 1e-7)
    rt.close()
    rows = [[RE + 1.0e4, 0.8726646259971648, -2.2689280275926285, 2000.0, 0.0],
            [RE + 5.0e4, 0.6108652381980153, -2.111848394913139, 2000.0, 0.0],
            [RE + 5.0e4, 0.8726646259971648, -0.29670597283903605, 2000.0, 0.0]]
    launch = np.ascontiguousarray(np.array(rows).T)
    o2 = dict(reltol=1e-3, abstol=1e-6, dtmax=0, dtmin=0, maxiters=3000, t1=5e9, r_esc=10 * RE)
    ref = O.solve_batch(O.env(O.LAT2D), O.opts(**o2), launch, False, 1)
    assert O.interior_events(reset=True) >= 1
    rt = R.RayTracer("lat2d")
    rt.set_solver(**o2)
    _check_solve(rt.solve(launch), ref, 1e-7)
    rt.close()


def test_nan_policy_matches_oracle():
    """nan_policy = 1 (OrdinaryDiffEq: a non-finite error estimate ends the ray with DtNaN) against the oracle on the
    cfg3 grid, where 2-3 % of the rays meet a NaN trial stage."""
    launch = _pick(sweeps.sweep_3d(), 4000)
    rt = R.RayTracer("ray3d")
    rt.set_solver(r_esc=10 * RE, nan_policy=1)
    res = rt.solve(launch)
    rt.close()
    ref = O.solve_batch(O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, r_esc=10 * RE, nan_policy=1), launch, True, NCPU)
    _check_solve(res, ref, 1e-7)
    assert 0.005 < (res.status == 7).mean() < 0.1
