"""The Tsit5 restatement on smooth problems with known answers (the reference's scratch cells: harmonic
oscillator / bouncing-ball style event scratch.jl:570-642, Lorenz scratch.jl:516-529).  CPU oracle only."""
import math

import numpy as np

import oracle_lib as O


def test_fixed_step_order_five():
    errs = []
    for nsteps in (20, 40, 80):
        o = O.opts(adaptive=0, dt=2.0 / nsteps, t1=2.0, maxiters=10 ** 6)
        r = O.solve_test_ode(0, o, [1.0, 0.0])
        assert r.status == 1 and r.naccept == nsteps
        errs.append(math.hypot(r.u_end[0] - math.cos(2.0), r.u_end[1] + math.sin(2.0)))
    p1, p2 = math.log2(errs[0] / errs[1]), math.log2(errs[1] / errs[2])
    assert 4.7 < p1 < 5.6 and 4.7 < p2 < 5.6, (errs, p1, p2)


def test_adaptive_meets_tolerance_and_counts_rhs():
    for rtol in (1e-4, 1e-7, 1e-10):
        o = O.opts(reltol=rtol, abstol=rtol, dtmax=0, dtmin=0, t1=10.0)
        r = O.solve_test_ode(0, o, [1.0, 0.0])
        err = math.hypot(r.u_end[0] - math.cos(10.0), r.u_end[1] + math.sin(10.0))
        assert r.status == 1 and err < 50 * rtol
        assert r.nfe == 6 * (r.naccept + r.nreject) + 2  # FSAL + initial-dt probe (SURVEY §8d unit of work)
        assert r.t_end == 10.0


def test_event_root_on_dense_output():
    """u[0] - re changes sign: cos(t) = 0.5 first at t = pi/3 (falling).  The event point is on the side that has
    not crossed and within the interpolant's accuracy of the true root."""
    o = O.opts(reltol=1e-9, abstol=1e-9, dtmax=0, dtmin=0, t1=10.0)
    r = O.solve_test_ode(0, o, [1.0, 0.0], re_evt=0.5, cap=4096)
    assert r.status == 2
    assert abs(r.t_end - math.pi / 3) < 1e-7
    assert r.u_end[0] >= 0.5 and r.u_end[0] - 0.5 < 1e-9
    assert r.traj[-1, 0] == r.t_end and np.array_equal(r.traj[-1, 1:], r.u_end)
    assert np.all(np.diff(r.traj[:, 0]) > 0) and r.traj.shape[0] == r.naccept + 1


def test_event_fires_on_upward_crossing_too():
    o = O.opts(reltol=1e-8, abstol=1e-8, dtmax=0, dtmin=0, t1=10.0)
    r = O.solve_test_ode(0, o, [0.0, 1.0], re_evt=0.5)  # sin(t) rises through 0.5 at pi/6
    assert r.status == 2 and abs(r.t_end - math.pi / 6) < 1e-6 and r.u_end[0] <= 0.5


def test_event_inside_a_step_is_found_by_the_interior_check():
    """ContinuousCallback's default interp_points = 10: a step whose end points are both on the same side but which
    dips through the level in between still terminates (DiffEqBase determine_event_occurance).  -cos(t) is above 0.99
    only for 3.0001 < t < 3.2831, inside the fixed step [2.8, 3.5]."""
    O.interior_events(reset=True)
    o = O.opts(adaptive=0, dt=0.7, t1=7.0, maxiters=100)
    r = O.solve_test_ode(0, o, [-1.0, 0.0], re_evt=0.99)
    assert r.status == 2 and O.interior_events() == 1
    assert abs(r.t_end - math.acos(-0.99)) < 2e-3 and r.u_end[0] <= 0.99  # 4th-order interpolant on a 0.7-long step
    assert r.naccept == 5
    # the same level is never reached from a launch exactly on it with zero initial sign (prev_sign = 0: no event)
    r = O.solve_test_ode(0, O.opts(adaptive=0, dt=0.1, t1=1.0, maxiters=100), [0.99, 0.0], re_evt=0.99)
    assert r.status == 1


def test_controller_deadband_keeps_the_step():
    """step_accept_controller!: 1 <= q <= 1.2 (qsteady_min, qsteady_max) leaves dt unchanged; on a smooth problem at
    steady state consecutive accepted steps therefore repeat the same dt exactly."""
    o = O.opts(reltol=1e-6, abstol=1e-6, dtmax=0, dtmin=0, t1=50.0)
    r = O.solve_test_ode(0, o, [1.0, 0.0], cap=4096)
    dt = np.diff(r.traj[:, 0])
    assert r.status == 1 and np.count_nonzero(dt[1:-1] == dt[:-2]) > 0.3 * dt.size


def test_nan_policy():
    """A non-finite error estimate: nan_policy 0 retries with dt/5 (ends with DtLessThanMin), 1 ends with DtNaN."""
    E = O.env(O.RAY3D)
    u0 = [E.re + 1e6, math.pi / 4, 0.0, 1.0, 1.0, 1.0e-3, 0.0]
    kw = dict(reltol=1e-3, dtmax=0, dtmin=0, dt=1.0e12, t1=5e12)  # a first step far below the centre: NaN stages
    r0 = O.solve_ray_3d(E, O.opts(**kw), u0, 1000.0)
    r1 = O.solve_ray_3d(E, O.opts(nan_policy=1, **kw), u0, 1000.0)
    assert r1.status == 7 and r1.nreject == 1 and r1.naccept == 0
    assert r0.nreject >= 1 and r0.status != 7


def test_maxiters_dtmin_and_escape_classes():
    o = O.opts(reltol=1e-6, dtmax=0, dtmin=0, t1=100.0, maxiters=10)
    r = O.solve_test_ode(1, o, [1.0, 0.0, 0.0])
    assert r.status == 4 and r.naccept + r.nreject == 10
    o = O.opts(reltol=1e-12, abstol=1e-14, dtmax=0, dtmin=0.5, dt=1.0, t1=100.0)  # first attempt rejects, h -> h/5 <= dtmin
    r = O.solve_test_ode(1, o, [1.0, 0.0, 0.0])
    assert r.status == 5
    o = O.opts(reltol=1e-6, dtmax=0, dtmin=0, t1=100.0, r_esc=15.0)
    r = O.solve_test_ode(1, o, [1.0, 0.0, 0.0])
    assert r.status == 3 and r.u_end[0] > 15.0


def test_save_stride_keeps_end_point():
    o = O.opts(reltol=1e-6, abstol=1e-6, dtmax=0, dtmin=0, t1=5.0, save_stride=4)
    r = O.solve_test_ode(0, o, [1.0, 0.0], cap=4096)
    assert r.traj[0, 0] == 0.0 and r.traj[-1, 0] == 5.0
    assert r.traj.shape[0] in (r.naccept // 4 + 1, r.naccept // 4 + 2)


def test_reference_ray_cfg1():
    """config 1: the reference's own ray (RayTrace_lat.jl:333-338 with Tsit5 for CVODE_BDF).  Pins the step counts
    the survey's throw-away probe saw (5 020 accepted, 0 rejected; ours differ by a few steps: different libm)."""
    E = O.env(O.LAT2D)
    r = O.solve_ray_2d(E, O.opts(), [E.re + 1e6, math.pi / 4, 0.0, 1000.0], cap=8192)
    assert r.status == 1 and r.t_end == 5.0e9
    assert 5000 <= r.naccept <= 5040 and r.nreject <= 5
    assert abs(r.u_end[0] / E.re - 2.22) < 0.01 and abs(math.degrees(r.u_end[1]) - 6.05) < 0.2
    assert r.traj.shape == (r.naccept + 1, 9)
    gd = r.u_end[3] - 1000.0  # slot-4 dual use: group delay = u4 - f0 (RayTrace_lat.jl:257,273)
    assert 0.3 < gd < 0.4


def test_trajectory_capacity_keeps_the_final_sample_and_flags_truncation():
    """max_samples too small: rows 0 .. cap-2 are the first regular samples, the last row is the final sample (event
    point / tspan end), and the ray is flagged (no silent loss of the end of a trajectory)."""
    o = O.opts(reltol=1e-9, abstol=1e-9, dtmax=0, dtmin=0, t1=10.0)
    full = O.solve_test_ode(0, o, [1.0, 0.0], re_evt=0.5, cap=4096)
    assert not full.truncated and full.traj.shape[0] == full.naccept + 1
    cut = O.solve_test_ode(0, o, [1.0, 0.0], re_evt=0.5, cap=10)
    assert cut.truncated and cut.traj.shape[0] == 10
    assert np.array_equal(cut.traj[:9], full.traj[:9]) and np.array_equal(cut.traj[9], full.traj[-1])
    assert cut.traj[9, 0] == cut.t_end and cut.status == 2
    fit = O.solve_test_ode(0, o, [1.0, 0.0], re_evt=0.5, cap=full.traj.shape[0])  # exactly fits: not truncated
    assert not fit.truncated and np.array_equal(fit.traj, full.traj)
