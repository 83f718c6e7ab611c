"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle, bit for bit.

Tolerances (BASELINE.json north_star): fixed-step 1e-10 relative, adaptive 1e-7 relative, termination class exact.
Both implementations evaluate the same IEEE operations in the same order (tests assert EXACT equality, which
implies the tolerances; the tolerance assertions are kept so that a future relaxation of exactness still has the
contractual bound written down)."""
import numpy as np
import pytest

import oracle_lib as O
import raytrace_b200 as R
from raytrace_b200 import sweeps

pytestmark = pytest.mark.gpu

RE = 6.3712e6
PRESETS_2D = [("lat2d", O.LAT2D), ("raymain", O.RAYMAIN), ("scratch", O.SCRATCH)]


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return np.array_equal(a.view(np.uint64) if a.dtype == np.float64 else a,
                          b.view(np.uint64) if b.dtype == np.float64 else b) or (
        a.dtype == np.float64 and np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)], b[~np.isnan(b)]))


def _points_2d(n, seed=0):
    rng = np.random.default_rng(seed)
    r = RE * rng.uniform(1.05, 7.0, n)
    lam = np.deg2rad(rng.uniform(5, 80, n))
    chi = rng.uniform(-np.pi, np.pi, n)
    f = np.exp(rng.uniform(np.log(300.0), np.log(3e4), n))
    return r, lam, chi, f


@pytest.mark.parametrize("name,pid", PRESETS_2D)
def test_phase_refractive_index_2d_bitexact(name, pid):
    rt = R.RayTracer(name)
    E = O.env(pid)
    r, lam, chi, f = _points_2d(3000, 1)
    mu, dmu = rt.phase_refractive_index(r, lam, chi, f)
    ref = np.array([O.phase_refractive_index_2d(E, *p) for p in zip(r, lam, chi, f)])
    assert _same(mu, ref[:, 0]) and _same(dmu, ref[:, 1])
    rt.close()


@pytest.mark.parametrize("name,pid", PRESETS_2D)
def test_haselgrove_2d_bitexact(name, pid):
    rt = R.RayTracer(name)
    E = O.env(pid)
    r, lam, chi, f = _points_2d(1500, 2)
    u = np.vstack([r, lam, chi, f])
    du = rt.haselgrove(u)
    ref = np.array([O.haselgrove_2d(E, u[:, k]) for k in range(u.shape[1])]).T
    assert _same(du, ref)
    rt.close()


def _points_3d(n, seed=3):
    rng = np.random.default_rng(seed)
    r = RE * rng.uniform(1.05, 7.0, n)
    th = np.deg2rad(rng.uniform(10, 170, n))
    ph = rng.uniform(0, 2 * np.pi, n)
    rho = rng.normal(size=(3, n)) * rng.uniform(0.5, 100, n)
    f = np.exp(rng.uniform(np.log(300.0), np.log(3e4), n))
    return r, th, ph, rho, f


def test_refractive_index_3d_bitexact():
    rt = R.RayTracer("ray3d")
    E = O.env(O.RAY3D)
    r, th, ph, rho, f = _points_3d(3000)
    mu, dmu, psi, b = rt.refractive_index(r, th, ph, rho[0], rho[1], rho[2], f)
    ref = np.array([O.refractive_index_3d(E, r[k], th[k], ph[k], rho[0, k], rho[1, k], rho[2, k], f[k]) for k in range(r.size)])
    assert _same(mu, ref[:, 0]) and _same(dmu, ref[:, 1]) and _same(psi, ref[:, 2]) and _same(b, ref[:, 3:6].T)
    rt.close()


def test_haselgrove_3d_bitexact_including_nan_lanes():
    rt = R.RayTracer("ray3d")
    E = O.env(O.RAY3D)
    r, th, ph, rho, f = _points_3d(1500, 4)
    rho[2, :10] = 0.0  # the reference's own u0 has rho_phi = 0 -> 0/0 (RayTrace_3D.jl:390, :304-310)
    u = np.vstack([r, th, ph, rho, np.zeros_like(r)])
    du = rt.haselgrove(u, [f])
    ref = np.array([O.haselgrove_3d(E, u[:, k], f[k]) for k in range(r.size)]).T
    assert np.isnan(ref[2, :10]).all()
    assert _same(du, ref)
    rt.close()


def _check_solve(res, ref, rtol):
    assert np.array_equal(res.status, ref["status"]), "termination classes differ"
    assert np.array_equal(res.naccept, ref["naccept"]) and np.array_equal(res.nreject, ref["nreject"])
    assert np.array_equal(res.nfe, ref["nfe"]) and np.array_equal(res.nreflect, ref["nreflect"])
    fin = np.isfinite(ref["u_end"])
    assert np.array_equal(fin, np.isfinite(res.u_end))
    scale = np.maximum(np.abs(ref["u_end"][fin]), 1e-300)
    assert np.max(np.abs(res.u_end[fin] - ref["u_end"][fin]) / scale, initial=0.0) <= rtol
    assert _same(res.u_end, ref["u_end"]) and _same(res.t_end, ref["t_end"])


@pytest.mark.parametrize("coop", [-1, 0])
def test_solve_2d_fixed_step_bitexact(coop):
    launch = sweeps.sweep_2d(16, 8)
    rt = R.RayTracer("lat2d")
    rt.set_solver(adaptive=0, dt=2.0e5, t1=6.0e7, maxiters=100000, coop_threshold=coop)
    res = rt.solve(launch)
    ref = O.solve_batch(O.env(O.LAT2D), O.opts(adaptive=0, dt=2.0e5, t1=6.0e7), launch, False, 8)
    _check_solve(res, ref, 1e-10)
    assert np.all(res.naccept[res.status == 1] == 300)
    f0 = launch[3]
    assert _same(res.group_delay, ref["u_end"][3] - f0)
    rt.close()


# coop_threshold: -1 = one thread per ray throughout; 0 = default (these small batches then run entirely with 3 lanes
# per ray); 150 = switch from one mode to the other in the middle of the solve
@pytest.mark.parametrize("coop", [-1, 0, 150])
@pytest.mark.parametrize("name,pid", PRESETS_2D)
def test_solve_2d_adaptive_bitexact(name, pid, coop):
    launch = sweeps.sweep_2d(24, 12)
    if pid != O.LAT2D:
        launch[1] = np.pi / 2 - launch[1]  # colatitude presets
    rt = R.RayTracer(name)
    rt.set_solver(adaptive=1, reltol=1e-7, abstol=1e-6, dtmax=1e6, dtmin=1e-8, t1=4.0e8, maxiters=3000, r_esc=10 * RE,
                  slice_attempts=37, coop_threshold=coop)
    res = rt.solve(launch)
    ref = O.solve_batch(O.env(pid), O.opts(t1=4.0e8, maxiters=3000, r_esc=10 * RE), launch, False, 8)
    _check_solve(res, ref, 1e-7)
    assert res.stats.slices > 1
    rt.close()


@pytest.mark.parametrize("coop", [-1, 0, 150])
def test_solve_3d_adaptive_bitexact_and_statuses(coop):
    launch = sweeps.sweep_3d(8, 4, 10)
    launch[5, :3] = 0.0  # DtNaN rays (rho_phi = 0)
    rt = R.RayTracer("ray3d")
    rt.set_solver(maxiters=5000, r_esc=10 * RE, slice_attempts=16, coop_threshold=coop)
    res = rt.solve(launch)
    ref = O.solve_batch(O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, maxiters=5000, r_esc=10 * RE), launch, True, 8)
    _check_solve(res, ref, 1e-7)
    assert set(np.unique(res.status)) >= {2, 7}
    assert _same(res.group_delay, ref["u_end"][6])
    rt.close()


@pytest.mark.parametrize("coop", [-1, 0])
def test_trajectories_match_oracle_2d(coop):
    launch = sweeps.sweep_2d(4, 3)
    rt = R.RayTracer("lat2d")
    rt.set_solver(t1=3.0e8, save_mode=1, max_samples=512, save_stride=1, coop_threshold=coop)
    res = rt.solve(launch)
    E, o = O.env(O.LAT2D), O.opts(t1=3.0e8, save_mode=1, max_samples=512)
    for k in range(launch.shape[1]):
        ref = O.solve_ray_2d(E, o, launch[:4, k], launch[4, k], cap=512)
        assert _same(res.trajectory(k), ref.traj)
    rt.close()


@pytest.mark.parametrize("coop", [-1, 0])
def test_trajectories_match_oracle_3d_with_stride(coop):
    launch = sweeps.sweep_3d(2, 2, 4)
    rt = R.RayTracer("ray3d")
    rt.set_solver(save_mode=1, max_samples=64, save_stride=3, coop_threshold=coop)
    res = rt.solve(launch)
    E, o = O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, save_mode=1, max_samples=64, save_stride=3)
    for k in range(launch.shape[1]):
        ref = O.solve_ray_3d(E, o, launch[:7, k], launch[7, k], launch[8, k], cap=64)
        assert _same(res.trajectory(k), ref.traj)
    rt.close()


# ---- the branch-free primitives, the scheduler, and full-size properties -----------------------------------------
def test_branch_free_primitives_bit_equal_exact_ones():
    """rt_fast.cuh against IEEE division / sqrt and rt_math.h on pseudo-random operands, on the device."""
    import ctypes as C
    cnt = (C.c_ulonglong * 30)()
    assert R.api._dev_lib().rt_selftest_fastmath(0, C.c_longlong(200), C.c_ulonglong(2024), cnt, 30) == 0
    for k in range(10):
        compared, mism, flagged = cnt[3 * k], cnt[3 * k + 1], cnt[3 * k + 2]
        assert compared > 5_000_000 and mism == 0, (k, compared, mism, flagged)


def test_branch_free_rhs_vouches_only_for_exact_bits_on_wild_states():
    """The trial stages of steps about to be rejected land anywhere (negative radii, L < -1000, poles): wherever
    the branch-free right-hand side says its result is exact it has the exact code's bits — or is NaN where the
    exact result is NaN in all seven components (the known-NaN shortcuts of rtf_haselgrove_3d)."""
    rng = np.random.default_rng(11)
    n = 300_000
    sign = np.where(rng.random(n) < 0.6, -1.0, 1.0)
    r = sign * 10.0 ** rng.uniform(4.0, 11.5, n)
    th = np.where(rng.random(n) < 0.5, rng.uniform(0, np.pi, n), rng.normal(0, 300, n))
    th[: n // 20] = rng.choice([0.0, np.pi, 1e-9, np.pi - 1e-9, 3.1147], n // 20)  # poles
    rho = rng.normal(0, 1, (n, 3)) * 10.0 ** rng.uniform(-2, 5, (n, 1))
    u = np.column_stack([r, th, rng.uniform(-3, 3, n), rho, rng.uniform(0, 1, n)])
    u[:100, 0] = np.nan
    u[100:200, 0] = np.inf
    u[200:300, 5] = 0.0
    f = rng.choice([1000.0, 5000.0, 25000.0], n)
    rt = R.RayTracer("ray3d", dev=True)  # diagnostics entry point: developer build of the library
    df, ok, de = rt.debug_rhs3d(u, f)
    rt.close()
    nan_e = np.isnan(de).all(axis=1)
    same = (df.view(np.uint64) == de.view(np.uint64)).all(axis=1) | (np.isnan(df).all(axis=1) & nan_e)
    assert not (ok & ~same).any(), (int((ok & ~same).sum()), u[ok & ~same][:3])
    # the shortcuts are exercised, and most wild states stay off the exact path
    assert (ok & nan_e).sum() > 20_000 and ok.mean() > 0.6, (int((ok & nan_e).sum()), float(ok.mean()))
    assert ok[:100].all() and nan_e[:100].all()  # NaN radius


def test_trajectory_post_processing_on_the_device():
    """x = r cos(lam), y = r sin(lam), dt series (RayTrace_lat.jl:351-352,369-372) from the saved
    trajectories, against the oracle's sin / cos on the copied samples (bit-exact)."""
    for preset, launch in (("lat2d", sweeps.sweep_2d(3, 4)), ("raymain", sweeps.sweep_2d(2, 3)), ("ray3d", sweeps.sweep_3d(2, 2, 3))):
        rt = R.RayTracer(preset)
        rt.set_solver(maxiters=300, save_mode=1, max_samples=256)
        res = rt.solve(launch, derived=True)
        rt.close()
        tr, d = res.traj, res.traj_derived
        assert d.shape == (tr.shape[0], 4 if preset == "ray3d" else 3) and tr.shape[0] > launch.shape[1] * 20
        r, a = tr[:, 1], tr[:, 2]
        sa, ca = O.math("sin", a), O.math("cos", a)
        if preset == "ray3d":
            sp, cp = O.math("sin", tr[:, 3]), O.math("cos", tr[:, 3])
            want = [(r * sa) * cp, (r * sa) * sp, r * ca]
        elif preset == "lat2d":
            want = [r * ca, r * sa]
        else:
            want = [r * sa, r * ca]
        for k, w in enumerate(want):
            assert _same(d[:, k], w), (preset, k)
        dt = np.zeros(tr.shape[0])
        dt[:-1] = tr[1:, 0] - tr[:-1, 0]
        dt[res.traj_offsets[1:] - 1] = 0.0  # the last sample of each ray
        assert _same(d[:, -1], dt), preset
        k = launch.shape[1] - 1
        assert np.array_equal(res.derived(k), d[res.traj_offsets[k]:res.traj_offsets[k + 1]])


def test_refractive_index_surface_sweep():
    """RayTrace_lat.jl:380-399 as one batch call: mu(psi), dmu/dpsi(psi) at the reference's test point."""
    r, lam, chi, psi = sweeps.index_surface()
    assert psi.size == 6284 and abs(r - np.hypot(9.031e6, 1.537e6)) < 1e-6
    rt = R.RayTracer("lat2d")
    mu, dmu = rt.phase_refractive_index(r, lam, chi, 5000.0)
    rt.close()
    E = O.env(O.LAT2D)
    for k in (0, 1571, 3142, 6283):
        m, d = O.phase_refractive_index_2d(E, r, lam, chi[k], 5000.0)[:2]
        assert _same(np.array([mu[k], dmu[k]]), np.array([m, d])), k
    assert np.isfinite(mu).all() and mu.min() > 1.0


def test_plasmasphere_jl_composition_as_the_density_inside_mu():
    """density_model = 1 puts diffusive_equilibrium(r, CA92(L)) + ionosphere_eq(r)
    (plasmasphere.jl:96-117,171) under the refractive index; the integrate kernel then runs the exact code."""
    r, lam, chi, f = _points_2d(1500, 21)
    env = R.env_preset("lat2d")
    env.density_model = 1
    rt = R.RayTracer("lat2d", env=env)
    E = O.env(O.LAT2D)
    mu0 = np.array([O.phase_refractive_index_2d(E, *p)[0] for p in zip(r[:50], lam[:50], chi[:50], f[:50])])
    E.density_model = 1
    mu, dmu = rt.phase_refractive_index(r, lam, chi, f)
    ref = np.array([O.phase_refractive_index_2d(E, *p) for p in zip(r, lam, chi, f)])
    assert _same(mu, ref[:, 0]) and _same(dmu, ref[:, 1]) and not _same(mu[:50], mu0)
    launch = sweeps.sweep_2d(3, 4)
    rt.set_solver(maxiters=400)
    res = rt.solve(launch)
    rt.close()
    _check_solve(res, O.solve_batch(E, O.opts(maxiters=400), launch, False, 4), 1e-10)

    env3 = R.env_preset("ray3d")
    env3.density_model = 1
    rt = R.RayTracer("ray3d", env=env3)
    E3 = O.env(O.RAY3D)
    E3.density_model = 1
    r, th, ph, rho, f = _points_3d(1000, 22)
    u = np.vstack([r, th, ph, rho, np.zeros_like(r)])
    du = rt.haselgrove(u, [f])
    assert _same(du, np.array([O.haselgrove_3d(E3, u[:, k], f[k]) for k in range(r.size)]).T)
    launch = sweeps.sweep_3d(3, 2, 4)
    for coop in (-1, 10 ** 9):
        rt.set_solver(maxiters=300, r_esc=10 * RE, coop_threshold=coop)
        res = rt.solve(launch)
        _check_solve(res, O.solve_batch(E3, O.opts(reltol=1e-3, dtmax=0, dtmin=0, maxiters=300, r_esc=10 * RE), launch, True, 4), 1e-7)
    rt.close()
    bad = R.env_preset("lat2d")
    bad.density_model = 7
    with pytest.raises(RuntimeError, match="density_model"):
        R.RayTracer("lat2d", env=bad)


def test_meridian_maps_bitexact():
    """plasmasphere.jl:157-175 and RayTrace_3D.jl:504-517 as one batch kernel, against the oracle's loop."""
    x = np.linspace(-4 * RE, 4 * RE, 81)  # includes x = 0 (y/x = +-Inf) and the origin (0/0)
    y = np.linspace(-4 * RE, 4 * RE, 61)
    rt = R.RayTracer("lat2d")
    m = rt.meridian_maps(x, y)
    rt.close()
    ref = O.meridian_maps(O.env(O.LAT2D), x, y)
    for i, k in enumerate(("L", "ne_total", "log10_ne_total", "Bmag")):
        assert m[k].shape == (81, 61) and _same(m[k].ravel(), ref[i]), k
    ix, iy = 60, 30  # x = 2 re, y = 0: L = 2, |B| = B0 / 8 * 4
    assert abs(m["L"][ix, iy] - 2.0) < 1e-12 and abs(m["Bmag"][ix, iy] - O.env(O.LAT2D).B0 / 2.0) < 1e-18
    with pytest.raises(RuntimeError, match="RAY3D"):
        R.RayTracer("ray3d").meridian_maps(x, y)


def test_early_result_copy_into_pinned_arrays_matches_the_plain_copy(monkeypatch):
    """rt_solve_3d with pinned result arrays sends the finished rays' records to the host while the long rays still
    run and patches those in afterwards (run_solve / solve_common); RT_EARLY_MIN=0 enables it for a small list."""
    import ctypes as C
    import torch
    launch = sweeps.sweep_3d(12, 3, 24)
    n = launch.shape[1]
    monkeypatch.setenv("RT_EARLY_MIN", "0")  # read by the developer build only, at its first solve
    rt = R.RayTracer("ray3d", dev=True)
    rt.set_solver(maxiters=600, r_esc=10 * RE, coop_threshold=64)
    plain = rt.solve(launch)  # pageable numpy arrays: one copy at the end
    res = R.api.SolveResult(n, 7)
    keep = []
    for name in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
        t = torch.from_numpy(getattr(res, name)).pin_memory()
        t.fill_(-7)
        keep.append(t)
        setattr(res, name, t.numpy())
    cres = res._c()
    lp = torch.from_numpy(np.ascontiguousarray(launch)).pin_memory()
    cols = [C.c_void_p(lp[k].data_ptr()) for k in range(9)]
    for _ in range(2):
        assert R.api._dev_lib().rt_solve_3d(rt._ctx, C.c_int64(n), *cols, C.byref(cres)) == 0
        for k in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
            assert _same(getattr(res, k), getattr(plain, k)), k
    assert (plain.naccept + plain.nreject).max() == 600  # long rays were still running at the snapshot
    rt.close()


def test_resolvable_finite_difference_steps_are_an_option_off_the_parity_path():
    """fd_step_angle / fd_step_r > 0 replace the files' 1e-11 (under which d(mu)/dr is identically 0
    and d(mu)/d(theta) is rounding noise); oracle and kernels agree bit for bit with the option on as well."""
    env = R.env_preset("ray3d")
    env.fd_step_angle, env.fd_step_r = 1e-6, 10.0
    E = O.env(O.RAY3D)
    E.fd_step_angle, E.fd_step_r = 1e-6, 10.0
    rt = R.RayTracer("ray3d", env=env)
    r, th, ph, rho, f = _points_3d(600, 31)
    u = np.vstack([r, th, ph, rho, np.zeros_like(r)])
    du = rt.haselgrove(u, [f])
    assert _same(du, np.array([O.haselgrove_3d(E, u[:, k], f[k]) for k in range(r.size)]).T)
    du0 = R.RayTracer("ray3d").haselgrove(u, [f])
    assert not _same(du[3], du0[3]) and _same(du[0], du0[0])  # d(rho_r)/dt now has its (1/mu) d(mu)/dr term
    launch = sweeps.normalize_rho(rt, sweeps.sweep_3d(3, 2, 4))  # |rho| = mu at launch
    mu = rt.refractive_index(*launch[:6], launch[7])[0]
    assert np.allclose(np.sqrt((launch[3:6] ** 2).sum(axis=0)), mu, rtol=1e-14)
    rt.set_solver(maxiters=300, r_esc=10 * RE)
    res = rt.solve(launch)
    _check_solve(res, O.solve_batch(E, O.opts(reltol=1e-3, dtmax=0, dtmin=0, maxiters=300, r_esc=10 * RE), launch, True, 4), 1e-7)
    rt.close()
    env2 = R.env_preset("lat2d")
    env2.fd_step_angle, env2.fd_step_r = 1e-6, 10.0
    E2 = O.env(O.LAT2D)
    E2.fd_step_angle, E2.fd_step_r = 1e-6, 10.0
    rt = R.RayTracer("lat2d", env=env2)
    r, lam, chi, f = _points_2d(600, 32)
    u = np.vstack([r, lam, chi, f])
    assert _same(rt.haselgrove(u), np.array([O.haselgrove_2d(E2, u[:, k]) for k in range(u.shape[1])]).T)
    rt.close()


def test_edge_sizes_and_every_termination_class():
    """Empty, single-ray and ragged launch lists (n % 128 != 0, n % 40 != 0), and every status the solver can
    return — Success (t1 reached), Terminated (ground), Escaped, MaxIters, DtLessThanMin — against the oracle."""
    rt3 = R.RayTracer("ray3d")
    full = sweeps.sweep_3d(5, 3, 9)  # 135 rays
    res = rt3.solve(full[:, :0])
    assert res.status.size == 0 and res.u_end.shape == (7, 0)
    E3 = O.env(O.RAY3D)
    cases = [
        dict(maxiters=40),                                   # MaxIters for every ray
        dict(maxiters=400, r_esc=RE + 1.2e6),                # upward rays escape at once, the others reach the ground
        dict(maxiters=400, dtmin=5e6, dtmax=1e9),            # DtLessThanMin after the first rejections
        dict(maxiters=400, t1=2.0e-3),                       # Success: tspan ends first
    ]
    seen = set()
    for kw in cases:
        okw = dict(dict(reltol=1e-3, dtmax=0, dtmin=0, maxiters=400), **kw)
        for n in (1, 129, 135):
            for coop in (-1, 10 ** 9):
                rt3.set_solver(**dict(dict(r_esc=0.0, dtmin=0.0, dtmax=0.0, t1=5.0e9), **kw, coop_threshold=coop))
                res = rt3.solve(full[:, :n])
                _check_solve(res, O.solve_batch(E3, O.opts(**okw), full[:, :n], True, 2), 1e-7)
                seen |= set(int(v) for v in np.unique(res.status))
    names = {R.STATUS_NAMES[v] for v in seen}
    # (DtNaN: the sweep's chi = 0 column has rho_theta = 0, the reference's own 0/0, RayTrace_3D.jl:304-310)
    assert {"Success", "Terminated", "Escaped", "MaxIters", "DtLessThanMin", "DtNaN"} <= names, names
    with pytest.raises(ValueError, match="launch list"):
        rt3.solve(sweeps.sweep_2d(2, 2))
    rt3.close()
    rt2 = R.RayTracer("lat2d")
    launch = sweeps.sweep_2d(7, 19)  # 133 rays
    for n in (1, 133):
        for coop in (-1, 10 ** 9):
            rt2.set_solver(maxiters=150, coop_threshold=coop)
            res = rt2.solve(launch[:, :n])
            _check_solve(res, O.solve_batch(O.env(O.LAT2D), O.opts(maxiters=150), launch[:, :n], False, 2), 1e-10)
    rt2.close()


def test_two_chains_and_one_chain_give_identical_results():
    """Rays that start upward go to the priority chain (scheduling only)."""
    launch = sweeps.sweep_3d(12, 3, 24)
    out = []
    for single in (0, 1):
        rt = R.RayTracer("ray3d")
        rt.set_solver(maxiters=600, r_esc=10 * RE, single_chain=single, coop_threshold=64)
        res = rt.solve(launch)
        assert (res.stats.prio_rays > 0) == (single == 0)
        out.append(res)
        rt.close()
    a, b = out
    for k in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
        assert _same(getattr(a, k), getattr(b, k)), k
    ref = O.solve_batch(O.env(O.RAY3D), O.opts(reltol=1e-3, dtmax=0, dtmin=0, maxiters=600, r_esc=10 * RE), launch, True, 8)
    _check_solve(a, ref, 1e-7)
    assert 4 in set(np.unique(a.status))  # the long rays hit the iteration cap here


def test_cfg3_full_size_properties_and_partition():
    """BASELINE.json configs[2] at full size (10^6 rays): size-independent properties.  Every ray ends on the Earth's
    surface (event root to rounding); counters are consistent; the result does not depend on how the launch list is
    scheduled (one chain / two chains) or partitioned (multi-GPU sharding is a launch-list partition); mu does not
    depend on longitude (RayTrace_3D.jl:54-66), so the 100 longitudes of one (colatitude, chi) cell agree up to the
    step-size controller's phi-dependent error weight."""
    import ctypes as C
    full = sweeps.sweep_3d()
    n = full.shape[1]
    rtd = R.RayTracer("ray3d", dev=True)  # the developer build keeps the census of exact-path calls
    rtd.set_solver(r_esc=10 * RE)
    fb = (C.c_ulonglong * 8)()
    R.api._dev_lib().rt_debug_fallback_counts(rtd._ctx, fb, 1)
    resd = rtd.solve(full)
    R.api._dev_lib().rt_debug_fallback_counts(rtd._ctx, fb, 1)
    rtd.close()
    rt = R.RayTracer("ray3d")
    rt.set_solver(r_esc=10 * RE)
    res = rt.solve(full)
    assert _same(res.u_end, resd.u_end) and _same(res.t_end, resd.t_end)  # product and developer builds agree
    assert set(np.unique(res.status)) == {2}  # every ray of this sweep ends on the Earth
    assert int(res.nfe.astype(np.int64).sum()) == 301897484  # the workload bench.py quotes (the oracle run over the full list gives the same)
    assert np.array_equal(res.nfe, 6 * (res.naccept + res.nreject) + 2)
    assert np.max(np.abs(res.u_end[0] - RE)) < 1e-6 and np.all(res.t_end > 0)  # metres: bisection to 1 ulp in theta
    assert _same(res.group_delay, res.u_end[6])
    assert fb[0] < 1e-3 * res.nfe.astype(np.int64).sum()  # the exact path is the rare path
    cube = lambda a: a.reshape(100, 100, 100)  # (colat, lon, chi)
    short = cube(res.naccept + res.nreject).max(axis=1) < 64  # cells of rays that go straight down
    t = cube(res.t_end)
    spread = (t.max(axis=1) - t.min(axis=1)) / t.mean(axis=1)
    assert short.mean() > 0.9 and np.max(spread[short]) < 5e-2  # solver tolerance is 1e-3
    # scheduling invariance: one chain, different slice length
    rt.set_solver(r_esc=10 * RE, single_chain=1, slice_attempts=7)
    one = rt.solve(full)
    for k in ("u_end", "t_end", "status", "naccept", "nreject", "nreflect"):
        assert _same(getattr(one, k), getattr(res, k)), k
    # partition invariance: the 3rd of 8 contiguous parts
    rt.set_solver(r_esc=10 * RE, single_chain=0, slice_attempts=0)
    lo, hi = sweeps.partition(n, 2, 8)
    part = rt.solve(np.ascontiguousarray(full[:, lo:hi]))
    for k in ("t_end", "status", "naccept", "nfe"):
        assert _same(getattr(part, k), getattr(res, k)[lo:hi]), k
    assert _same(part.u_end, res.u_end[:, lo:hi])
    rt.close()


def test_cfg2_size_2d_sweep_capped_matches_oracle_sample():
    """BASELINE.json configs[1] launch grid at full size (10^5 rays), iteration cap lowered so that the sweep takes
    seconds; a strided sample of it against the oracle, and statuses/counters consistent everywhere."""
    full = sweeps.sweep_2d()
    rt = R.RayTracer("lat2d")
    rt.set_solver(maxiters=400, r_esc=10 * RE)
    res = rt.solve(full)
    assert np.array_equal(res.nfe, 6 * (res.naccept + res.nreject) + 2)
    assert np.all((res.status == 4) == (res.naccept + res.nreject >= 400))
    idx = np.arange(0, full.shape[1], 997)
    ref = O.solve_batch(O.env(O.LAT2D), O.opts(maxiters=400, r_esc=10 * RE), np.ascontiguousarray(full[:, idx]), False, 8)
    assert _same(res.u_end[:, idx], ref["u_end"]) and _same(res.t_end[idx], ref["t_end"])
    assert np.array_equal(res.status[idx], ref["status"]) and np.array_equal(res.nfe[idx], ref["nfe"])
    rt.close()


def test_plasmasphere_jl_batch_bitexact():
    """plasmasphere.jl functions on the GPU against the oracle, all three CA92 segments."""
    rng = np.random.default_rng(7)
    n = 4000
    r = RE * rng.uniform(1.02, 9.5, n)
    lam = np.deg2rad(rng.uniform(-70, 70, n))
    rt = R.RayTracer("lat2d")
    got = rt.plasmasphere(r, lam)
    e = rt.env
    L = O.lib()
    Lref = np.array([L.orc_L_rlam(O._d(a), O._d(b), O._d(RE)) for a, b in zip(r, lam)])
    eq = np.array([L.orc_plasmasphere_eq(O._d(x), O._d(e.Lppi), O._d(e.Lppo), O._d(0), O._d(90), O._d(2)) for x in Lref])
    de = np.array([L.orc_diffusive_equilibrium(O._d(a), O._d(b)) for a, b in zip(r, eq)])
    io = np.array([L.orc_ionosphere_eq(O._d(a), O._d(RE)) for a in r])
    assert {0, 1, 2} <= set(np.digitize(Lref, [e.Lppi, e.Lppo]))
    assert _same(got["L"], Lref) and _same(got["plasmasphere_eq"], eq)
    assert _same(got["diffusive_equilibrium"], de) and _same(got["ionosphere_eq"], io) and _same(got["ne_total"], de + io)
    rt.close()
    rt3 = R.RayTracer("ray3d")
    with pytest.raises(RuntimeError, match="RAY3D"):
        rt3.plasmasphere(r[:4], lam[:4])
    rt3.close()


def test_multi_gpu_entry_point_block_cyclic_partition():
    """rt_solve_*_multi: one process, one context + host thread per device, block-cyclic partition inside the library,
    records in launch order in the caller's arrays.  Two contexts on device 0 stand in for two GPUs (rays are
    independent, so the result must equal the single-context solve bit for bit); ragged sizes, pinned and pageable."""
    import torch
    full = sweeps.sweep_3d(12, 3, 24)
    rt = R.RayTracer("ray3d")
    rt.set_solver(maxiters=600, r_esc=10 * RE)
    one = rt.solve(full)
    rt.close()
    ndev = torch.cuda.device_count()
    for devices, block in (([0, 0], 128), ([0, 0, 0], 200), ([0], 0), (list(range(min(ndev, 8))), 0)):
        m = R.MultiRayTracer("ray3d", devices=devices, block=block)
        m.set_solver(maxiters=600, r_esc=10 * RE)
        for n in (full.shape[1], 777, 1, 0):
            res = R.api.SolveResult(n, 7)
            if n == 777:  # pinned result arrays: the early-copy path of every device
                for name in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
                    setattr(res, name, torch.from_numpy(getattr(res, name)).pin_memory().numpy())
            m.solve(np.ascontiguousarray(full[:, :n]), res)
            for k in ("t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
                assert _same(getattr(res, k), getattr(one, k)[:n]), (devices, n, k)
            assert _same(res.u_end, one.u_end[:, :n])
            b, counts = m.partition(n)
            assert sum(counts) == n and b >= 1
            for g in range(len(devices)):
                idx = sweeps.block_cyclic_indices(n, b, len(devices), g)
                assert idx.size == counts[g]
                if idx.size:
                    assert m.global_index(n, g, idx.size - 1) == idx[-1] and m.global_index(n, g, 0) == idx[0]
        assert m.stats(0).kernel_launches >= 0
        m.close()
    m2 = R.MultiRayTracer("lat2d", devices=[0, 0], block=16)
    m2.set_solver(maxiters=150)
    launch = sweeps.sweep_2d(7, 19)
    res = m2.solve(launch)
    m2.close()
    _check_solve(res, O.solve_batch(O.env(O.LAT2D), O.opts(maxiters=150), launch, False, 4), 1e-10)


@pytest.mark.parametrize("coop", [-1, 10 ** 9])
def test_trajectory_truncation_keeps_final_sample_and_matches_oracle(coop):
    """max_samples smaller than the ray's step count (staged rows, warp flush, reserved final row): the kept rows and
    the truncation flags equal the oracle's; slices of 3 attempts leave partial stages at every launch boundary."""
    launch = sweeps.sweep_3d(3, 2, 7)
    E, cap = O.env(O.RAY3D), 12
    rt = R.RayTracer("ray3d")
    rt.set_solver(save_mode=1, max_samples=cap, save_stride=1, slice_attempts=3, coop_threshold=coop, r_esc=10 * RE)
    res = rt.solve(launch)
    o = O.opts(reltol=1e-3, dtmax=0, dtmin=0, save_mode=1, max_samples=cap, r_esc=10 * RE)
    ntr = 0
    for k in range(launch.shape[1]):
        ref = O.solve_ray_3d(E, o, launch[:7, k], launch[7, k], launch[8, k], cap=cap)
        assert _same(res.trajectory(k), ref.traj), k
        assert bool(res.truncated[k]) == ref.truncated, k
        if ref.truncated:
            ntr += 1
            assert ref.traj.shape[0] == cap and res.trajectory(k)[-1, 0] == res.t_end[k]  # the event row survived
    assert ntr > launch.shape[1] // 2
    rt.close()
    launch = sweeps.sweep_2d(5, 3)
    rt = R.RayTracer("lat2d")
    rt.set_solver(t1=3.0e8, save_mode=1, max_samples=40, save_stride=2, slice_attempts=5, coop_threshold=coop)
    res = rt.solve(launch)
    E2, o2 = O.env(O.LAT2D), O.opts(t1=3.0e8, save_mode=1, max_samples=40, save_stride=2)
    for k in range(launch.shape[1]):
        ref = O.solve_ray_2d(E2, o2, launch[:4, k], launch[4, k], cap=40)
        assert _same(res.trajectory(k), ref.traj) and bool(res.truncated[k]) == ref.truncated, k
    assert res.truncated.any()
    rt.close()


def test_trajectory_pack_on_the_device_and_bad_capacity():
    import ctypes as C
    import torch
    launch = sweeps.sweep_3d(6, 5, 8)
    rt = R.RayTracer("ray3d")
    rt.set_solver(save_mode=1, max_samples=128, r_esc=10 * RE)
    res = rt.solve(launch)
    d_off, d_rows, total = C.c_void_p(), C.c_void_p(), C.c_int64()
    assert R.api._lib().rt_traj_pack_device(rt._ctx, C.byref(d_off), C.byref(d_rows), C.byref(total)) == 0
    assert total.value == res.traj.shape[0] == res.traj_offsets[-1]
    off = np.zeros(launch.shape[1] + 1, np.int64)
    rows = np.zeros((total.value, 8))
    cudart = C.CDLL("libcudart.so.12")
    assert cudart.cudaMemcpy(off.ctypes.data_as(C.c_void_p), d_off, C.c_size_t(off.nbytes), C.c_int(2)) == 0
    assert cudart.cudaMemcpy(rows.ctypes.data_as(C.c_void_p), d_rows, C.c_size_t(rows.nbytes), C.c_int(2)) == 0
    assert np.array_equal(off, res.traj_offsets) and _same(rows, res.traj)
    counts = np.diff(res.traj_offsets)
    assert np.all(counts[res.truncated == 1] == 128) and np.all(res.truncated[counts < 128] == 0)
    with pytest.raises(RuntimeError, match="max_samples >= 2"):
        rt.set_solver(save_mode=1, max_samples=1)
    rt.close()


def test_analytic_partials_and_root_selection_options():
    """SURVEY §8 f4, off the parity path.  deriv_mode = 1: d(mu)/dr and d(mu)/d(theta) by forward-mode differentiation
    agree with central differences at steps FP64 can resolve (fd_step_angle = 1e-6 rad, fd_step_r = 10 m) to 1e-6
    relative, and leave the components that do not use them untouched.  root_mode = 1: the propagating root without
    abs() — equal to the as-written mu wherever mu^2_+ > 0, sqrt(mu^2_-) or NaN elsewhere."""
    import tools_golden as TG
    r, th, ph, rho, f = _points_3d(800, 41)
    u = np.vstack([r, th, ph, rho, np.zeros_like(r)])
    e_fd = R.env_preset("ray3d")
    e_fd.fd_step_angle, e_fd.fd_step_r = 1e-6, 10.0
    e_ad = R.env_preset("ray3d")
    e_ad.deriv_mode = 1
    rt_fd, rt_ad = R.RayTracer("ray3d", env=e_fd), R.RayTracer("ray3d", env=e_ad)
    d_fd, d_ad = rt_fd.haselgrove(u, [f]), rt_ad.haselgrove(u, [f])
    for k in (0, 1, 2, 5, 6):
        assert np.allclose(d_ad[k], d_fd[k], rtol=1e-12, atol=0.0), k  # no r / theta partial in these
    scale = np.abs(d_fd[4]) + np.abs(rho[1] * d_fd[0]) / r
    assert np.max(np.abs(d_ad[4] - d_fd[4]) / scale) < 1e-6                      # (1/mu) d(mu)/d(theta) term
    scale3 = np.abs(d_fd[3]) + np.abs(rho[1] * d_fd[1])
    assert np.max(np.abs(d_ad[3] - d_fd[3]) / scale3) < 1e-6                     # (1/mu) d(mu)/dr term
    d0 = R.RayTracer("ray3d").haselgrove(u, [f])
    assert not np.allclose(d_ad[3], d0[3], rtol=1e-3)                            # as written that term is exactly 0
    launch = sweeps.sweep_3d(3, 2, 6)
    for coop in (-1, 10 ** 9):
        rt_ad.set_solver(maxiters=400, r_esc=10 * RE, coop_threshold=coop)
        rt_fd.set_solver(maxiters=400, r_esc=10 * RE, coop_threshold=coop)
        a, b = rt_ad.solve(launch), rt_fd.solve(launch)
        down = a.status == 2  # rays that reach the ground: most follow the same path to within the solver tolerance
        # (1e-3); a few sit on a reflect / do-not-reflect boundary and differ wholesale, hence the median
        assert np.array_equal(a.status, b.status) and down.sum() > 10
        assert np.median(np.abs(a.t_end[down] - b.t_end[down]) / b.t_end[down]) < 5e-3
    rt_fd.close()
    rt_ad.close()
    # 2D
    r, lam, chi, f = _points_2d(800, 42)
    u2 = np.vstack([r, lam, chi, f])
    e_fd, e_ad = R.env_preset("lat2d"), R.env_preset("lat2d")
    e_fd.fd_step_angle, e_fd.fd_step_r = 1e-6, 10.0
    e_ad.deriv_mode = 1
    a, b = R.RayTracer("lat2d", env=e_ad).haselgrove(u2), R.RayTracer("lat2d", env=e_fd).haselgrove(u2)
    assert np.allclose(a[[0, 1, 3]], b[[0, 1, 3]], rtol=1e-12, atol=0.0)
    assert np.max(np.abs(a[2] - b[2]) / (np.abs(b[2]) + np.abs(b[1]))) < 1e-6
    # root selection
    e_rt = R.env_preset("lat2d")
    e_rt.root_mode = 1
    rt_rt = R.RayTracer("lat2d", env=e_rt)
    mu1, dmu1 = rt_rt.phase_refractive_index(r, lam, chi, f)
    mu0, _ = R.RayTracer("lat2d").phase_refractive_index(r, lam, chi, f)
    ref = [TG.pri_lat(*p) for p in zip(r, lam, chi, f)]
    m2p = np.array([q["mu2_plus"] for q in ref])
    m2m = np.array([(q["B"] - q["F"]) / (2 * q["A"]) for q in ref])
    pos = m2p > 0
    assert 0.05 < pos.mean() < 0.95                      # both cases occur in the sample
    assert np.allclose(mu1[pos], mu0[pos], rtol=1e-13)
    neg = ~pos & (m2m > 0)
    assert neg.any() and np.allclose(mu1[neg], np.sqrt(m2m[neg]), rtol=1e-9)
    assert np.isnan(mu1[~pos & ~(m2m > 0)]).all()
    # the corrected d(mu)/d(psi) is the derivative of the selected root: check against a psi difference of mu itself
    h = 1e-6
    mu_p, _ = rt_rt.phase_refractive_index(r, lam, chi + h / 2, f)
    mu_m, _ = rt_rt.phase_refractive_index(r, lam, chi - h / 2, f)
    ok = np.isfinite(mu1) & np.isfinite(mu_p) & np.isfinite(mu_m)
    num = (mu_p - mu_m) / h                                # psi = pi/2 + dip + chi: d/dchi = d/dpsi
    assert np.max(np.abs(dmu1[ok] - num[ok]) / (np.abs(num[ok]) + 1e-3 * mu1[ok])) < 1e-5
    rt_rt.close()
    bad = R.env_preset("raymain")
    bad.deriv_mode = 1
    with pytest.raises(RuntimeError, match="LAT2D and RAY3D"):
        R.RayTracer("raymain", env=bad)
