"""Host-side mirror of the reference's interface for the hot path, over the C ABI (ctypes).

The reference is a set of Julia scripts; its interface for this path is the set of functions
  phase_refractive_index(r, lam, chi, freq) -> [mu, dmudpsi]            RayTrace_lat.jl:44-194
  refractive_index(r, th, ph, rho_r, rho_th, rho_ph, freq) -> [mu, dmudpsi, psi, Bvec]   RayTrace_3D.jl:93-219
  haselgrove!(du, u, p, t)                                               RayTrace_lat.jl:251, RayTrace_3D.jl:327
  ODEProblem(haselgrove!, u0, tspan, p); solve(prob, alg; reltol, callback, dtmax, dtmin)
                                                                         RayTrace_lat.jl:337-338, RayTrace_3D.jl:394-395
  initialize_plasmasphere(Lppi, d, Rbar, mlt) -> Lppo                   plasmasphere.jl:48-71
and this module keeps those names and argument meanings, batched: every scalar argument may be an array
(one entry per ray).  All compute runs in libraytrace_b200.so on the GPU; there is no CPU fallback and
importing this module without the built library raises.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

__all__ = [
    "RayTracer", "ODEProblem", "ODESolution", "Tsit5", "solve", "EnvParams", "SolverOpts", "initialize_plasmasphere",
    "env_preset", "solver_preset", "read_ray_start", "write_ray_start", "STATUS_NAMES", "PRESETS", "lib_path",
    "MultiRayTracer",
]

PRESETS = {"lat2d": 0, "ray3d": 1, "raymain": 2, "scratch": 3}
STATUS_NAMES = {1: "Success", 2: "Terminated", 3: "Escaped", 4: "MaxIters", 5: "DtLessThanMin", 6: "Unstable", 7: "DtNaN"}


class EnvParams(C.Structure):
    """rt_env_params (include/raytrace_b200.h)."""

    _fields_ = [("preset", C.c_int32), ("density_model", C.c_int32)] + [
        (k, C.c_double)
        for k in ("c", "re", "B0", "e", "me", "mp", "eps", "Lppi", "Lppo", "ne_Lppi", "ca_coef", "ca_trough", "iono_n0",
                  "iono_k", "iono_h", "fd_step_angle", "fd_step_r")
    ] + [("deriv_mode", C.c_int32), ("root_mode", C.c_int32)]


class SolverOpts(C.Structure):
    """rt_solver_opts (include/raytrace_b200.h)."""

    _fields_ = [
        ("adaptive", C.c_int32), ("save_mode", C.c_int32), ("max_samples", C.c_int32), ("save_stride", C.c_int32),
        ("maxiters", C.c_int64), ("t1", C.c_double), ("dt", C.c_double), ("reltol", C.c_double),
        ("abstol", C.c_double), ("dtmax", C.c_double), ("dtmin", C.c_double), ("r_esc", C.c_double),
        ("slice_attempts", C.c_int32), ("coop_threshold", C.c_int32), ("single_chain", C.c_int32), ("nan_policy", C.c_int32),
    ]


class _Result(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect")]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_int64), ("slices", C.c_int64), ("integrate_ms", C.c_double),
                ("total_ms", C.c_double), ("total_nfe", C.c_int64), ("prio_rays", C.c_int64)]


_LIB = None
_LIB_DEV = None


def lib_path():
    return _build.LIB


def _load(path):
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python ray-trace_b200/build.py` (needs nvcc). "
            "There is no CPU fallback.")
    L = C.CDLL(path)
    L.rt_last_error.restype = C.c_char_p
    L.rt_last_error.argtypes = [C.c_void_p]
    return L


def _lib():
    """The product library: the reference-facing C ABI and nothing else."""
    global _LIB
    if _LIB is None:
        _LIB = _load(lib_path())
    return _LIB


def _dev_lib():
    """The -DRT_DEV build of the same sources: + diagnostics (include/raytrace_b200_dev.h) and environment switches.
    Only diagnostics tests and tools/ use it (RayTracer(..., dev=True))."""
    global _LIB_DEV
    if _LIB_DEV is None:
        _LIB_DEV = _load(_build.LIB_DEV)
    return _LIB_DEV


def _check(rc, ctx=None, lib=None):
    if rc != 0:
        msg = (lib or _lib()).rt_last_error(ctx).decode()
        raise RuntimeError(f"raytrace_b200 error {rc}: {msg}")


def _f64(x, n=None):
    a = np.ascontiguousarray(np.atleast_1d(np.asarray(x, dtype=np.float64)))
    if n is not None and a.size != n:
        a = np.ascontiguousarray(np.broadcast_to(a, (n,)))
    return a


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _preset_id(p):
    return PRESETS[p] if isinstance(p, str) else int(p)


def env_preset(preset="lat2d"):
    e = EnvParams()
    _check(_lib().rt_env_preset(C.c_int(_preset_id(preset)), C.byref(e)))
    return e


def solver_preset(preset="lat2d"):
    o = SolverOpts()
    _check(_lib().rt_solver_preset(C.c_int(_preset_id(preset)), C.byref(o)))
    return o


def initialize_plasmasphere(Lppi, d, Rbar, mlt, re=6.3712e6, return_ne=False):
    """initialize_plasmasphere(Lppi, d, R̄, mlt) -> Lppo  (plasmasphere.jl:48-71)."""
    Lppo, ne = C.c_double(), C.c_double()
    _check(_lib().rt_initialize_plasmasphere(C.c_double(Lppi), C.c_double(d), C.c_double(Rbar), C.c_double(mlt),
                                             C.c_double(re), C.byref(Lppo), C.byref(ne)))
    return (Lppo.value, ne.value) if return_ne else Lppo.value


class SolveResult:
    """Terminal records of a batch solve (structure of arrays) plus optional trajectories."""

    def __init__(self, n, dim):
        self.n, self.dim = n, dim
        self.u_end = np.zeros((dim, n))
        self.t_end = np.zeros(n)
        self.group_delay = np.zeros(n)
        self.status = np.zeros(n, np.int32)
        self.naccept = np.zeros(n, np.int32)
        self.nreject = np.zeros(n, np.int32)
        self.nfe = np.zeros(n, np.int32)
        self.nreflect = np.zeros(n, np.int32)
        self.traj_offsets = None
        self.traj = None
        self.traj_derived = None
        self.truncated = None  # [n] int32, 1 where a trajectory did not fit max_samples (rt_traj_truncated)
        self.stats = None

    def _c(self):
        return _Result(*[_ptr(getattr(self, k)) for k, _ in _Result._fields_])

    def derived(self, i):
        """Rows (x, y, dt) — 3D: (x, y, z, dt) — of ray i's trajectory (RayTracer.solve(..., derived=True))."""
        lo, hi = self.traj_offsets[i], self.traj_offsets[i + 1]
        return self.traj_derived[lo:hi]

    def trajectory(self, i):
        """Rows (t, u..., [mu, dmudpsi, dip, psi]) of ray i — sol.t / sol[k,:] / saved_mu."""
        lo, hi = self.traj_offsets[i], self.traj_offsets[i + 1]
        return self.traj[lo:hi]


class RayTracer:
    """One context = one GPU + one environment preset + solver options."""

    def __init__(self, preset="lat2d", device=0, env=None, opts=None, dev=False):
        self._ctx = C.c_void_p()
        self._L = _dev_lib() if dev else _lib()
        _check(self._L.rt_create(C.byref(self._ctx), C.c_int(device)), lib=self._L)
        self.preset = _preset_id(preset)
        self.env = env if env is not None else env_preset(self.preset)
        self.opts = opts if opts is not None else solver_preset(self.preset)
        self.set_env(self.env)
        self.set_solver(self.opts)

    def close(self):
        if self._ctx:
            self._L.rt_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def dim(self):
        return 7 if self.env.preset == 1 else 4

    def set_env(self, env):
        _check(self._L.rt_set_env(self._ctx, C.byref(env)), self._ctx, self._L)
        self.env = env

    def set_solver(self, opts=None, **kw):
        o = opts if opts is not None else self.opts
        for k, v in kw.items():
            setattr(o, k, v)
        _check(self._L.rt_set_solver(self._ctx, C.byref(o)), self._ctx, self._L)
        self.opts = o

    def set_stream(self, cuda_stream_ptr):
        _check(self._L.rt_set_stream(self._ctx, C.c_void_p(cuda_stream_ptr)), self._ctx, self._L)

    def stats(self):
        s = Stats()
        _check(self._L.rt_get_stats(self._ctx, C.byref(s)), self._ctx, self._L)
        return s

    def measure_fp64_peak(self):
        v = C.c_double()
        _check(self._L.rt_measure_fp64_peak(self._ctx, C.byref(v)), self._ctx, self._L)
        return v.value

    # -- reference function names, batched ---------------------------------------------------------
    def phase_refractive_index(self, r, lam, chi, freq):
        """phase_refractive_index(r, λ, χ, freq) -> (mu, dmudpsi)  (RayTrace_lat.jl:44)."""
        r = _f64(r)
        n = max(np.size(r), np.size(lam), np.size(chi), np.size(freq))
        r, lam, chi, freq = _f64(r, n), _f64(lam, n), _f64(chi, n), _f64(freq, n)
        mu, dmu = np.zeros(n), np.zeros(n)
        _check(self._L.rt_phase_refractive_index_2d(self._ctx, C.c_int64(n), _ptr(r), _ptr(lam), _ptr(chi), _ptr(freq),
                                                   _ptr(mu), _ptr(dmu)), self._ctx, self._L)
        return mu, dmu

    def refractive_index(self, r, th, ph, rho_r, rho_th, rho_ph, freq):
        """refractive_index(r,θ,ϕ,ρ_r,ρ_θ,ρ_ϕ,freq) -> (mu, dmudpsi, psi, Bvec[3,n])  (RayTrace_3D.jl:93)."""
        args = [r, th, ph, rho_r, rho_th, rho_ph, freq]
        n = max(np.size(a) for a in args)
        a = [_f64(x, n) for x in args]
        mu, dmu, psi, b = np.zeros(n), np.zeros(n), np.zeros(n), np.zeros((3, n))
        _check(self._L.rt_refractive_index_3d(self._ctx, C.c_int64(n), *[_ptr(x) for x in a], _ptr(mu), _ptr(dmu),
                                             _ptr(psi), _ptr(b)), self._ctx, self._L)
        return mu, dmu, psi, b

    def haselgrove(self, u, p=None, t=None):
        """haselgrove!(du,u,p,t) for a batch: u is [dim] or [dim, n]; p = [freq] in 3D.  Returns du."""
        u = np.asarray(u, dtype=np.float64)
        single = u.ndim == 1
        u2 = np.ascontiguousarray(u.reshape(self.dim, -1))
        n = u2.shape[1]
        du = np.zeros_like(u2)
        if self.dim == 7:
            f = _f64(p[0] if isinstance(p, (list, tuple)) else p, n)
            _check(self._L.rt_haselgrove_3d(self._ctx, C.c_int64(n), _ptr(u2), _ptr(f), _ptr(du)), self._ctx, self._L)
        else:
            _check(self._L.rt_haselgrove_2d(self._ctx, C.c_int64(n), _ptr(u2), _ptr(du)), self._ctx, self._L)
        return du[:, 0] if single else du

    def plasmasphere(self, r, lam):
        """plasmasphere.jl, batched: returns dict(L=L_rλ(r,λ) :122, plasmasphere_eq :73, diffusive_equilibrium :96,
        ionosphere_eq :108, ne_total :171), each [n], densities in cm^-3."""
        n = max(np.size(r), np.size(lam))
        r, lam = _f64(r, n), _f64(lam, n)
        out = np.zeros((5, n))
        _check(self._L.rt_plasmasphere_density(self._ctx, C.c_int64(n), _ptr(r), _ptr(lam), _ptr(out)), self._ctx, self._L)
        return dict(zip(("L", "plasmasphere_eq", "diffusive_equilibrium", "ionosphere_eq", "ne_total"), out))

    def meridian_maps(self, x, y):
        """plasmasphere.jl:157-175 / RayTrace_3D.jl:504-517 on the grid x[nx] × y[ny] (metres): dict of [nx, ny]
        arrays L, ne_total (cm^-3), log10_ne_total, Bmag — the inputs of the reference's heatmaps and L contours."""
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
        y = np.ascontiguousarray(np.asarray(y, dtype=np.float64))
        out = np.zeros((4, x.size * y.size))
        _check(self._L.rt_meridian_maps(self._ctx, C.c_int64(x.size), C.c_int64(y.size), _ptr(x), _ptr(y), _ptr(out)), self._ctx, self._L)
        return {k: out[i].reshape(x.size, y.size) for i, k in enumerate(("L", "ne_total", "log10_ne_total", "Bmag"))}

    def debug_rhs3d(self, u, freq):
        """Diagnostics: both versions of the 3D right-hand side on chosen states u [n, 7] at frequency freq [n]:
        returns (du_fast [n, 7], ok [n] bool — the branch-free code vouches for its result —, du_exact [n, 7])."""
        u = np.ascontiguousarray(np.asarray(u, dtype=np.float64))
        if u.ndim != 2 or u.shape[1] != 7:
            raise ValueError("u must be [n, 7]")
        n = u.shape[0]
        inp = np.ascontiguousarray(np.concatenate([u, _f64(freq, n)[:, None]], axis=1))
        df, de, ok = np.zeros((n, 7)), np.zeros((n, 7)), np.zeros(n, dtype=np.int32)
        _check(self._L.rt_debug_rhs3d(self._ctx, C.c_int(n), _ptr(inp), _ptr(df), ok.ctypes.data_as(C.c_void_p), _ptr(de)),
               self._ctx, self._L)
        return df, ok.astype(bool), de

    # -- the solve ---------------------------------------------------------------------------------
    def solve(self, launch, with_trajectories=None, derived=False):
        """Trace a launch list.  2D: columns (r, θ, χ, f, t) [5, n];  3D: (r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, T0, f, t) [9, n]."""
        launch = np.ascontiguousarray(np.asarray(launch, dtype=np.float64))
        ncol = 9 if self.dim == 7 else 5
        if launch.ndim != 2 or launch.shape[0] != ncol:
            raise ValueError(f"launch list must be [{ncol}, n]")
        n = launch.shape[1]
        res = SolveResult(n, self.dim)
        cres = res._c()
        cols = [_ptr(launch[k]) for k in range(ncol)]
        fn = self._L.rt_solve_3d if self.dim == 7 else self._L.rt_solve_2d
        _check(fn(self._ctx, C.c_int64(n), *cols, C.byref(cres)), self._ctx, self._L)
        res.stats = self.stats()
        want = self.opts.save_mode if with_trajectories is None else with_trajectories
        if want and self.opts.save_mode and n > 0:
            total = C.c_int64()
            _check(self._L.rt_traj_size(self._ctx, C.byref(total)), self._ctx, self._L)
            ncolt = self._L.rt_traj_columns(self._ctx)
            res.traj_offsets = np.zeros(n + 1, np.int64)
            res.traj = np.zeros((total.value, ncolt))
            _check(self._L.rt_traj_copy(self._ctx, _ptr(res.traj_offsets), _ptr(res.traj)), self._ctx, self._L)
            res.truncated = np.zeros(n, np.int32)
            _check(self._L.rt_traj_truncated(self._ctx, _ptr(res.truncated)), self._ctx, self._L)
            if derived:  # Cartesian track and step sizes, computed on the device (RayTrace_lat.jl:351-352,369-372)
                res.traj_derived = np.zeros((total.value, self._L.rt_traj_derived_columns(self._ctx)))
                _check(self._L.rt_traj_derived(self._ctx, _ptr(res.traj_offsets), _ptr(res.traj_derived)), self._ctx, self._L)
        return res


class MultiRayTracer:
    """The same solve over several GPUs of one box from ONE process (rt_multi_* / rt_solve_*_multi): one context and one
    host thread per device, block-cyclic launch-list partition inside the library, results in launch order in the
    caller's arrays.  devices: list of CUDA device indices (None = all visible)."""

    def __init__(self, preset="lat2d", devices=None, env=None, opts=None, block=0):
        self._m = C.c_void_p()
        L = _lib()
        L.rt_multi_last_error.restype = C.c_char_p
        L.rt_multi_last_error.argtypes = [C.c_void_p]
        L.rt_multi_global_index.restype = C.c_int64
        L.rt_multi_ctx.restype = C.c_void_p
        n = 0 if devices is None else len(devices)
        arr = None if devices is None else (C.c_int * n)(*devices)
        _check(L.rt_multi_create(C.byref(self._m), C.c_int(n), arr))
        self.ngpu = L.rt_multi_count(self._m)
        self.preset = _preset_id(preset)
        self.env = env if env is not None else env_preset(self.preset)
        self.opts = opts if opts is not None else solver_preset(self.preset)
        self._mcheck(L.rt_multi_set_env(self._m, C.byref(self.env)))
        self._mcheck(L.rt_multi_set_solver(self._m, C.byref(self.opts)))
        if block:
            self._mcheck(L.rt_multi_set_block(self._m, C.c_int64(block)))

    def _mcheck(self, rc):
        if rc != 0:
            raise RuntimeError(f"raytrace_b200 error {rc}: {_lib().rt_multi_last_error(self._m).decode()}")

    def close(self):
        if self._m:
            _lib().rt_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def dim(self):
        return 7 if self.env.preset == 1 else 4

    def set_solver(self, opts=None, **kw):
        o = opts if opts is not None else self.opts
        for k, v in kw.items():
            setattr(o, k, v)
        self._mcheck(_lib().rt_multi_set_solver(self._m, C.byref(o)))
        self.opts = o

    def partition(self, n):
        """(block length, rays per device) of a solve of n rays."""
        b = C.c_int64()
        cnt = (C.c_int64 * self.ngpu)()
        self._mcheck(_lib().rt_multi_partition(self._m, C.c_int64(n), C.byref(b), cnt))
        return b.value, list(cnt)

    def global_index(self, n, gpu, local):
        return int(_lib().rt_multi_global_index(self._m, C.c_int64(n), C.c_int(gpu), C.c_int64(local)))

    def stats(self, gpu):
        s = Stats()
        self._mcheck(_lib().rt_multi_get_stats(self._m, C.c_int(gpu), C.byref(s)))
        return s

    def solve(self, launch, res=None):
        """Trace a launch list [5 or 9, n] (host array; pinned arrays in `res` — a SolveResult — get the early copy)."""
        launch = np.ascontiguousarray(np.asarray(launch, dtype=np.float64))
        ncol = 9 if self.dim == 7 else 5
        if launch.ndim != 2 or launch.shape[0] != ncol:
            raise ValueError(f"launch list must be [{ncol}, n]")
        n = launch.shape[1]
        res = res if res is not None else SolveResult(n, self.dim)
        cres = res._c()
        fn = _lib().rt_solve_3d_multi if self.dim == 7 else _lib().rt_solve_2d_multi
        self._mcheck(fn(self._m, C.c_int64(n), *[_ptr(launch[k]) for k in range(ncol)], C.byref(cres)))
        return res


# ---- SciML-shaped convenience layer: ODEProblem / solve / ODESolution ------------------------------------
class Tsit5:
    """Algorithm tag, as in `solve(prob, Tsit5())` (scratch.jl:614)."""


class _Haselgrove:
    def __init__(self, preset):
        self.preset = _preset_id(preset)


def haselgrove(preset="lat2d"):
    """The RHS handle to put into ODEProblem: haselgrove!(du,u,p,t) of the named reference file."""
    return _Haselgrove(preset)


__all__.append("haselgrove")


class ODEProblem:
    """ODEProblem(haselgrove!, u0, tspan, p)  (RayTrace_lat.jl:337, RayTrace_3D.jl:394).

    u0 may be [dim] (one ray, as in the reference) or [dim, n] (a sweep); p = [] in 2D, [freq] in 3D
    (freq scalar or [n])."""

    def __init__(self, f, u0, tspan, p=()):
        if not isinstance(f, _Haselgrove):
            raise TypeError("f must be raytrace_b200.haselgrove(preset): arbitrary Python RHS functions are not "
                            "supported (the RHS runs fused on the GPU)")
        self.f, self.tspan, self.p = f, (float(tspan[0]), float(tspan[1])), p
        self.u0 = np.asarray(u0, dtype=np.float64)


class ODESolution:
    """sol.t, sol.u, sol[i, :] (1-based component i like Julia's sol[i,:] is NOT emulated: 0-based), sol.retcode."""

    def __init__(self, res, k):
        tr = res.trajectory(k) if res.traj is not None else None
        self.retcode = STATUS_NAMES[int(res.status[k])]
        self.status = int(res.status[k])
        self.u_end = res.u_end[:, k].copy()
        self.t_end = float(res.t_end[k])
        self.group_delay = float(res.group_delay[k])
        self.naccept, self.nreject, self.nfe = int(res.naccept[k]), int(res.nreject[k]), int(res.nfe[k])
        self.nreflect = int(res.nreflect[k])
        dim = res.dim
        if tr is not None:
            self.t = tr[:, 0].copy()
            self.u = tr[:, 1:1 + dim].copy()
            self.saved_mu = tr[:, 1 + dim:].copy()  # saveμ_cb: [mu, dmudpsi, dip, psi] (2D only)
            self.truncated = bool(res.truncated[k])  # samples were dropped before the final one (max_samples)
        else:
            self.truncated = False
            self.t, self.u, self.saved_mu = np.array([self.t_end]), self.u_end[None, :], None

    def __getitem__(self, key):
        i, sl = key
        return self.u[sl, i]


def solve(prob, alg=None, reltol=None, abstol=None, dtmax=None, dtmin=None, maxiters=None, adaptive=True, dt=None,
          callback=True, save_everystep=True, max_samples=None, device=0, r_esc=None, tracer=None):
    """solve(prob, Tsit5(); reltol, abstol, dtmax, dtmin, maxiters, adaptive, dt, callback) -> ODESolution(s).

    Defaults follow SciML (reltol 1e-3, abstol 1e-6, maxiters 1e5) as at RayTrace_3D.jl:395; pass reltol=1e-7,
    dtmax=1e6, dtmin=1e-8 for the RayTrace_lat.jl:338 call.  `callback=True` is the reference's re_cb (terminate
    at Earth's surface), which is always on in the kernel; callback=False is not supported.
    max_samples (rows kept per ray when save_everystep): default maxiters + 2, i.e. nothing can be dropped, as long as
    all trajectories fit 8 GB of device memory, else as many rows as do; sol.truncated tells when rows were dropped
    (the final sample is always kept)."""
    if alg is not None and not isinstance(alg, Tsit5):
        raise TypeError("only Tsit5() is implemented (north_star replaces CVODE_BDF by Tsit5)")
    if not callback:
        raise ValueError("the Earth-hit callback cannot be disabled")
    preset = prob.f.preset
    rt = tracer or RayTracer(preset, device=device)
    o = solver_preset(preset)
    o.reltol = 1e-3 if reltol is None else reltol
    o.abstol = 1e-6 if abstol is None else abstol
    o.dtmax = 0.0 if dtmax is None else dtmax
    o.dtmin = 0.0 if dtmin is None else dtmin
    o.maxiters = 100000 if maxiters is None else int(maxiters)
    o.adaptive = 1 if adaptive else 0
    o.dt = 0.0 if dt is None else dt
    o.t1 = prob.tspan[1]
    o.r_esc = 0.0 if r_esc is None else r_esc
    u0 = prob.u0
    single = u0.ndim == 1
    u0 = u0.reshape(rt.dim, -1)
    n = u0.shape[1]
    if max_samples is None:
        row_bytes = 8 * (8 if rt.dim == 7 else 9)
        max_samples = int(max(2, min(o.maxiters + 2, (8 << 30) // (row_bytes * max(n, 1)))))
    o.save_mode = 1 if save_everystep else 0
    o.max_samples = max_samples if save_everystep else 0
    rt.set_solver(o)
    t0 = np.full(n, prob.tspan[0])
    if rt.dim == 7:
        f = _f64(prob.p[0], n)
        launch = np.vstack([u0, f[None, :], t0[None, :]])
    else:
        launch = np.vstack([u0, t0[None, :]])
    res = rt.solve(launch)
    sols = [ODESolution(res, k) for k in range(n)]
    if tracer is None:
        rt.close()
    return sols[0] if single else sols


# ---- ray_start.dat: the launch list README.md:11 plans (r, θ, χ, f, t); 3D: r θ ϕ ρ_r ρ_θ ρ_ϕ T0 f t ----------
def read_ray_start(path):
    """Read an ASCII launch list (# comments, whitespace-separated FP64) -> array [ncol, n] (ncol 5 or 9)."""
    rows = []
    with open(path) as fh:
        for ln, line in enumerate(fh, 1):
            line = line.split("#", 1)[0].strip()
            if not line:
                continue
            vals = [float(v) for v in line.split()]
            if rows and len(vals) != len(rows[0]):
                raise ValueError(f"{path}:{ln}: expected {len(rows[0])} columns, found {len(vals)}")
            if len(vals) not in (5, 9):
                raise ValueError(f"{path}:{ln}: a launch row has 5 (2D) or 9 (3D) columns, found {len(vals)}")
            rows.append(vals)
    if not rows:
        return np.zeros((5, 0))
    return np.ascontiguousarray(np.array(rows, dtype=np.float64).T)


def write_ray_start(path, launch, header=None):
    launch = np.asarray(launch, dtype=np.float64)
    with open(path, "w") as fh:
        if header:
            fh.write("# " + header + "\n")
        fh.write("# r theta chi f t\n" if launch.shape[0] == 5 else "# r theta phi rho_r rho_theta rho_phi T0 f t\n")
        for k in range(launch.shape[1]):
            fh.write(" ".join(repr(float(v)) for v in launch[:, k]) + "\n")


def _solve_device(self, n, col_ptrs, res_ptrs):
    """rt_solve_{2d,3d}_device: every pointer is a device address (ints), e.g. torch tensors' data_ptr().
    col_ptrs: 5 (2D) or 9 (3D) launch columns; res_ptrs: dict with the rt_result member names."""
    cres = _Result(*[C.c_void_p(res_ptrs[k]) for k, _ in _Result._fields_])
    fn = self._L.rt_solve_3d_device if self.dim == 7 else self._L.rt_solve_2d_device
    _check(fn(self._ctx, C.c_int64(n), *[C.c_void_p(p) for p in col_ptrs], C.byref(cres)), self._ctx, self._L)
    return self.stats()


RayTracer.solve_device = _solve_device


def _solve_device_static(self, n, col_ptrs, res_ptrs):
    """rt_solve_3d_device_static: solve_device without device allocations (the integrator's working set lives in a
    static device array of the library); 3D only, n <= STATIC_MAX_RAYS, save_mode 0.  Not in the reference: added for
    callers that may not call cudaMalloc (include/raytrace_b200.h)."""
    cres = _Result(*[C.c_void_p(res_ptrs[k]) for k, _ in _Result._fields_])
    _check(self._L.rt_solve_3d_device_static(self._ctx, C.c_int64(n), *[C.c_void_p(p) for p in col_ptrs], C.byref(cres)),
           self._ctx, self._L)
    return self.stats()


STATIC_MAX_RAYS = 16384  # RT_STATIC_MAX_RAYS
RayTracer.solve_device_static = _solve_device_static
