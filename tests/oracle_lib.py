"""ctypes binding of the CPU oracle (oracle/librt_oracle.so).  Test infrastructure only.

Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "librt_oracle.so")

LAT2D, RAY3D, RAYMAIN, SCRATCH = 0, 1, 2, 3


class OrcEnv(C.Structure):
    _fields_ = [("preset", C.c_int), ("density_model", C.c_int)] + [
        (k, C.c_double)
        for k in ("c", "re", "B0", "e", "me", "mp", "eps", "Kp_max", "Lppi", "d", "Rbar", "mlt", "Lppo", "ne_Lppi",
                  "fd_step_angle", "fd_step_r")
    ]


class SolverOpts(C.Structure):
    """Mirror of rt_solver_opts in include/raytrace_b200.h."""

    _fields_ = [
        ("adaptive", C.c_int32),
        ("save_mode", C.c_int32),
        ("max_samples", C.c_int32),
        ("save_stride", C.c_int32),
        ("maxiters", C.c_int64),
        ("t1", C.c_double),
        ("dt", C.c_double),
        ("reltol", C.c_double),
        ("abstol", C.c_double),
        ("dtmax", C.c_double),
        ("dtmin", C.c_double),
        ("r_esc", C.c_double),
        ("slice_attempts", C.c_int32),
        ("coop_threshold", C.c_int32),
        ("single_chain", C.c_int32),
        ("nan_policy", C.c_int32),
    ]


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_initialize_plasmasphere.restype = C.c_double
        for n in ("orc_ddr_2d", "orc_ddlam_2d", "orc_ddchi_2d", "orc_ddf_2d", "orc_dd_3d", "orc_ddrho_3d",
                  "orc_plasmasphere_eq", "orc_diffusive_equilibrium", "orc_ionosphere_eq", "orc_L_rlam"):
            getattr(_lib, n).restype = C.c_double
        _lib.orc_interior_events.restype = C.c_longlong
        assert _lib.orc_sizeof_env() == C.sizeof(OrcEnv)
    return _lib


def env(preset=LAT2D):
    e = OrcEnv()
    lib().orc_env_default(C.c_int(preset), C.byref(e))
    return e


def opts(adaptive=1, t1=5.0e9, dt=0.0, reltol=1e-7, abstol=1e-6, dtmax=1e6, dtmin=1e-8, maxiters=100000,
         r_esc=0.0, save_mode=0, max_samples=0, save_stride=1, slice_attempts=0, nan_policy=0):
    return SolverOpts(adaptive, save_mode, max_samples, save_stride, maxiters, t1, dt, reltol, abstol, dtmax,
                      dtmin, r_esc, slice_attempts, 0, 0, nan_policy)


def _d(x):
    return C.c_double(float(x))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def phase_refractive_index_2d(E, r, lam, chi, f):
    out = np.zeros(2)
    lib().orc_phase_refractive_index_2d(C.byref(E), _d(r), _d(lam), _d(chi), _d(f), _p(out))
    return out


def dd_2d(E, which, r, lam, chi, f):
    fn = getattr(lib(), {"r": "orc_ddr_2d", "lam": "orc_ddlam_2d", "chi": "orc_ddchi_2d", "f": "orc_ddf_2d"}[which])
    return fn(C.byref(E), _d(r), _d(lam), _d(chi), _d(f))


def haselgrove_2d(E, u):
    u = np.ascontiguousarray(u, dtype=np.float64)
    du = np.zeros(4)
    lib().orc_haselgrove_2d(C.byref(E), _p(u), _p(du))
    return du


def magnetic_field(E, r, th, ph):
    b = np.zeros(3)
    lib().orc_magnetic_field(C.byref(E), _d(r), _d(th), _d(ph), _p(b))
    return b


def refractive_index_3d(E, r, th, ph, rr, rt, rp, f):
    out = np.zeros(6)
    lib().orc_refractive_index_3d(C.byref(E), _d(r), _d(th), _d(ph), _d(rr), _d(rt), _d(rp), _d(f), _p(out))
    return out


def dd_3d(E, which, *a):
    return lib().orc_dd_3d(C.byref(E), C.c_int(which), *[_d(x) for x in a])


def ddrho_3d(E, k, *a):
    return lib().orc_ddrho_3d(C.byref(E), C.c_int(k), *[_d(x) for x in a])


def haselgrove_3d(E, u, f):
    u = np.ascontiguousarray(u, dtype=np.float64)
    du = np.zeros(7)
    lib().orc_haselgrove_3d(C.byref(E), _p(u), _d(f), _p(du))
    return du


class RayOut:
    def __init__(self, u_end, t_end, ints, traj, truncated=False):
        self.u_end, self.t_end, self.truncated = u_end, t_end, truncated
        self.status, self.naccept, self.nreject, self.nfe, self.nreflect = [int(v) for v in ints]
        self.traj = traj


def solve_ray_2d(E, o, u0, t0=0.0, cap=0):
    u0 = np.ascontiguousarray(u0, dtype=np.float64)
    ue, te, ints = np.zeros(4), C.c_double(), np.zeros(5, np.int32)
    traj = np.zeros((cap, 9)) if cap else None
    ns = C.c_int64(0)
    lib().orc_solve_ray_2d(C.byref(E), C.byref(o), _p(u0), _d(t0), _p(ue), C.byref(te), _p(ints),
                           _p(traj) if cap else None, C.c_int64(cap), C.byref(ns))
    return RayOut(ue, te.value, ints, traj[: abs(ns.value)] if cap else None, ns.value < 0)


def solve_ray_3d(E, o, u0, f, t0=0.0, cap=0):
    u0 = np.ascontiguousarray(u0, dtype=np.float64)
    ue, te, ints = np.zeros(7), C.c_double(), np.zeros(5, np.int32)
    traj = np.zeros((cap, 8)) if cap else None
    ns = C.c_int64(0)
    lib().orc_solve_ray_3d(C.byref(E), C.byref(o), _p(u0), _d(f), _d(t0), _p(ue), C.byref(te), _p(ints),
                           _p(traj) if cap else None, C.c_int64(cap), C.byref(ns))
    return RayOut(ue, te.value, ints, traj[: abs(ns.value)] if cap else None, ns.value < 0)


def solve_batch(E, o, launch, dim3, nthreads=1):
    """launch: [ncol][n] component-major (2D: r,theta,chi,f,t0; 3D: r,theta,phi,rho_r,rho_t,rho_p,T0,f,t0)."""
    launch = np.ascontiguousarray(launch, dtype=np.float64)
    n = launch.shape[1]
    dim = 7 if dim3 else 4
    ue, te, ints = np.zeros((dim, n)), np.zeros(n), np.zeros((5, n), np.int32)
    rc = lib().orc_solve_batch(C.byref(E), C.byref(o), C.c_int(int(dim3)), C.c_int64(n), _p(launch), _p(ue),
                               _p(te), _p(ints), C.c_int(nthreads))
    assert rc == 0
    return {"u_end": ue, "t_end": te, "status": ints[0], "naccept": ints[1], "nreject": ints[2], "nfe": ints[3],
            "nreflect": ints[4]}


def solve_test_ode(which, o, u0, t0=0.0, re_evt=0.0, cap=0):
    u0 = np.ascontiguousarray(u0, dtype=np.float64)
    dim = 3 if which else 2
    ue, te, ints = np.zeros(dim), C.c_double(), np.zeros(5, np.int32)
    traj = np.zeros((cap, 1 + dim)) if cap else None
    ns = C.c_int64(0)
    lib().orc_solve_test_ode(C.c_int(which), C.byref(o), _p(u0), _d(t0), _d(re_evt), _p(ue), C.byref(te),
                             _p(ints), _p(traj) if cap else None, C.c_int64(cap), C.byref(ns))
    return RayOut(ue, te.value, ints, traj[: abs(ns.value)] if cap else None, ns.value < 0)


def interior_events(reset=False):
    """Events the oracle located through ContinuousCallback's interior safety check (Theta = i/9) since the last reset
    (single-thread runs only)."""
    return int(lib().orc_interior_events(C.c_int(int(reset))))


MATH = {"exp": 0, "exp10": 1, "log": 2, "log10": 3, "sin": 4, "cos": 5, "tan": 6, "atan": 7, "acos": 8, "pow_m4p5": 9,
        "pow_b1": 10, "pow_b2": 11}


def meridian_maps(E, x, y):
    """plasmasphere.jl:157-175 / RayTrace_3D.jl:504-517 on the grid x[nx] x y[ny] (metres):
    [L, ne_total, log10 ne_total, |B|][nx*ny]."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    out = np.zeros((4, x.size * y.size))
    lib().orc_meridian_maps(C.byref(E), C.c_int64(x.size), C.c_int64(y.size), _p(x), _p(y), _p(out))
    return out


def math(name, x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    lib().orc_math(C.c_int(MATH[name]), C.c_int64(x.size), _p(x), _p(y))
    return y
