"""The C-ABI library loads, exports every symbol include/raytrace_b200.h declares, its host-side entry points agree
with the oracle, and it fails loudly (no CPU fallback) without a GPU.  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib as O
import raytrace_b200 as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "raytrace_b200.h")).read()
    return sorted(set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = C.CDLL(R.lib_path())
    names = declared_symbols()
    assert len(names) >= 22, names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/raytrace_b200.h but not exported"
    assert lib.rt_version() >= 100


def test_struct_layouts_match_header():
    assert C.sizeof(R.EnvParams) == 8 + 17 * 8 + 8
    assert C.sizeof(R.SolverOpts) == 4 * 4 + 8 + 7 * 8 + 16
    assert C.sizeof(R.SolverOpts) == C.sizeof(O.SolverOpts)


@pytest.mark.parametrize("name,pid", [("lat2d", O.LAT2D), ("ray3d", O.RAY3D), ("raymain", O.RAYMAIN), ("scratch", O.SCRATCH)])
def test_env_preset_matches_oracle_environment(name, pid):
    e, E = R.env_preset(name), O.env(pid)
    for k in ("c", "re", "B0", "e", "me", "mp", "eps", "Lppi", "Lppo", "ne_Lppi"):
        assert getattr(e, k) == getattr(E, k), k
    assert e.ca_trough == 5800.0 + 300.0 * E.mlt
    want = 0.0510054794520548 if name == "ray3d" else 0.12757508166977483
    assert abs(e.ca_coef - want) < 1e-15


def test_solver_presets_follow_the_reference_calls():
    o = R.solver_preset("lat2d")  # RayTrace_lat.jl:338
    assert (o.reltol, o.dtmax, o.dtmin, o.t1, o.maxiters) == (1e-7, 1e6, 1e-8, 5.0e9, 100000)
    o = R.solver_preset("ray3d")  # RayTrace_3D.jl:395: defaults
    assert (o.reltol, o.abstol, o.dtmax, o.t1) == (1e-3, 1e-6, 0.0, 5.0e9)
    assert R.solver_preset("scratch").t1 == 5.0e10  # scratch.jl:390


def test_initialize_plasmasphere_host_entry_point():
    Lppo, ne = R.initialize_plasmasphere(4.22, 0, 90, 2, return_ne=True)
    assert Lppo == 4.387901808136615 and abs(ne - 403.7856649204069) < 1e-10
    ne_o = C.c_double()
    assert O.lib().orc_initialize_plasmasphere(O._d(4.22), O._d(0), O._d(90), O._d(2), O._d(6.3712e6), C.byref(ne_o)) == Lppo
    # a different Kp moves the plasmapause
    assert R.initialize_plasmasphere(5.6 - 0.46 * 5, 0, 90, 2) < Lppo


def test_no_cpu_fallback_without_a_device(have_gpu):
    if have_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CUDA device"):
        R.RayTracer("lat2d")


def test_invalid_arguments_are_api_errors_not_crashes():
    lib = C.CDLL(R.lib_path())
    assert lib.rt_env_preset(C.c_int(9), C.byref(R.EnvParams())) == -1
    assert lib.rt_solver_preset(C.c_int(0), None) == -1
    assert lib.rt_create(None, C.c_int(0)) == -1
    lib.rt_last_error.restype = C.c_char_p
    assert b"rt_create" in lib.rt_last_error(None)
    assert lib.rt_set_env(None, None) == -1 and lib.rt_traj_columns(None) == -1


def test_ray_start_roundtrip(tmp_path):
    from raytrace_b200 import sweeps
    for L in (sweeps.sweep_2d(3, 4), sweeps.sweep_3d(2, 2, 3)):
        p = tmp_path / "ray_start.dat"
        R.write_ray_start(str(p), L, header="test")
        back = R.read_ray_start(str(p))
        assert back.shape == L.shape and np.array_equal(back, L)  # repr() round-trips FP64 exactly
    (tmp_path / "bad.dat").write_text("1 2 3\n")
    with pytest.raises(ValueError):
        R.read_ray_start(str(tmp_path / "bad.dat"))
    (tmp_path / "empty.dat").write_text("# nothing\n\n")
    assert R.read_ray_start(str(tmp_path / "empty.dat")).shape == (5, 0)


def test_sweeps_are_the_named_configs():
    from raytrace_b200 import sweeps
    a = sweeps.sweep_2d()
    assert a.shape == (5, 100000) and a[3].min() == 1e3 and abs(a[3].max() - 3e4) < 1e-9
    b = sweeps.sweep_3d(10, 10, 10)
    assert b.shape == (9, 1000) and np.all(b[3:6] != 0.0)  # no zero rho component (RayTrace_3D.jl:304-310)
    assert np.allclose(np.sum(b[3:6] ** 2, axis=0), 1.0)
    parts = [sweeps.partition(1000003, r, 8) for r in range(8)]
    assert parts[0][0] == 0 and parts[-1][1] == 1000003 and all(parts[i][1] == parts[i + 1][0] for i in range(7))


def test_launch_angle_helper_reproduces_the_scratch_cell():
    """scratch.jl:725-731 with hbi = 1e6, b = -pi/3."""
    from raytrace_b200 import sweeps
    import math
    re, hbi, b = 6.3712e6, 1.0e6, -math.pi / 3
    alpha = math.acos((re * math.sin(b) ** 2.0 + math.cos(b) * ((re + hbi) ** 2.0 - re ** 2.0 * math.sin(b) ** 2.0) ** 0.5) / (re + hbi))
    a, chi = sweeps.launch_angle(b)
    assert abs(a - alpha) < 1e-15 and abs(chi - (b - alpha)) < 1e-15 and 0 < a < 0.3


def test_block_cyclic_partition_mirror_and_multi_create_without_gpu(have_gpu):
    """The multi-GPU partition of rt_solve_*_multi (blocks of B rays go round the devices), as sweeps mirrors it; and
    rt_multi_create fails loudly without a device like rt_create does."""
    from raytrace_b200 import sweeps
    for n, B, N in ((1000, 128, 3), (5, 2, 2), (0, 4, 2), (16384 * 8 + 5, 16384, 8)):
        parts = [sweeps.block_cyclic_indices(n, B, N, r) for r in range(N)]
        assert sorted(np.concatenate(parts).tolist()) == list(range(n))
        for r, p in enumerate(parts):
            j = np.arange(p.size)
            assert np.array_equal(p, (j // B) * (B * N) + r * B + j % B)
    if not have_gpu:
        m = C.c_void_p()
        lib = C.CDLL(R.lib_path())
        assert lib.rt_multi_create(C.byref(m), C.c_int(2), None) == -2 and not m.value  # RT_ERR_NO_DEVICE
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            R.MultiRayTracer("ray3d", devices=[0, 1])


def test_product_library_exports_only_the_reference_facing_surface():
    """libraytrace_b200.so = include/raytrace_b200.h exactly; the diagnostics of include/raytrace_b200_dev.h exist in the
    -DRT_DEV build only; the product binary carries none of the RT_* switch names."""
    import subprocess
    from raytrace_b200 import build as B

    def exported(path):
        out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
        return {ln.split()[-1] for ln in out.splitlines() if " T " in ln and ln.split()[-1].startswith("rt_")}

    dev_src = open(os.path.join(ROOT, "include", "raytrace_b200_dev.h")).read()
    dev_names = set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", dev_src))
    assert len(dev_names) == 6
    prod, dev = exported(B.LIB), exported(B.LIB_DEV)
    assert prod == set(declared_symbols()), prod ^ set(declared_symbols())
    assert dev == prod | dev_names
    # (the statically linked CUDA runtime imports getenv for its own CUDA_* variables; the library's own switch names
    # are what must be absent)
    blob = open(B.LIB, "rb").read()
    assert b"RT_COOP_POLICY" not in blob and b"RT_EARLY_MIN" not in blob
    assert b"RT_COOP_POLICY" in open(B.LIB_DEV, "rb").read()
