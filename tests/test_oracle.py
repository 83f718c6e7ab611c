"""CPU tests of the oracle: reference-stated identities, the survey's probe values, and golden fixtures from the
independent restatement (tools/make_golden.py).  No GPU."""
import json
import math
import os

import numpy as np
import pytest

import oracle_lib as O

RE = 6.3712e6
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "refractive_index.json")))


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def test_presolve_matches_golden_and_probe():
    E = O.env(O.LAT2D)
    assert E.Lppi == pytest.approx(4.22, rel=1e-15)
    assert E.Lppo == GOLD["constants"]["Lppo"] == 4.387901808136615  # SURVEY §3.1 [PROBE]
    assert rel(E.ne_Lppi, 403.7856649204069) < 1e-14
    assert O.env(O.RAY3D).B0 == 3.12e-5 and E.B0 == 3.0696381e-5  # RayTrace_3D.jl:40 vs RayTrace_lat.jl:15


def test_probe_values_at_reference_u0():
    """values at u0 = [re+1e6, pi/4, 0, 1000] (RayTrace_lat.jl:333)."""
    E = O.env(O.LAT2D)
    mu, dmu = O.phase_refractive_index_2d(E, RE + 1e6, math.pi / 4, 0.0, 1000.0)
    assert rel(mu, 150.94166037288866) < 1e-13
    assert rel(dmu, 115.05618007269247) < 1e-12
    assert O.dd_2d(E, "r", RE + 1e6, math.pi / 4, 0.0, 1000.0) == 0.0  # FP64 absorption, RayTrace_lat.jl:200-207
    assert rel(O.dd_2d(E, "lam", RE + 1e6, math.pi / 4, 0.0, 1000.0), -40.479962) < 1e-3  # noise-dominated FD
    assert rel(O.dd_2d(E, "chi", RE + 1e6, math.pi / 4, 0.0, 1000.0), -39.651741) < 1e-3
    assert rel(O.dd_2d(E, "f", RE + 1e6, math.pi / 4, 0.0, 1000.0), -0.14482655) < 1e-6
    du = O.haselgrove_2d(E, [RE + 1e6, math.pi / 4, 0.0, 1000.0])
    assert rel(du[0], 6.6250761885723545e-3) < 1e-12 and rel(du[1], -6.850992787092621e-10) < 1e-12
    assert rel(du[3], 1.3513701443800207e-10) < 1e-9 and rel(du[2], -2.41043e-10) < 1e-3


def test_probe_values_3d_and_nan_quirk():
    E = O.env(O.RAY3D)
    o = O.refractive_index_3d(E, RE + 1e6, math.pi / 4, 0.0, 1.0, 1.0, 1e-3, 1000.0)
    assert rel(o[0], 146.16017762798208) < 1e-13 and rel(o[1], 73.60788197035532) < 1e-12
    assert rel(o[2], 2.819841349194276) < 1e-14
    du = O.haselgrove_3d(E, [RE + 1e6, math.pi / 4, 0.0, 1.0, 1.0, 1e-3, 0.0], 1000.0)
    assert rel(du[0], -9.628103804669998e-3) < 1e-12 and rel(du[6], 9.035737902855376e-11) < 1e-9
    # the file's own u0 has rho_phi = 0 -> 0/0 in ddrho_phi (RayTrace_3D.jl:390, :304-310)
    du = O.haselgrove_3d(E, [RE + 1e6, math.pi / 4, 0.0, 1.0, 1.0, 0.0, 0.0], 1000.0)
    assert np.isnan(du[2:5]).all() and du[5] == 0.0 and np.isfinite(du[[0, 1, 6]]).all()


@pytest.mark.parametrize("key,pid", [("lat2d", O.LAT2D), ("raymain", O.RAYMAIN)])
def test_golden_2d(key, pid):
    E = O.env(pid)
    for row in GOLD[key]:
        mu, dmu = O.phase_refractive_index_2d(E, *row["in"])
        assert rel(mu, row["out"]["mu"]) < 5e-13, row["in"]
        assert rel(dmu, row["out"]["dmudpsi"]) < 1e-9, row["in"]  # cancellation in dF/dpsi amplifies ulps


def test_golden_3d():
    E = O.env(O.RAY3D)
    for row in GOLD["ray3d"]:
        o = O.refractive_index_3d(E, *row["in"])
        assert rel(o[0], row["out"]["mu"]) < 5e-13 and rel(o[2], row["out"]["psi"]) < 1e-14
        assert rel(o[3], row["out"]["Br"]) < 1e-14 and rel(o[4], row["out"]["Bt"]) < 1e-14 and o[5] == 0.0
        b = O.magnetic_field(E, *row["in"][:3])
        assert b[0] == o[3] and b[1] == o[4]


def test_seasonal_parenthesisation_quirk():
    """RayTrace_3D.jl:109 divides cos(.) by 365 instead of its argument: 0.0510 vs 0.1276."""
    assert rel(GOLD["constants"]["seasonal_3d"], 0.0510054794520548) < 1e-14
    assert rel(GOLD["constants"]["seasonal_lat"], 0.12757508166977483) < 1e-14
    # the two files therefore give different densities on the same L-shell: compare mu at an inner point
    r, th = 2 * RE, 1.0
    m3 = O.refractive_index_3d(O.env(O.RAY3D), r, th, 0.0, 1.0, 1.0, 1.0, 2000.0)[0]
    assert np.isfinite(m3)


def test_identity_psi_half_pi():
    """scratch.jl:145-220 comments: at psi = pi/2, A = S, B = RL + PS, F = |RL - PS| and d(mu)/d(psi) = 0, hence
    mu^2_+ = max(RL, PS)/S (the comment's "RL/S" is the RL > PS case).  psi = pi/2 <=> chi = -dip."""
    E = O.env(O.LAT2D)
    lam, r, f = 0.6, 2.5 * RE, 4000.0
    dip = math.atan(2 * math.tan(lam))
    mu, dmu = O.phase_refractive_index_2d(E, r, lam, -dip, f)
    import tools_golden as TG
    ref = TG.pri_lat(r, lam, -dip, f)  # R, L, P from the independent restatement
    S = (ref["R"] + ref["L"]) / 2
    assert rel(mu ** 2, abs(max(ref["R"] * ref["L"], ref["P"] * S) / S)) < 1e-9
    assert abs(dmu) < 1e-6 * abs(mu)


def test_identity_psi_zero_gives_R():
    """RayTrace_lat.jl:172 / RayMain.jl:242: psi = 0 -> mu^2 = R for the '+' root (|.| taken as the file does)."""
    E = O.env(O.LAT2D)
    lam, r, f = 0.4, 3.0 * RE, 2500.0
    dip = math.atan(2 * math.tan(lam))
    chi = -math.pi / 2 - dip  # psi = 0
    mu, _ = O.phase_refractive_index_2d(E, r, lam, chi, f)
    import tools_golden as TG
    ref = TG.pri_lat(r, lam, chi, f)
    assert rel(mu ** 2, abs(ref["R"])) < 1e-6 or rel(mu ** 2, abs(ref["L"])) < 1e-6


def test_dipole_identity():
    """RayTrace_lat.jl:58-65: 4 - 3 sin^2(theta) == 1 + 3 cos^2(theta); |B| from components == closed form."""
    E = O.env(O.RAY3D)
    for th in np.linspace(0.2, 2.9, 13):
        b = O.magnetic_field(E, 2 * RE, th, 0.0)
        lam = math.pi / 2 - th
        closed = E.B0 * (RE / (2 * RE)) ** 3 * math.sqrt(1 + 3 * math.sin(lam) ** 2)
        assert rel(math.hypot(b[0], b[1]), closed) < 1e-14


def test_ddr_is_exactly_zero_by_absorption_not_fiat():
    E = O.env(O.LAT2D)
    assert O.dd_2d(E, "r", 3 * RE, 0.5, 0.2, 2000.0) == 0.0
    # the zero comes from r -+ 5e-12 rounding back to r (ulp(3 re) ~ 3.7e-9), not from a special case:
    r = 3 * RE
    assert r - 1.0e-11 / 2.0 == r and r + 1.0e-11 / 2.0 == r
    assert 1e-3 - 1.0e-11 / 2.0 != 1e-3  # ... which stops being true for |r| << 1 (slow path in the kernel)


def test_plasmasphere_jl_functions():
    L = O.lib()
    assert rel(L.orc_ionosphere_eq(O._d(1000000 + RE), O._d(RE)), 33334.97520765754) < 1e-13  # plasmasphere.jl:132
    assert L.orc_L_rlam(O._d(2 * RE), O._d(0.0), O._d(RE)) == 2.0
    Lppi, Lppo = 4.22, 4.387901808136615
    a = L.orc_plasmasphere_eq(O._d(4.22), O._d(Lppi), O._d(Lppo), O._d(0), O._d(90), O._d(2))
    assert rel(a, 403.7856649204069) < 1e-13  # continuity at Lppi: segment 1 == ne_Lppi
    b = L.orc_plasmasphere_eq(O._d(4.2200001), O._d(Lppi), O._d(Lppo), O._d(0), O._d(90), O._d(2))
    assert rel(b, a) < 1e-5
    # diffusive equilibrium: unity at the base rb, decreasing outward (plasmasphere.jl:96-106)
    assert L.orc_diffusive_equilibrium(O._d(7370000.0), O._d(100.0)) == 100.0
    assert L.orc_diffusive_equilibrium(O._d(2 * 7370000.0), O._d(100.0)) < 100.0


def test_composition_density_model_and_meridian_maps():
    """density_model = 1 (plasmasphere.jl:171) changes mu through the density only; the maps of
    plasmasphere.jl:157-175 agree with the scalar functions they are built from."""
    E0, E1 = O.env(O.LAT2D), O.env(O.LAT2D)
    E1.density_model = 1
    r, lam, chi, f = 2.5 * RE, 0.4, 0.2, 4000.0
    m0, m1 = O.phase_refractive_index_2d(E0, r, lam, chi, f), O.phase_refractive_index_2d(E1, r, lam, chi, f)
    assert m0[0] != m1[0] and 0.3 < m1[0] / m0[0] < 1.0  # diffusive equilibrium thins the plasma: smaller mu
    x, y = np.array([2 * RE, 3 * RE]), np.array([0.0, RE])
    M = O.meridian_maps(E0, x, y)
    assert M.shape == (4, 4) and M[0, 0] == 2.0 and M[3, 0] == E0.B0 / 2.0  # (2 re, 0): L = 2, B0 (1/2)^3 (1 + 3)
    L = O.lib()
    k = 3  # (3 re, re)
    rr, th = math.hypot(3 * RE, RE), math.atan(1.0 / 3.0)
    assert rel(M[0, k], (rr / RE) / math.cos(th) ** 2) < 1e-14
    ne_eq = L.orc_plasmasphere_eq(O._d(M[0, k]), O._d(E0.Lppi), O._d(E0.Lppo), O._d(E0.d), O._d(E0.Rbar), O._d(E0.mlt))
    want = L.orc_diffusive_equilibrium(O._d(rr), O._d(ne_eq)) + L.orc_ionosphere_eq(O._d(rr), O._d(RE))
    assert rel(M[1, k], want) < 1e-13 and rel(M[2, k], math.log10(want)) < 1e-13


def test_variant_presets_differ_as_the_files_do():
    r, th, chi, f = RE + 1e6, math.pi / 3, 0.3, 5000.0
    m_main = O.phase_refractive_index_2d(O.env(O.RAYMAIN), r, th, chi, f)
    m_scr = O.phase_refractive_index_2d(O.env(O.SCRATCH), r, th, chi, f)
    m_lat = O.phase_refractive_index_2d(O.env(O.LAT2D), r, math.pi / 2 - th, chi, f)
    assert not np.allclose(m_main, m_scr) and not np.allclose(m_main, m_lat)
    du_m = O.haselgrove_2d(O.env(O.RAYMAIN), [r, th, chi, f])
    du_s = O.haselgrove_2d(O.env(O.SCRATCH), [r, th, chi, f])
    assert np.isfinite(du_m).all() and np.isfinite(du_s).all()


# ---- the Haselgrove right-hand side against the independent restatement's goldens (tests/golden/haselgrove_rhs.json) ----
RHS_GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "haselgrove_rhs.json")))
EPS = 2.220446049250313e-16


def _fd_noise(mu, nulp=48):
    """|error| of a width-1e-11 central difference of mu when each mu carries `nulp`/2 ulp of libm disagreement."""
    return nulp * EPS * mu / 1.0e-11


@pytest.mark.parametrize("key,pid", [("lat2d", O.LAT2D), ("raymain", O.RAYMAIN)])
def test_golden_haselgrove_2d(key, pid):
    """du[1], du[2] (mu and dmu/dpsi only) at 1e-9 (dmu/dpsi's own cancellation), du[4] within the 1e-5-wide
    difference in f's rounding noise, du[3] within the 1e-11-wide difference's rounding noise (RayTrace_lat.jl:210-222,270-273; RayMain.jl:341-344)."""
    E = O.env(pid)
    assert len(RHS_GOLD[key]) >= 20
    for row in RHS_GOLD[key]:
        u, g, mu = row["u"], row["du"], row["mu"]
        du = O.haselgrove_2d(E, u)
        assert rel(du[0], g[0]) < 1e-9 and rel(du[1], g[1]) < 1e-9, u
        # (1 + (f/mu) dmu/df) / c: the 1e-5-wide difference turns nulp ulp of mu into f nulp eps / 1e-5
        assert abs(du[3] - g[3]) <= (u[3] * 48 * EPS / 1.0e-5) / 2.99792458e8 + 1e-9 * abs(g[3]), u
        bound = _fd_noise(mu) * abs(math.cos(u[2])) / (u[0] * mu * mu) + 1e-9 * abs(g[2])
        assert abs(du[2] - g[2]) <= bound, (u, du[2], g[2], bound)
        assert np.sign(du[0]) == np.sign(g[0]) and np.sign(du[1]) == np.sign(g[1])


def test_golden_haselgrove_3d():
    """du[1..4] and du[6] are free of noisy differences (Kimura-form d(mu)/d(rho_k); d(mu)/dr and d(mu)/dphi are
    exactly 0); du[5] carries the 1e-11-wide theta difference, du[7] the 1e-5-wide f difference
    (RayTrace_3D.jl:223-323,350-356)."""
    E = O.env(O.RAY3D)
    assert len(RHS_GOLD["ray3d"]) >= 20
    for row in RHS_GOLD["ray3d"]:
        u, f, g, mu = row["u"], row["freq"], row["du"], row["mu"]
        du = O.haselgrove_3d(E, u, f)
        for k in (0, 1, 2, 3, 5):
            assert rel(du[k], g[k]) < 1e-8, (k, u, du[k], g[k])
        assert abs(du[6] - g[6]) <= (f * 48 * EPS / 1.0e-5) / 2.99792458e8 + 1e-9 * abs(g[6]), u
        bound = _fd_noise(mu) / (u[0] * mu) + 1e-8 * abs(g[4])
        assert abs(du[4] - g[4]) <= bound, (u, du[4], g[4], bound)
