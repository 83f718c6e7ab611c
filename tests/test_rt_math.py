"""Accuracy of the shared deterministic math header (host build, through the oracle library) against mpmath."""
import mpmath as mp
import numpy as np
import pytest

import oracle_lib as O

mp.mp.prec = 200
RNG = np.random.default_rng(7)
N = 4000

CASES = [
    ("exp", lambda: RNG.uniform(-700, 700, N), mp.exp, 1.0),
    ("exp", lambda: RNG.uniform(-3, 3, N), mp.exp, 1.0),
    ("exp10", lambda: RNG.uniform(-300, 300, N), lambda v: mp.mpf(10) ** v, 1.05),
    ("exp10", lambda: RNG.uniform(-2, 6, N), lambda v: mp.mpf(10) ** v, 1.05),
    ("log", lambda: np.exp(RNG.uniform(-700, 700, N)), mp.log, 1.0),
    ("log", lambda: RNG.uniform(0.5, 2.0, N), mp.log, 1.0),
    ("log10", lambda: np.exp(RNG.uniform(-30, 30, N)), mp.log10, 2.0),
    ("sin", lambda: RNG.uniform(-10, 10, N), mp.sin, 1.0),
    ("cos", lambda: RNG.uniform(-10, 10, N), mp.cos, 1.0),
    ("sin", lambda: RNG.uniform(-1e6, 1e6, N), mp.sin, 1.0),
    ("cos", lambda: RNG.uniform(-1e12, 1e12, N), mp.cos, 1.0),
    ("tan", lambda: RNG.uniform(-1.5, 1.5, N), mp.tan, 2.0),
    ("tan", lambda: RNG.uniform(-1e6, 1e6, N), mp.tan, 2.0),
    # next to the poles, where tan = sin / cos divides by a cancelled cosine (the reduction's tail keeps it accurate)
    ("tan", lambda: (RNG.integers(-1000, 1000, N) + 0.5) * np.pi + RNG.uniform(-1e-6, 1e-6, N), mp.tan, 2.5),
    ("atan", lambda: RNG.uniform(-5, 5, N), mp.atan, 1.0),
    ("atan", lambda: np.exp(RNG.uniform(-30, 30, N)), mp.atan, 1.0),
    ("acos", lambda: RNG.uniform(-1, 1, N), mp.acos, 1.0),
    ("acos", lambda: 1 - np.exp(RNG.uniform(-30, 0, N)), mp.acos, 1.0),
    ("pow_m4p5", lambda: RNG.uniform(1.0, 12.0, N), lambda v: v ** mp.mpf(-4.5), 4.0),
    # the step controller's EEst^0.14 and qold^0.08 (rt_pow_pos = exp(p log x)): the documented bound is
    # 1 + |p log x| ulp; error estimates between e^-12 and e^6 are what a solve at reltol 1e-3 .. 1e-9 produces,
    # e^-60 .. e^30 is far outside (the factor is clamped to [1/5, 10] there anyway)
    ("pow_b1", lambda: np.exp(RNG.uniform(-12, 6, N)), lambda v: mp.power(v, mp.mpf(0.14)), 1.0 + 0.14 * 12),
    ("pow_b2", lambda: np.exp(RNG.uniform(-12, 6, N)), lambda v: mp.power(v, mp.mpf(0.08)), 1.0 + 0.08 * 12),
    ("pow_b1", lambda: np.exp(RNG.uniform(-60, 30, N)), lambda v: mp.power(v, mp.mpf(0.14)), 1.0 + 0.14 * 60),
    ("pow_b2", lambda: np.exp(RNG.uniform(-60, 30, N)), lambda v: mp.power(v, mp.mpf(0.08)), 1.0 + 0.08 * 60),
]


@pytest.mark.parametrize("name,gen,ref,bound", CASES)
def test_ulp_error(name, gen, ref, bound):
    x = gen()
    y = O.math(name, x)
    worst = 0.0
    for xi, yi in zip(x, y):
        t = ref(mp.mpf(float(xi)))
        if t == 0:
            continue
        u = float(np.spacing(abs(float(t))))
        worst = max(worst, float(abs((mp.mpf(float(yi)) - t) / u)))
    assert worst <= bound, f"{name}: {worst} ulp"


def test_special_values():
    inf, nan = np.inf, np.nan
    assert np.array_equal(O.math("exp", [0.0, -inf, inf, 710.0, -746.0]), [1.0, 0.0, inf, inf, 0.0])
    assert np.isnan(O.math("exp", [nan])[0]) and np.isnan(O.math("exp10", [nan])[0])
    assert np.array_equal(O.math("exp10", [0.0, 1.0, 2.0, 3.0, -1.0, 22.0, 309.0, -400.0]),
                          [1.0, 10.0, 100.0, 1000.0, 0.1, 1e22, inf, 0.0])
    assert np.array_equal(O.math("log", [1.0, 0.0, inf]), [0.0, -inf, inf]) and np.isnan(O.math("log", [-1.0])[0])
    assert np.isnan(O.math("sin", [inf, nan, 2.0 ** 45])).all()  # documented domain: |x| < 2^45
    assert np.isfinite(O.math("cos", [2.0 ** 45 - 1.0])).all()
    assert np.array_equal(O.math("acos", [1.0, -1.0]), [0.0, np.pi]) and np.isnan(O.math("acos", [1.0000001])[0])
    assert O.math("atan", [inf])[0] == np.pi / 2 and O.math("atan", [-inf])[0] == -np.pi / 2
    # subnormal results and arguments
    assert O.math("exp", [-745.0])[0] == 5e-324 and O.math("log", [5e-324])[0] == pytest.approx(-744.4400719213812, rel=1e-15)
