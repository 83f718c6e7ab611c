"""Generate tests/golden/*.json with an INDEPENDENT restatement of the reference formulas.

The reference is Julia and cannot run here (no julia binary), and it ships no golden vectors.
These fixtures therefore come from a second, separately written restatement of RayTrace_lat.jl:44-194 /
RayTrace_3D.jl:54-219 in plain Python floats + the C library's libm (math module), NOT from the oracle and NOT from
rt_math.h.  They pin the oracle at the libm-agreement level (a few ulp in mu; finite differences of width 1e-11 are
noise-dominated and carry loose tolerances).  Run:  python tools/make_golden.py
"""
import json
import math
import os

c = 2.99792458e8
re = 6.3712e6
e = 1.602e-19
me = 9.1093e-31
mp = 1.6726219e-27
eps = 8.854e-12
Kp_max, d, Rbar, mlt = 3, 0, 90, 2
Lppi = 5.6 - 0.46 * Kp_max
pi = math.pi


def seasonal_lat():
    return 0.15 * (math.cos((2 * pi * (d + 9)) / 365) - 0.5 * math.cos((4 * pi * (d + 9)) / 365)) + 0.00127 * Rbar - 0.0635


def seasonal_3d():
    return 0.15 * (math.cos((2 * pi * (d + 9))) / 365 - 0.5 * math.cos((4 * pi * (d + 9))) / 365) + 0.00127 * Rbar - 0.0635


ne_Lppi = 10 ** ((-0.3145 * Lppi + 3.9043) + seasonal_lat() * math.exp((2 - Lppi) / 1.5))


def presolve():
    best, Lb = None, None
    n = int((10 * re - re) // 1000) + 1
    for i in range(n):
        L = (re + 1000.0 * i) / re
        a = abs(ne_Lppi * 10 ** (-1.0 * (L - Lppi) / 0.1) - ((5800 + 300 * mlt) * L ** (-4.5) + (1 - math.exp((2 - L) / 10))))
        if best is None or a < best:
            best, Lb = a, L
    return Lb


Lppo = presolve()


def density(r, L, coef):
    iono = 1.8e5 * math.exp(-4.183119 * ((r / re) - 1.0471))
    if L <= Lppi:
        pl = 10 ** ((-0.3145 * L + 3.9043) + coef * math.exp((2 - L) / 1.5))
    elif L <= Lppo:
        pl = ne_Lppi * 10 ** ((Lppi - L) / 0.1)
    else:
        pl = (5800 + 300 * mlt) * L ** (-4.5) + (1 - math.exp((2 - L) / 10))
    return (iono + pl) * 1e6


def stix(n_e, Bmag, w, psi):
    we2 = n_e * e ** 2 / (eps * me)
    wp2 = n_e * e ** 2 / (eps * mp)
    We, Wp = e * Bmag / me, e * Bmag / mp
    R = 1 - (we2 / w ** 2) * (w / (w - We)) - (wp2 / w ** 2) * (w / (w + Wp))
    L = 1 - (we2 / w ** 2) * (w / (w + We)) - (wp2 / w ** 2) * (w / (w - Wp))
    P = 1 - we2 / w ** 2 - wp2 / w ** 2
    D, S = (R - L) / 2, (R + L) / 2
    s, co = math.sin(psi), math.cos(psi)
    A = S * s ** 2 + P * co ** 2
    B = R * L * s ** 2 + P * S * (1 + co ** 2)
    F = math.sqrt((R * L - P * S) ** 2 * s ** 4 + 4 * (P * D * co) ** 2)
    mu = math.sqrt(abs((B + F) / (2 * A)))
    dA = 2 * (S - P) * s * co
    dB = 2 * (R * L - P * S) * s * co
    dF = 1 / (2 * F) * ((R * L - P * S) ** 2 * 4 * s ** 3 * co - 8 * (P * D) ** 2 * s * co)
    dmu = 1 / (2 * mu) * ((dB + dF) / (2 * A) - 2 * dA * (B + F) / (2 * A ** 2))
    return dict(mu=mu, dmudpsi=dmu, R=R, L=L, P=P, A=A, B=B, F=F, mu2_plus=(B + F) / (2 * A))


def pri_lat(r, lam, chi, f, B0=3.0696381e-5):
    dip = math.atan(2 * math.tan(lam))
    psi = pi / 2 + dip + chi
    Bmag = B0 * (re ** 3 / r ** 3) * math.sqrt(1 + 3 * math.sin(lam) ** 2)
    L = r / (re * math.cos(lam) ** 2)
    out = stix(density(r, L, seasonal_lat()), Bmag, 2 * pi * f, psi)
    out.update(dip=dip, psi=psi, Lshell=L)
    return out


def pri_raymain(r, th, chi, f, B0=3.0696381e-5):
    dip = math.atan(2 / math.tan(th))
    psi = pi / 2 + dip + chi
    Bmag = B0 * (re ** 3 / r ** 3) * math.sqrt(1 + 3 * math.cos(th) ** 2)
    n_e = 1e6 * 1.8e5 * math.exp(-4.183119 * (r / re - 1.0471))
    return stix(n_e, Bmag, 2 * pi * f, psi)


def ri_3d(r, th, ph, rr, rt, rp, f, B0=3.12e-5):
    lam = pi / 2 - th
    Br = -2 * B0 * (re / r) ** 3 * math.sin(lam)
    Bt = -B0 * (re / r) ** 3 * math.cos(lam)
    Bmag = math.sqrt(Br ** 2 + Bt ** 2)
    L = r / (re * math.cos(pi / 2 - th) ** 2)
    psi = math.acos((Br * rr + Bt * rt) / (Bmag * math.sqrt(rr ** 2 + rt ** 2 + rp ** 2)))
    out = stix(density(r, L, seasonal_3d()), Bmag, 2 * pi * f, psi)
    out.update(psi=psi, Br=Br, Bt=Bt)
    return out


# ---- the Haselgrove right-hand sides, written from the .jl text (not from the oracle) ----------------------------
def rhs_lat(u):
    """haselgrove!(du,u,p,t), RayTrace_lat.jl:251-273, with ddr/ddλ/ddχ/ddf of :198-246 as written."""
    r, lam, chi, freq = u
    mu_of = lambda *a: pri_lat(*a)["mu"]
    h = 1.0e-11
    dmudr = (mu_of(r + h / 2.0, lam, chi, freq) - mu_of(r - h / 2.0, lam, chi, freq)) / h
    dmudlam = (mu_of(r, lam + h / 2.0, chi, freq) - mu_of(r, lam - h / 2.0, chi, freq)) / h
    df = 1.0e-5
    dmudf = (mu_of(r, lam, chi, freq + df / 2.0) - mu_of(r, lam, chi, freq - df / 2.0)) / df
    v = pri_lat(r, lam, chi, freq)
    mu, dmudpsi = v["mu"], v["dmudpsi"]
    return [1 / (mu ** 2) * (mu * math.cos(chi) + dmudpsi * math.sin(chi)),
            1 / (r * mu ** 2) * (mu * math.sin(chi) - dmudpsi * math.cos(chi)),
            1 / (r * mu ** 2) * (dmudlam * math.cos(chi) - (r * dmudr + mu) * math.sin(chi)),
            1 / c * (1 + (freq / mu) * dmudf)]


def rhs_raymain(u):
    """haselgrove!(du,u,p,t), RayMain.jl:321-344 (colatitude, opposite signs on the dmu/dpsi terms)."""
    r, th, chi, freq = u
    mu_of = lambda *a: pri_raymain(*a)["mu"]
    h = 1.0e-11
    dmudr = (mu_of(r + h / 2.0, th, chi, freq) - mu_of(r - h / 2.0, th, chi, freq)) / h
    dmudth = (mu_of(r, th + h / 2.0, chi, freq) - mu_of(r, th - h / 2.0, chi, freq)) / h
    df = 1.0e-5
    dmudf = (mu_of(r, th, chi, freq + df / 2.0) - mu_of(r, th, chi, freq - df / 2.0)) / df
    v = pri_raymain(r, th, chi, freq)
    mu, dmudpsi = v["mu"], v["dmudpsi"]
    return [1 / (mu ** 2) * (mu * math.cos(chi) - dmudpsi * math.sin(chi)),
            1 / (r * mu ** 2) * (mu * math.sin(chi) + dmudpsi * math.cos(chi)),
            1 / (r * mu ** 2) * (dmudth * math.cos(chi) - (r * dmudr + mu) * math.sin(chi)),
            1 / c * (1 + (freq / mu) * dmudf)]


def rhs_3d(u, freq):
    """haselgrove!(du,u,p,t), RayTrace_3D.jl:327-356 with ddr/ddθ/ddϕ/ddρr/ddρλ/ddρϕ/ddf of :223-323."""
    r, th, ph, rr, rt, rp, _T = u
    mu_of = lambda *a: ri_3d(*a)["mu"]
    h = 1.0e-11
    dmudr = (mu_of(r + h / 2.0, th, ph, rr, rt, rp, freq) - mu_of(r - h / 2.0, th, ph, rr, rt, rp, freq)) / h
    dmudth = (mu_of(r, th + h / 2.0, ph, rr, rt, rp, freq) - mu_of(r, th - h / 2.0, ph, rr, rt, rp, freq)) / h
    dmudph = (mu_of(r, th, ph + h / 2.0, rr, rt, rp, freq) - mu_of(r, th, ph - h / 2.0, rr, rt, rp, freq)) / h
    df = 1.0e-5
    dmudf = (mu_of(r, th, ph, rr, rt, rp, freq + df / 2.0) - mu_of(r, th, ph, rr, rt, rp, freq - df / 2.0)) / df
    v = ri_3d(r, th, ph, rr, rt, rp, freq)
    mu, dmudpsi, psi = v["mu"], v["dmudpsi"], v["psi"]
    Bvec = (v["Br"], v["Bt"], 0.0)
    Bmag = math.sqrt(Bvec[0] ** 2 + Bvec[1] ** 2 + Bvec[2] ** 2)

    def ddrho(k, rho_k):  # Kimura 1966 form, :261-311 (0/0 for rho_k = 0 is a NaN in Julia: callers avoid it)
        cos_a = (Bvec[k] * rho_k) / (Bmag * abs(rho_k))
        return dmudpsi * ((rho_k * math.cos(psi) - mu * cos_a) / (mu ** 2 * math.sin(psi)))

    du = [0.0] * 7
    du[0] = 1 / (mu ** 2) * (rr - mu * ddrho(0, rr))
    du[1] = 1 / (r * mu ** 2) * (rt - mu * ddrho(1, rt))
    du[2] = 1 / (r * mu ** 2 * math.sin(th)) * (rp - mu * ddrho(2, rp))
    du[3] = (1 / mu) * dmudr + rt * du[1] + rp * du[2] * math.sin(th)
    du[4] = (1 / r) * ((1 / mu) * dmudth - rt * du[0] + r * rp * du[2] * math.cos(th))
    du[5] = (1 / (r * math.sin(th))) * ((1 / mu) * dmudph - rp * du[0] * math.sin(th) - r * rp * du[1] * math.cos(th))
    du[6] = (1 / c) * (1 + (freq / mu) * dmudf)
    return du


def rhs_points(seed, n, dim3):
    """Deterministic pseudo-random evaluation points (a small LCG: no numpy, no platform RNG)."""
    x = seed

    def rnd():
        nonlocal x
        x = (6364136223846793005 * x + 1442695040888963407) % (1 << 64)
        return (x >> 11) / float(1 << 53)

    pts = []
    for _ in range(n):
        r = re * (1.1 + 5.0 * rnd())
        lat = math.radians(8.0 + 62.0 * rnd())
        f = math.exp(math.log(500.0) + rnd() * math.log(40.0))
        if dim3:
            chi = math.radians(-75.0 + 150.0 * rnd())
            al = 0.05 + 1.0 * rnd()
            sc = 0.5 + 60.0 * rnd()
            pts.append([r, pi / 2 - lat, 6.0 * rnd(), sc * math.cos(chi), sc * math.sin(chi) * math.cos(al),
                        sc * math.sin(chi) * math.sin(al), 0.0, f])
        else:
            pts.append([r, lat, -3.0 + 6.0 * rnd(), f])
    return pts


def make_rhs_golden():
    G = {"note": "du = haselgrove!(u) from the independent python+libm restatement (tools/make_golden.py). Components "
                 "built on the 1e-11 central differences in lambda/theta are rounding noise amplified by 1e11 "
                 "(SURVEY.md 0.4): compare those with the ulp-of-mu bound the test states, the others at 1e-12 / 1e-6."}
    G["lat2d"] = [{"u": p, "du": rhs_lat(p), "mu": pri_lat(*p)["mu"]} for p in rhs_points(11, 24, False)]
    G["lat2d"].insert(0, {"u": [re + 1e6, pi / 4, 0.0, 1000.0], "du": rhs_lat([re + 1e6, pi / 4, 0.0, 1000.0]),
                          "mu": pri_lat(re + 1e6, pi / 4, 0.0, 1000.0)["mu"]})
    G["raymain"] = []
    for p in rhs_points(12, 24, False):
        q = [p[0] if p[0] < 2.5 * re else re * (1.05 + (p[0] / re) / 8.0), pi / 2 - p[1], p[2], p[3]]  # ionosphere-only density
        G["raymain"].append({"u": q, "du": rhs_raymain(q), "mu": pri_raymain(*q)["mu"]})
    G["ray3d"] = [{"u": p[:7], "freq": p[7], "du": rhs_3d(p[:7], p[7]), "mu": ri_3d(*p[:6], p[7])["mu"]}
                  for p in rhs_points(13, 24, True)]
    p0 = [re + 1e6, pi / 4, 0.0, 1.0, 1.0, 1e-3, 0.0]
    G["ray3d"].insert(0, {"u": p0, "freq": 1000.0, "du": rhs_3d(p0, 1000.0), "mu": ri_3d(*p0[:6], 1000.0)["mu"]})
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "haselgrove_rhs.json")
    with open(out, "w") as fh:
        json.dump(G, fh, indent=1)
    print("wrote", out)


def main():
    pts2 = [(re + 1e6, pi / 4, 0.0, 1000.0), (re + 1e6, pi / 4, 0.0, 5000.0), (2 * re, 0.3, 0.5, 3000.0),
            (3.5 * re, 0.9, -1.2, 12000.0), (4.3 * re * math.cos(0.2) ** 2, 0.2, 2.0, 800.0), (6 * re, 0.1, -2.5, 20000.0),
            (1.2 * re, 1.2, 3.0, 30000.0), (5 * re, 0.5, 0.1, 1500.0)]
    pts3 = [(re + 1e6, pi / 4, 0.0, 1.0, 1.0, 1e-3, 1000.0), (2 * re, 1.0, 1.0, 0.5, -0.7, 0.2, 3000.0),
            (4.5 * re, 1.4, 2.0, -30.0, 12.0, 5.0, 700.0), (1.3 * re, 0.6, 4.0, 90.0, 40.0, -10.0, 15000.0),
            (6.5 * re, 1.5, 5.5, 3.0, 3.0, 3.0, 2500.0)]
    G = {
        "note": "independent python+libm restatement; see tools/make_golden.py",
        "constants": {"Lppi": Lppi, "Lppo": Lppo, "ne_Lppi": ne_Lppi, "seasonal_lat": seasonal_lat(), "seasonal_3d": seasonal_3d()},
        "lat2d": [{"in": list(p), "out": pri_lat(*p)} for p in pts2],
        "raymain": [{"in": [p[0], pi / 2 - p[1], p[2], p[3]], "out": pri_raymain(p[0], pi / 2 - p[1], p[2], p[3])} for p in pts2],
        "ray3d": [{"in": list(p), "out": ri_3d(*p)} for p in pts3],
    }
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "refractive_index.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w") as fh:
        json.dump(G, fh, indent=1)
    print("wrote", out)


# ---- whole rays: the independent right-hand sides under an integrator that shares nothing with the oracle's ---------
def integrate_independent(kind, u0, freq=None, rtol=1e-8, atol=None, t1=5.0e9):
    """scipy.integrate.solve_ivp (RK45 = Dormand-Prince 5(4): another tableau, another controller, another event
    root-finder than the oracle's Tsit5 restatement) over rhs_lat / rhs_3d, with the reference's terminal event
    u[1] - re (RayTrace_lat.jl:303-315).  Returns dict(t, y, event)."""
    from scipy.integrate import solve_ivp

    def ev(t, y):
        return y[0] - re

    ev.terminal = True
    if kind == "lat2d":
        f = lambda t, y: rhs_lat(list(y))
        atol = atol or [1e-3, 1e-12, 1e-12, 1e-12]
    else:
        f = lambda t, y: rhs_3d(list(y), freq)
        atol = atol or 1e-10
    s = solve_ivp(f, (0.0, t1), u0, method="RK45", rtol=rtol, atol=atol, events=ev)
    if len(s.t_events[0]):
        return {"t": float(s.t_events[0][0]), "y": [float(v) for v in s.y_events[0][0]], "event": True}
    return {"t": float(s.t[-1]), "y": [float(v) for v in s.y[:, -1]], "event": False}


def make_integration_golden():
    """cfg1 (the reference's own ray) and ten rays of the cfg3 grid, two of them long magnetospheric paths
    (> 900 attempts at the bench's tolerance).  About two minutes of pure-Python right-hand sides."""
    alpha = 0.1
    rays = []
    # (colatitude index, longitude index, chi index) on the 100 x 100 x 100 grid of sweeps.sweep_3d
    picks = [(0, 0, 0), (12, 34, 57), (24, 69, 14), (37, 3, 71), (49, 38, 28), (61, 72, 85), (74, 7, 42), (86, 41, 99),
             (72, 92, 0), (73, 13, 4)]
    for it, ip, ic in picks:
        th = math.radians(25.0 + (70.0 - 25.0) * it / 99.0)
        ph = 2 * pi * ip / 100.0
        chi = math.radians(-60.0 + 120.0 * ic / 99.0)
        u0 = [re + 1.0e6, th, ph, math.cos(chi), math.sin(chi) * math.cos(alpha), math.sin(chi) * math.sin(alpha), 0.0]
        rays.append({"grid_index": [it, ip, ic], "u0": u0, "freq": 1000.0, "out": integrate_independent("ray3d", u0, 1000.0)})
        print("3d", it, ip, ic, rays[-1]["out"]["t"], flush=True)
    u0 = [re + 1.0e6, pi / 4, 0.0, 1000.0]
    G = {"note": "scipy RK45 (rtol 1e-8) over the independent python right-hand sides; see tools/make_golden.py",
         "rtol": 1e-8, "cfg1": {"u0": u0, "out": integrate_independent("lat2d", u0)}, "cfg3": rays}
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "integration_rk45.json")
    with open(out, "w") as fh:
        json.dump(G, fh, indent=1)
    print("wrote", out)


if __name__ == "__main__":
    import sys
    main()
    make_rhs_golden()
    if "--integrate" in sys.argv:
        make_integration_golden()
