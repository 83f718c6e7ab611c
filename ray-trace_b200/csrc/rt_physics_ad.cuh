// rt_physics_ad.cuh — [EXT] the physically-correct option set of SURVEY.md §8 row f4, NEVER on the parity path.
//
// The reference differentiates mu numerically with steps FP64 cannot resolve (d(mu)/dr is identically 0, d(mu)/d(theta)
// is rounding noise amplified by 1e11: RayTrace_lat.jl:198-222, RayTrace_3D.jl:223-247) and its TODOs ask for analytic
// derivatives (RayMain.jl:97-107, RayTrace_3D.jl:22-27); it takes mu = sqrt(abs(mu^2_+)) with the comment "abs() is not
// physical! for test only" (RayTrace_lat.jl:157,174).  rt_env_params.deriv_mode = 1 evaluates d(mu)/dr and
// d(mu)/d(theta|lambda) exactly, by forward-mode differentiation of the same refractive-index code (dual numbers carrying
// the two partials through every operation: no step size, no cancellation); rt_env_params.root_mode = 1 selects the
// propagating root without abs — mu^2_+ when it is positive, else mu^2_- when that is positive, else NaN (evanescent) —
// and differentiates THAT root with the correct quotient rule (the formula at RayTrace_lat.jl:177-184 doubles the
// second term).  Both options route the integrate kernel through its out-of-line exact right-hand side.
#pragma once
#include "rt_physics.cuh"

struct RtDual {  // value and its partials with respect to (r, theta)
  double v, a, b;
};
#define RTD __device__ __forceinline__
RTD RtDual rd_const(double x) { return RtDual{x, 0.0, 0.0}; }
RTD RtDual operator+(RtDual x, RtDual y) { return RtDual{x.v + y.v, x.a + y.a, x.b + y.b}; }
RTD RtDual operator-(RtDual x, RtDual y) { return RtDual{x.v - y.v, x.a - y.a, x.b - y.b}; }
RTD RtDual operator-(RtDual x) { return RtDual{-x.v, -x.a, -x.b}; }
RTD RtDual operator*(RtDual x, RtDual y) { return RtDual{x.v * y.v, x.a * y.v + x.v * y.a, x.b * y.v + x.v * y.b}; }
RTD RtDual operator/(RtDual x, RtDual y) {
  const double q = x.v / y.v;
  return RtDual{q, (x.a - q * y.a) / y.v, (x.b - q * y.b) / y.v};
}
RTD RtDual operator+(RtDual x, double c) { return RtDual{x.v + c, x.a, x.b}; }
RTD RtDual operator+(double c, RtDual x) { return RtDual{c + x.v, x.a, x.b}; }
RTD RtDual operator-(RtDual x, double c) { return RtDual{x.v - c, x.a, x.b}; }
RTD RtDual operator-(double c, RtDual x) { return RtDual{c - x.v, -x.a, -x.b}; }
RTD RtDual operator*(RtDual x, double c) { return RtDual{x.v * c, x.a * c, x.b * c}; }
RTD RtDual operator*(double c, RtDual x) { return RtDual{c * x.v, c * x.a, c * x.b}; }
RTD RtDual operator/(RtDual x, double c) { return RtDual{x.v / c, x.a / c, x.b / c}; }
RTD RtDual operator/(double c, RtDual y) {
  const double q = c / y.v;
  return RtDual{q, -q * y.a / y.v, -q * y.b / y.v};
}
RTD RtDual rd_scale(RtDual x, double dv) { return RtDual{0.0, x.a * dv, x.b * dv}; }  // chain rule helper: f'(x) dx
RTD RtDual rd_sqrt(RtDual x) {
  const double s = RT_SQRT(x.v), h = 0.5 / s;
  return RtDual{s, x.a * h, x.b * h};
}
RTD RtDual rd_abs(RtDual x) { return x.v < 0.0 ? -x : x; }
RTD void rd_sincos(RtDual x, RtDual *s, RtDual *c) {
  double sv, cv;
  rt_sincos(x.v, &sv, &cv);
  *s = RtDual{sv, cv * x.a, cv * x.b};
  *c = RtDual{cv, -sv * x.a, -sv * x.b};
}
RTD RtDual rd_exp(RtDual x) {
  const double e = rt_exp(x.v);
  return RtDual{e, e * x.a, e * x.b};
}
RTD RtDual rd_exp10(RtDual x) {
  const double e = rt_exp10(x.v), k = e * 2.302585092994046;
  return RtDual{e, k * x.a, k * x.b};
}
RTD RtDual rd_acos(RtDual x) {
  const double k = -1.0 / RT_SQRT(1.0 - x.v * x.v);
  return RtDual{rt_acos(x.v), k * x.a, k * x.b};
}
RTD RtDual rd_atan(RtDual x) {
  const double k = 1.0 / (1.0 + x.v * x.v);
  return RtDual{rt_atan(x.v), k * x.a, k * x.b};
}

// CA92 + ionosphere (rt_density_cm3, RayTrace_lat.jl:69-83) in dual arithmetic; L^-4.5 as exp-free power rule
RTD RtDual rd_density_cm3(const RtDevEnv &E, RtDual r, RtDual L) {
  const RtDual ne_iono = E.iono_n0 * rd_exp(E.iono_k * ((r / E.re) - E.iono_h));
  RtDual ne_plasma = rd_const(0.0);
  if (L.v <= E.Lppi) {
    ne_plasma = rd_exp10((-0.3145 * L + 3.9043) + E.ca_coef * rd_exp((2.0 - L) / 1.5));
  } else if (L.v <= E.Lppo) {
    ne_plasma = E.ne_Lppi * rd_exp10((E.Lppi - L) / 0.1);
  } else {
    const double p = rt_pow_m4p5(L.v), dp = -4.5 * p / L.v;
    ne_plasma = E.ca_trough * RtDual{p, dp * L.a, dp * L.b} + (1.0 - rd_exp((2.0 - L) / 10.0));
  }
  if (E.density_model == 1) {  // plasmasphere.jl:96-117,171
    const double T = 2500.0, rb = 7370000.0, qq = rb / 7370.0;
    const RtDual G = rb * (1.0 - rb / r);
    const double S = ((1.506 * T) * (qq * qq)) * (1.0 / 1.0);
    return ne_plasma * rd_sqrt(rd_exp(-G / S)) + 1e5 * rd_exp(-10.0 * ((r / E.re) - 1.0471));
  }
  return ne_iono + ne_plasma;
}

struct RtMuAD {
  RtDual mu;       // mu with d/dr, d/dtheta
  double dmudpsi;  // d(mu)/d(psi) at fixed medium
  double psi, dip;
  double Br, Bt, Bmag, s, c;  // centre field and sin / cos(psi) (3D: Kimura-form d(mu)/d(rho_k))
};

// Stix block (RayTrace_lat.jl:86-184) on a dual medium; psi carries its own r / theta dependence through s, c
RTD void rd_stix(const RtDevEnv &E, RtDual n_e, RtDual Bmag, double omega, RtDual s, RtDual c, int root_mode, RtMuAD &o) {
  const RtDual w_e2 = (n_e * E.e2) / E.eps_me, w_p2 = (n_e * E.e2) / E.eps_mp;
  const RtDual W_e = (E.e * Bmag) / E.me, W_p = (E.e * Bmag) / E.mp;
  const double w2 = omega * omega;
  const RtDual xe = w_e2 / w2, xp = w_p2 / w2;
  const RtDual R = 1.0 - xe * (omega / (omega - W_e)) - xp * (omega / (omega + W_p));
  const RtDual L = 1.0 - xe * (omega / (omega + W_e)) - xp * (omega / (omega - W_p));
  const RtDual P = 1.0 - xe - xp;
  const RtDual D = (R - L) / 2.0, S = (R + L) / 2.0;
  const RtDual s2 = s * s, c2 = c * c;
  const RtDual A = S * s2 + P * c2;
  const RtDual B = (R * L) * s2 + (P * S) * (1.0 + c2);
  const RtDual X = R * L - P * S;
  const RtDual PDc = (P * D) * c;
  const RtDual F = rd_sqrt((X * X) * (s2 * s2) + 4.0 * (PDc * PDc));
  // d/dpsi at fixed medium (values only)
  const double dA = ((2.0 * (S.v - P.v)) * s.v) * c.v, dB = ((2.0 * X.v) * s.v) * c.v;
  const double PD = P.v * D.v;
  const double dF = (1.0 / (2.0 * F.v)) * ((((X.v * X.v) * 4.0) * (s2.v * s.v)) * c.v - ((8.0 * (PD * PD)) * s.v) * c.v);
  const RtDual m2p = (B + F) / (2.0 * A);
  if (root_mode == 0) {  // as written: sqrt(abs(mu^2_+)) and the file's d(mu)/d(psi) formula
    o.mu = rd_sqrt(rd_abs(m2p));
    o.dmudpsi = (1.0 / (2.0 * o.mu.v)) * ((dB + dF) / (2.0 * A.v) - ((2.0 * dA) * (B.v + F.v)) / (2.0 * (A.v * A.v)));
    return;
  }
  const RtDual m2m = (B - F) / (2.0 * A);
  const bool plus = m2p.v > 0.0, minus = !plus && m2m.v > 0.0;
  if (!plus && !minus) {  // evanescent: no propagating root
    o.mu = RtDual{rt_nan(), rt_nan(), rt_nan()};
    o.dmudpsi = rt_nan();
    return;
  }
  const RtDual m2 = plus ? m2p : m2m;
  const double sg = plus ? 1.0 : -1.0;
  o.mu = rd_sqrt(m2);
  // d(mu^2)/dpsi = (dB +- dF) / 2A - (B +- F) dA / 2A^2 ; d(mu) = d(mu^2) / 2 mu
  o.dmudpsi = (1.0 / (2.0 * o.mu.v)) * ((dB + sg * dF) / (2.0 * A.v) - (dA * (B.v + sg * F.v)) / (2.0 * (A.v * A.v)));
}

// phase_refractive_index(r, lambda, chi, freq) of RayTrace_lat.jl:44-194 with d/dr, d/dlambda
RTD void rd_mu2d_lat(const RtDevEnv &E, double r0, double lam0, double chi, double freq, RtMuAD &o) {
  const RtDual r{r0, 1.0, 0.0}, lam{lam0, 0.0, 1.0};
  RtDual st, ct;
  rd_sincos(lam, &st, &ct);
  const RtDual tn = st / ct;
  const RtDual dip = rd_atan(2.0 * tn);
  const RtDual psi = (RT_PI / 2.0 + dip) + chi;
  const RtDual r3 = (r * r) * r;
  const RtDual Bmag = (E.B0 * (E.re3 / r3)) * rd_sqrt(1.0 + (3.0 * st) * st);
  const RtDual Lsh = r / (E.re * (ct * ct));
  const RtDual n_e = rd_density_cm3(E, r, Lsh) * 1e6;
  RtDual s, c;
  rd_sincos(psi, &s, &c);
  rd_stix(E, n_e, Bmag, (2.0 * RT_PI) * freq, s, c, E.root_mode, o);
  o.psi = psi.v;
  o.dip = dip.v;
}

// refractive_index(r, theta, phi, rho, freq) of RayTrace_3D.jl:93-219 with d/dr, d/dtheta
RTD void rd_mu3d(const RtDevEnv &E, double r0, double th0, double rho_r, double rho_t, double rho_p, double freq, RtMuAD &o) {
  const RtDual r{r0, 1.0, 0.0}, th{th0, 0.0, 1.0};
  const RtDual lam = RT_PI / 2.0 - th;
  RtDual sl, cl;
  rd_sincos(lam, &sl, &cl);
  const RtDual q = E.re / r;
  const RtDual q3 = (q * q) * q;
  const RtDual Br = ((-2.0 * E.B0) * q3) * sl;
  const RtDual Bt = -((E.B0 * q3) * cl);
  const RtDual Bmag = rd_sqrt(Br * Br + Bt * Bt);
  const RtDual Lsh = r / (E.re * (cl * cl));
  const RtDual n_e = rd_density_cm3(E, r, Lsh) * 1e6;
  const double mumag = RT_SQRT((rho_r * rho_r + rho_t * rho_t) + rho_p * rho_p);
  const RtDual cos_psi = (Br * rho_r + Bt * rho_t) / (Bmag * mumag);
  const RtDual psi = rd_acos(cos_psi);
  RtDual s, c;
  rd_sincos(psi, &s, &c);
  rd_stix(E, n_e, Bmag, (2.0 * RT_PI) * freq, s, c, E.root_mode, o);
  o.psi = psi.v;
  o.Br = Br.v;
  o.Bt = Bt.v;
  o.Bmag = Bmag.v;
  o.s = s.v;
  o.c = c.v;
}

// mu only (values), for the f -/+ difference, honouring root_mode
RTD double rd_mu2d_lat_value(const RtDevEnv &E, double r, double lam, double chi, double freq) {
  RtMuAD o;
  rd_mu2d_lat(E, r, lam, chi, freq, o);
  return o.mu.v;
}
RTD double rd_mu3d_value(const RtDevEnv &E, double r, double th, double rr, double rt, double rp, double freq) {
  RtMuAD o;
  rd_mu3d(E, r, th, rr, rt, rp, freq, o);
  return o.mu.v;
}

// haselgrove!(du,u,p,t), RayTrace_lat.jl:251-273, with d(mu)/dr and d(mu)/dlambda from the dual evaluation when
// deriv_mode = 1 (else central differences with the environment's steps), d(mu)/df as written (1e-5-wide difference)
__device__ __noinline__ void rt_haselgrove_2d_lat_ext(const RtDevEnv &E, const double *u, double *du, RtExtras2D *ex) {
  const double r = u[0], lam = u[1], chi = u[2], freq = u[3];
  RtMuAD o;
  rd_mu2d_lat(E, r, lam, chi, freq, o);
  double dmudr = o.mu.a, dmudl = o.mu.b;
  if (E.deriv_mode == 0) {
    dmudl = (rd_mu2d_lat_value(E, r, lam + E.h_ang / 2.0, chi, freq) - rd_mu2d_lat_value(E, r, lam - E.h_ang / 2.0, chi, freq)) / E.h_ang;
    dmudr = (rd_mu2d_lat_value(E, r + E.h_r / 2.0, lam, chi, freq) - rd_mu2d_lat_value(E, r - E.h_r / 2.0, lam, chi, freq)) / E.h_r;
  }
  const double dmudf = (rd_mu2d_lat_value(E, r, lam, chi, freq + 1.0e-5 / 2.0) - rd_mu2d_lat_value(E, r, lam, chi, freq - 1.0e-5 / 2.0)) / 1.0e-5;
  double sc, cc;
  rt_sincos(chi, &sc, &cc);
  const double mu = o.mu.v, dmudpsi = o.dmudpsi;
  du[0] = (1.0 / (mu * mu)) * (mu * cc + dmudpsi * sc);
  du[1] = (1.0 / (r * (mu * mu))) * (mu * sc - dmudpsi * cc);
  du[2] = (1.0 / (r * (mu * mu))) * (dmudl * cc - (r * dmudr + mu) * sc);
  du[3] = (1.0 / E.c) * (1.0 + (freq / mu) * dmudf);
  if (ex) {
    ex->mu = mu;
    ex->dmudpsi = dmudpsi;
    ex->dip = o.dip;
    ex->psi = o.psi;
  }
}

// haselgrove!(du,u,p,t), RayTrace_3D.jl:327-356, same options
__device__ __noinline__ void rt_haselgrove_3d_ext(const RtDevEnv &E, const double *u, double freq, double *du) {
  const double r = u[0], th = u[1], rr_ = u[3], rt_ = u[4], rp_ = u[5];
  RtMuAD o;
  rd_mu3d(E, r, th, rr_, rt_, rp_, freq, o);
  double dmudr = o.mu.a, dmudt = o.mu.b;
  if (E.deriv_mode == 0) {
    dmudt = (rd_mu3d_value(E, r, th + E.h_ang / 2.0, rr_, rt_, rp_, freq) - rd_mu3d_value(E, r, th - E.h_ang / 2.0, rr_, rt_, rp_, freq)) / E.h_ang;
    dmudr = (rd_mu3d_value(E, r + E.h_r / 2.0, th, rr_, rt_, rp_, freq) - rd_mu3d_value(E, r - E.h_r / 2.0, th, rr_, rt_, rp_, freq)) / E.h_r;
  }
  const double dmudf = (rd_mu3d_value(E, r, th, rr_, rt_, rp_, freq + 1.0e-5 / 2.0) - rd_mu3d_value(E, r, th, rr_, rt_, rp_, freq - 1.0e-5 / 2.0)) / 1.0e-5;
  const double mu = o.mu.v;
  // Kimura-form d(mu)/d(rho_k), RayTrace_3D.jl:261-311 (B_phi = 0)
  const double den = (mu * mu) * o.s;
  const double ca0 = (o.Br * rr_) / (o.Bmag * rt_abs(rr_)), ca1 = (o.Bt * rt_) / (o.Bmag * rt_abs(rt_));
  const double ca2 = (0.0 * rp_) / (o.Bmag * rt_abs(rp_));
  const double dmr = o.dmudpsi * ((rr_ * o.c - mu * ca0) / den);
  const double dmt = o.dmudpsi * ((rt_ * o.c - mu * ca1) / den);
  const double dmp = o.dmudpsi * ((rp_ * o.c - mu * ca2) / den);
  double st, ct;
  rt_sincos(th, &st, &ct);
  du[0] = (1.0 / (mu * mu)) * (rr_ - mu * dmr);
  du[1] = (1.0 / (r * (mu * mu))) * (rt_ - mu * dmt);
  du[2] = (1.0 / ((r * (mu * mu)) * st)) * (rp_ - mu * dmp);
  du[3] = ((1.0 / mu) * dmudr + rt_ * du[1]) + (rp_ * du[2]) * st;
  du[4] = (1.0 / r) * (((1.0 / mu) * dmudt - rt_ * du[0]) + ((r * rp_) * du[2]) * ct);
  du[5] = (1.0 / (r * st)) * (((1.0 / mu) * 0.0 - (rp_ * du[0]) * st) - ((r * rp_) * du[1]) * ct);  // mu does not depend on phi
  du[6] = (1.0 / E.c) * (1.0 + (freq / mu) * dmudf);
}
