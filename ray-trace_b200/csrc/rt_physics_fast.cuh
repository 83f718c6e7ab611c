// rt_physics_fast.cuh — the right-hand sides of rt_physics.cuh as ONE branch-free instruction stream.
//
// Same arithmetic, same order, same bits as rt_haselgrove_3d / rt_haselgrove_2d<LAT2D> (which follow
// RayTrace_3D.jl:93-359 and RayTrace_lat.jl:44-296) whenever the returned flag is true; the caller re-evaluates
// with the exact code when it is false (non-finite trial stages, poles, zero numerators, out-of-range arguments).
// What differs from the exact code is scheduling only:
//   * division, sqrt and the elementary functions are the guard-predicated sequences of rt_fast.cuh, so the
//     five refractive-index evaluations of one right-hand side form independent chains the compiler interleaves;
//   * quantities that depend on r alone (dipole (re/r)^3, ionosphere term) or on rho alone (|rho|) are evaluated
//     once instead of at each of the three theta points — same operands, same operations, same bits;
//   * divisions with a common divisor share one reciprocal refinement; divisions by constants use reciprocals
//     prepared per environment (RtDevEnv::y_*);
//   * terms that are identically +-0 by construction (B_phi = 0; (mu - mu)/h for the r and phi differences)
//     are not evaluated; the flag guards the operand ranges under which that is exact.
#pragma once
#include "rt_fast.cuh"
#include "rt_physics.cuh"

#ifndef RT_ROLL
#define RT_ROLL 1
#endif
#if RT_ROLL
#define RT_PAIR_LOOP _Pragma("unroll 1")
#else
#define RT_PAIR_LOOP _Pragma("unroll")
#endif
// the f -/+ pair alone (two Stix evaluations on the same medium, few live values) can be unrolled separately
#ifndef RT_ROLL_F
#define RT_ROLL_F RT_ROLL
#endif
#if RT_ROLL_F
#define RT_PAIR_LOOP_F _Pragma("unroll 1")
#else
#define RT_PAIR_LOOP_F _Pragma("unroll")
#endif

// exponent field of x in [0x200, 0x5ff]: products of two such numbers neither vanish nor become NaN
RTF bool rtf_quarter_range(double x) {
  return (unsigned)((__double2hiint(x) & 0x7fffffff) - 0x20000000) < 0x40000000u;
}

// a / b for a positive constant b where a may be exactly zero (then the quotient is a itself)
RTF double rtf_div_c0(double a, double b, double y, bool &ok) {
  bool okq = true;
  double q = rtf_div_c(a, b, y, okq);
  const bool z = (a == 0.0);
  ok = ok && (okq || z);
  return z ? a : q;
}

// rt_density_cm3 without the ionosphere term's exp (hoisted: depends on r only)
// ovf: the plasmasphere segment's 10^x overflows to +Inf for certain (x above rt_exp10's threshold, every input of
// x computed exactly) -> the density is +Inf and the caller knows the right-hand side without the exact path
RTF double rtf_density_cm3(const RtDevEnv &E, double ne_iono, double L, bool &ok, bool &ovf) {
  const bool seg1 = L <= E.Lppi;
  const bool seg2 = E.Lppi < L && L <= E.Lppo;
  const bool seg3 = E.Lppo < L;
  const double t2 = 2.0 - L;
  const double dsel = seg1 ? 1.5 : 10.0, ysel = seg1 ? E.y_1p5 : E.y_10;
  bool okd = true, okx = true;
  double arg_e = rtf_div_c(t2, dsel, ysel, okd);
  arg_e = seg2 ? 0.0 : arg_e;  // unused by segment 2
  double ex = rtf_exp(arg_e, okx);
  // exp's argument beyond the branch-free range, where the exact value is not needed:
  //  trough, arg <= -700 (L > 7002, near the poles): 0 <= exp < 2^-1000 and its only use is 1.0 - exp == 1.0;
  //  plasmasphere, arg >= 700 (L < -1048): exp >= 1e304 or +Inf, so with ca_coef >= 1e-200 the exponent of 10
  //  is >= 1e104 (its other term is positive for negative L) and 10^x = +Inf for certain
  // (integer tests on the high word: FP64 compares cost pipe slots)
  const int arg_hi = __double2hiint(arg_e);  // 700.0 = 0x4085E000'00000000; arg_e is finite when okd
  const bool tiny = seg3 && okd && ((unsigned)arg_hi >= 0xC085E000u);
  const bool huge = seg1 && okd && (arg_hi >= 0x4085E000) && (E.ca_coef >= 1e-200);
  ex = tiny ? 0.0 : ex;
  ok = ok && ((okd && (okx || tiny || huge)) || seg2);
  double a1 = (-0.3145 * L + 3.9043) + E.ca_coef * ex;
  bool ok2 = true;
  double a2 = rtf_div_c(E.Lppi - L, 0.1, E.y_0p1, ok2);
  ok = ok && (ok2 || !seg2);
  double arg10 = seg1 ? a1 : (seg2 ? a2 : 0.0);
  // rt_exp10 returns +Inf above 308.2547...; 308.5 has a round high word (arg10 is finite or +Inf here)
  ovf = ok && seg1 && (huge || __double2hiint(arg10) >= 0x40734800);
  double p10 = rtf_exp10(arg10, ok);
  // (skipping L^-4.5 and acos' x >= 0.5 correction by a warp vote when no lane needs them was measured: 72 FP64
  // instructions fewer per right-hand side, no gain — the branches cut the blocks the scheduler interleaves over)
  bool ok3 = true;
  double pw = rtf_pow_m4p5(seg3 ? L : 1.0, ok3);
  ok = ok && (ok3 || !seg3);
  double ne_plasma = 0.0;
  ne_plasma = seg1 ? p10 : ne_plasma;
  ne_plasma = seg2 ? E.ne_Lppi * p10 : ne_plasma;
  ne_plasma = seg3 ? E.ca_trough * pw + (1.0 - ex) : ne_plasma;
  return ne_iono + ne_plasma;
}

RTF void rtf_plasma_from(const RtDevEnv &E, double n_e, double Bmag, RtPlasma &p, bool &ok) {
  const double ne2 = n_e * E.e2, eB = E.e * Bmag;
  p.w_e2 = rtf_div_c(ne2, E.eps_me, E.y_eps_me, ok);
  p.w_p2 = rtf_div_c(ne2, E.eps_mp, E.y_eps_mp, ok);
  p.W_e = rtf_div_c(eB, E.me, E.y_me, ok);
  p.W_p = rtf_div_c(eB, E.mp, E.y_mp, ok);
}

// rt_stix with shared reciprocals
// (w2 = omega omega and yw = rtf_rcp(w2) are passed in: the three theta points of one right-hand side share them)
template <bool WITH_D>
RTF RtMu rtf_stix(const RtPlasma &p, double omega, double w2, double yw, bool &ok) {
  const double xe = rtf_div_y(p.w_e2, w2, yw, ok), xp = rtf_div_y(p.w_p2, w2, yw, ok);
  const double R = 1.0 - xe * rtf_div(omega, omega - p.W_e, ok) - xp * rtf_div(omega, omega + p.W_p, ok);
  const double L = 1.0 - xe * rtf_div(omega, omega + p.W_e, ok) - xp * rtf_div(omega, omega - p.W_p, ok);
  const double P = 1.0 - xe - xp;
  const double D = (R - L) * 0.5;  // "/ 2.0": exact either way
  const double S = (R + L) * 0.5;
  const double s = p.s, c = p.c;
  const double s2 = s * s, c2 = c * c;
  const double A = S * s2 + P * c2;
  const double B = (R * L) * s2 + (P * S) * (1.0 + c2);
  const double X = R * L - P * S;
  const double PD = P * D;
  const double PDc = PD * c;
  const double F2 = (X * X) * (s2 * s2) + 4.0 * (PDc * PDc);
  const double F = rtf_sqrt(F2, ok);
  const double A2 = 2.0 * A;
  const double yA = rtf_rcp(A2);
  const double mu2 = rtf_div_y(B + F, A2, yA, ok);
  RtMu out;
  out.mu = rtf_sqrt(rt_abs(mu2), ok);
  out.dmudpsi = 0.0;
  if (WITH_D) {
    const double dA = ((2.0 * (S - P)) * s) * c;
    const double dB = ((2.0 * X) * s) * c;
    const double dF = rtf_inv(2.0 * F, ok) * ((((X * X) * 4.0) * (s2 * s)) * c - ((8.0 * (PD * PD)) * s) * c);
    bool ok1 = true;
    double t1 = rtf_div_y(dB + dF, A2, yA, ok1);
    const bool z1 = (dB + dF) == 0.0;  // psi = 0, pi/2, pi: the reference's own check values (scratch.jl:212-220)
    t1 = z1 ? (dB + dF) * yA : t1;     // +-0 with the quotient's sign
    ok = ok && (ok1 || z1);
    bool ok2 = true;
    const double n2 = (2.0 * dA) * (B + F);
    double t2 = rtf_div(n2, 2.0 * (A * A), ok2);
    const bool z2 = (n2 == 0.0);
    t2 = z2 ? n2 : t2;  // divisor 2 A^2 > 0
    ok = ok && (ok2 || z2);
    out.dmudpsi = rtf_inv(2.0 * out.mu, ok) * (t1 - t2);
  }
  return out;
}

template <bool WITH_D>
RTF RtMu rtf_stix(const RtPlasma &p, double omega, bool &ok) {
  const double w2 = omega * omega;
  return rtf_stix<WITH_D>(p, omega, w2, rtf_rcp(w2), ok);
}

// =====================================================================================================
// 3D
// =====================================================================================================
struct RtRad3D {  // shared by the three theta points of one right-hand side
  double bq2, bq;  // (-2 B0) (re/r)^3, B0 (re/r)^3
  double ne_iono, mumag, yr;
  bool iono_ovf;
};

RTF void rtf_rad3d(const RtDevEnv &E, double r, double rho_r, double rho_t, double rho_p, RtRad3D &R, bool &ok) {
  R.yr = rtf_rcp(r);
  const double q = rtf_div_y(E.re, r, R.yr, ok);
  const double q3 = (q * q) * q;
  R.bq2 = (-2.0 * E.B0) * q3;
  R.bq = E.B0 * q3;
  const double rE = rtf_div_c(r, E.re, E.y_re, ok);
  const double arg_i = E.iono_k * (rE - E.iono_h);
  // far below the centre (r < -1e9 m) the ionosphere's exp overflows: rt_exp returns +Inf above 709.78 (712 has a
  // round high word), the density is +-Inf or NaN whatever the plasmasphere term is, and mu is NaN (see below)
  R.iono_ovf = ok && (__double2hiint(arg_i) >= 0x40864000);
  R.ne_iono = E.iono_n0 * rtf_exp(arg_i, ok);
  R.mumag = rtf_sqrt((rho_r * rho_r + rho_t * rho_t) + rho_p * rho_p, ok);
}

// rt_geo3d at colatitude th (B_phi = 0: its products are +-0 and drop out for finite rho_phi)
RTF void rtf_geo3d(const RtDevEnv &E, const RtRad3D &R, double r, double th, double rho_r, double rho_t, RtGeo3D &g,
                   bool &ok, bool &ovf) {
  const double lam = RT_PI / 2.0 - th;
  double sl, cl;
  rtf_sincos(lam, &sl, &cl, ok);
  g.Br = R.bq2 * sl;
  const double Bl = R.bq * cl;
  g.Bt = -Bl;
  g.Bp = 0.0;
  g.Bmag = rtf_sqrt(g.Br * g.Br + g.Bt * g.Bt, ok);
  const double L = rtf_div(r, E.re * (cl * cl), ok);
  const double n_e = rtf_density_cm3(E, R.ne_iono, L, ok, ovf) * 1e6;
  const double dot = g.Br * rho_r + g.Bt * rho_t;
  const double cos_psi = rtf_div(dot, g.Bmag * R.mumag, ok);
  g.psi = rtf_acos(cos_psi, ok);
  rtf_plasma_from(E, n_e, g.Bmag, g.p, ok);
  rtf_sincos(g.psi, &g.p.s, &g.p.c, ok);
}

// the Haselgrove combination, RayTrace_3D.jl:261-311,350-356, given the five refractive indices
RTF void rtf_tail3d(const RtDevEnv &E, const RtRad3D &Rd, const double *u, double freq, const RtGeo3D &gc, double mu,
                    double dmudpsi, double mu_tl, double mu_tr, double mu_fl, double mu_fr, double *du, bool &ok) {
  const double r = u[0], th = u[1], rr_ = u[3], rt_ = u[4], rp_ = u[5];
  const double dmudt = rtf_div_c0(mu_tr - mu_tl, 1.0e-11, E.y_dth, ok);
  const double dmudf = rtf_div_c0(mu_fr - mu_fl, 1.0e-5, E.y_df, ok);
  // d(mu)/d(rho_k): common divisor (mu mu) sin(psi)
  const double mm = mu * mu;
  const double den = mm * gc.p.s;
  const double yden = rtf_rcp(den);
  const double cos_a0 = rtf_div(gc.Br * rr_, gc.Bmag * rt_abs(rr_), ok);
  const double cos_a1 = rtf_div(gc.Bt * rt_, gc.Bmag * rt_abs(rt_), ok);
  const double dmr = dmudpsi * rtf_div_y(rr_ * gc.p.c - mu * cos_a0, den, yden, ok);
  const double dmt = dmudpsi * rtf_div_y(rt_ * gc.p.c - mu * cos_a1, den, yden, ok);
  const double dmp = dmudpsi * rtf_div_y(rp_ * gc.p.c, den, yden, ok);  // cos(alpha_phi) = +-0
  double st, ct;
  rtf_sincos(th, &st, &ct, ok);
  const double rmm = r * mm;
  const double ymu = rtf_rcp(mu);
  const double imu = rtf_inv_y(mu, ymu, ok);
  du[0] = rtf_inv(mm, ok) * (rr_ - mu * dmr);
  du[1] = rtf_inv(rmm, ok) * (rt_ - mu * dmt);
  du[2] = rtf_inv(rmm * st, ok) * (rp_ - mu * dmp);
  du[3] = (0.0 + rt_ * du[1]) + (rp_ * du[2]) * st;  // (1/mu) d(mu)/dr = +0 (r -+ 5e-12 == r, 1/mu finite)
  du[4] = rtf_inv_y(r, Rd.yr, ok) * ((imu * dmudt - rt_ * du[0]) + ((r * rp_) * du[2]) * ct);
  du[5] = rtf_inv(r * st, ok) * ((0.0 - (rp_ * du[0]) * st) - ((r * rp_) * du[1]) * ct);  // (1/mu) d(mu)/dphi = +0
  du[6] = E.inv_c * (1.0 + rtf_div_y(freq, mu, ymu, ok) * dmudf);
}

// returns true when du holds the exact result
template <bool COOP>
RTF bool rtf_haselgrove_3d(const RtDevEnv &E, const double *u, double freq, double *du) {
  const double r = u[0], th = u[1], rr_ = u[3], rt_ = u[4], rp_ = u[5];
  bool ok = true;
  // 2^16 <= |r| < 2^513 (so that r -+ 5e-12 == r and d(mu)/dr is (mu - mu)/1e-11; trial stages overshoot to
  // negative radii routinely); rho_phi finite, non-zero
  ok = ok && ((unsigned)((__double2hiint(r) & 0x7fffffff) - 0x40f00000) < (0x60000000u - 0x40f00000u));
  ok = ok && rtf_quarter_range(rp_) && (E.density_model == 0) && (E.fd_default != 0);  // other density models / steps: exact code only
  // Two inputs whose result is known without arithmetic — all seven derivatives NaN — and which the garbage trial
  // stages of a step about to be rejected hit often enough to matter (one lane on the exact path stalls its warp
  // for ~10 right-hand sides):
  //  (1) r is NaN: |B| is NaN, so is every Stix parameter and mu, and every derivative has mu as a factor;
  //  (2) the centre density is +-Inf or NaN (negative radius: the plasmasphere's 10^x overflows for L < -11, the
  //      ionosphere's exp for r < -1e9 m): X_e and X_p are +-Inf or NaN, so P is +-Inf or NaN and R, L are each +-Inf or NaN whatever the
  //      gyrofrequency terms are; then either S or D is NaN (Inf - Inf), A = S s^2 + P c^2 or F^2 (through P D) is NaN,
  //      mu = sqrt|(B + F) / 2A| is NaN, and again every derivative has mu as a factor.
  // (NaN payloads never reach a result: the step is rejected on EEst != EEst.)
  const bool nan_in = (unsigned)(__double2hiint(r) & 0x7fffffff) > 0x7ff00000u;  // any quiet NaN
  bool ovf_c = false, ovf_x = false;
  const double two_pi = 2.0 * RT_PI;
  RtRad3D Rd;
  rtf_rad3d(E, r, rr_, rt_, rp_, Rd, ok);
  RtGeo3D gc;
  RtMu c;
  double mu_tl, mu_tr, mu_fl, mu_fr;
  if (COOP) {
    // lane rank 0 / 1 / 2 of the ray's group: theta, theta - 5e-12, theta + 5e-12 (geometry + Stix side by side),
    // then ranks 0 / 1: centre medium at f - 5e-6 / f + 5e-6 (rank 2 repeats f + 5e-6, unused)
    const int k = rt_group_rank();
    const double th_k = (k == 1) ? th - 1.0e-11 / 2.0 : ((k == 2) ? th + 1.0e-11 / 2.0 : th);
    RtGeo3D g;
    rtf_geo3d(E, Rd, r, th_k, rr_, rt_, g, ok, ovf_x);
    ovf_c = __shfl_sync(rt_group_mask(), ovf_x ? 1 : 0, rt_group_base()) != 0;
    RtMu v = rtf_stix<true>(g.p, two_pi * freq, ok);
    gc.Br = rt_group_get(g.Br, 0);
    gc.Bt = rt_group_get(g.Bt, 0);
    gc.Bmag = rt_group_get(g.Bmag, 0);
    gc.p.w_e2 = rt_group_get(g.p.w_e2, 0);
    gc.p.w_p2 = rt_group_get(g.p.w_p2, 0);
    gc.p.W_e = rt_group_get(g.p.W_e, 0);
    gc.p.W_p = rt_group_get(g.p.W_p, 0);
    gc.p.s = rt_group_get(g.p.s, 0);
    gc.p.c = rt_group_get(g.p.c, 0);
    const double f_k = (k == 0) ? freq - 1.0e-5 / 2.0 : freq + 1.0e-5 / 2.0;
    const double m2 = rtf_stix<false>(gc.p, two_pi * f_k, ok).mu;
    c.mu = rt_group_get(v.mu, 0);
    c.dmudpsi = rt_group_get(v.dmudpsi, 0);
    mu_tl = rt_group_get(v.mu, 1);
    mu_tr = rt_group_get(v.mu, 2);
    mu_fl = rt_group_get(m2, 0);
    mu_fr = rt_group_get(m2, 1);
  } else {
    // centre first (with d(mu)/d(psi)); the theta -/+ and f -/+ pairs are loops so that the build can choose
    // between rolled (small register footprint, more resident warps) and unrolled (more chains per warp)
    const double om = two_pi * freq, om2 = om * om, yom2 = rtf_rcp(om2);
    rtf_geo3d(E, Rd, r, th, rr_, rt_, gc, ok, ovf_c);
    c = rtf_stix<true>(gc.p, om, om2, yom2, ok);
    double mt[2], mf[2];
    RT_PAIR_LOOP
    for (int k = 0; k < 2; k++) {
      RtGeo3D g;
      rtf_geo3d(E, Rd, r, k ? th + 1.0e-11 / 2.0 : th - 1.0e-11 / 2.0, rr_, rt_, g, ok, ovf_x);
      const double m = rtf_stix<false>(g.p, om, om2, yom2, ok).mu;
      if (k == 0) mt[0] = m; else mt[1] = m;
    }
    RT_PAIR_LOOP_F
    for (int k = 0; k < 2; k++) {
      const double m = rtf_stix<false>(gc.p, two_pi * (k ? freq + 1.0e-5 / 2.0 : freq - 1.0e-5 / 2.0), ok).mu;
      if (k == 0) mf[0] = m; else mf[1] = m;
    }
    mu_tl = mt[0];
    mu_tr = mt[1];
    mu_fl = mf[0];
    mu_fr = mf[1];
  }
  ok = ok && rtf_quarter_range(gc.Bmag);
  // known NaN result: poison mu; rtf_tail3d is arithmetic and divisions only, every derivative has mu as a factor
  // (its flags fail on the way: the return value below overrides them)
  const bool known_nan = nan_in || ovf_c || Rd.iono_ovf;  // the same on the lanes of a group
  c.mu = known_nan ? rt_nan() : c.mu;
  rtf_tail3d(E, Rd, u, freq, gc, c.mu, c.dmudpsi, mu_tl, mu_tr, mu_fl, mu_fr, du, ok);
  if (COOP) ok = rt_group_all(ok);
  return ok || known_nan;
}

// =====================================================================================================
// 2D, RayTrace_lat.jl preset
// =====================================================================================================
struct RtRad2D {
  double bq;  // B0 (re^3 / r^3)
  double ne_iono;
};

RTF void rtf_rad2d(const RtDevEnv &E, double r, RtRad2D &R, bool &ok) {
  const double r3 = (r * r) * r;
  R.bq = E.B0 * rtf_div(E.re3, r3, ok);
  const double rE = rtf_div_c(r, E.re, E.y_re, ok);
  R.ne_iono = E.iono_n0 * rtf_exp(E.iono_k * (rE - E.iono_h), ok);
}

// rt_geo2d<LAT2D>
RTF void rtf_geo2d_lat(const RtDevEnv &E, const RtRad2D &R, double r, double th, double chi, RtPlasma &p, double &dip,
                       double &psi, bool &ok) {
  double st, ct;
  rtf_sincos(th, &st, &ct, ok);
  // tan(theta): zero at the equator is legitimate (sin = +-0 exactly only for theta = +-0)
  bool okt = true;
  double tn = rtf_div(st, ct, okt);
  const bool zt = (st == 0.0);
  tn = zt ? st : tn;  // cos(0) = 1
  ok = ok && (okt || zt);
  bool oka = true;
  const double tt = 2.0 * tn;
  double at = rtf_atan(tt, oka);
  at = zt ? 0.0 : at;  // rt_atan(+-0) = +0
  ok = ok && (oka || zt);
  dip = at;
  psi = (RT_PI / 2.0 + dip) + chi;
  const double Bmag = R.bq * rtf_sqrt(1.0 + (3.0 * st) * st, ok);
  const double Lshell = rtf_div(r, E.re * (ct * ct), ok);
  bool ovf_unused;
  const double n_e = rtf_density_cm3(E, R.ne_iono, Lshell, ok, ovf_unused) * 1e6;
  rtf_plasma_from(E, n_e, Bmag, p, ok);
  rtf_sincos(psi, &p.s, &p.c, ok);
}

template <bool COOP>
RTF bool rtf_haselgrove_2d_lat(const RtDevEnv &E, const double *u, double *du, RtExtras2D *ex) {
  const double r = u[0], th = u[1], chi = u[2], freq = u[3];
  bool ok = (E.density_model == 0) && (E.fd_default != 0);  // other density models / steps: exact code only
  ok = ok && ((unsigned)((__double2hiint(r) & 0x7fffffff) - 0x40f00000) < (0x60000000u - 0x40f00000u));
  const double two_pi = 2.0 * RT_PI;
  RtRad2D Rd;
  rtf_rad2d(E, r, Rd, ok);
  RtMu c;
  double mu_tl, mu_tr, mu_fl, mu_fr, dip, psi;
  if (COOP) {
    const int k = rt_group_rank();
    const double th_k = (k == 1) ? th - 1.0e-11 / 2.0 : ((k == 2) ? th + 1.0e-11 / 2.0 : th);
    RtPlasma p, pc;
    rtf_geo2d_lat(E, Rd, r, th_k, chi, p, dip, psi, ok);
    RtMu v = rtf_stix<true>(p, two_pi * freq, ok);
    pc.w_e2 = rt_group_get(p.w_e2, 0);
    pc.w_p2 = rt_group_get(p.w_p2, 0);
    pc.W_e = rt_group_get(p.W_e, 0);
    pc.W_p = rt_group_get(p.W_p, 0);
    pc.s = rt_group_get(p.s, 0);
    pc.c = rt_group_get(p.c, 0);
    const double f_k = (k == 0) ? freq - 1.0e-5 / 2.0 : freq + 1.0e-5 / 2.0;
    const double m2 = rtf_stix<false>(pc, two_pi * f_k, ok).mu;
    c.mu = rt_group_get(v.mu, 0);
    c.dmudpsi = rt_group_get(v.dmudpsi, 0);
    mu_tl = rt_group_get(v.mu, 1);
    mu_tr = rt_group_get(v.mu, 2);
    mu_fl = rt_group_get(m2, 0);
    mu_fr = rt_group_get(m2, 1);
    dip = rt_group_get(dip, 0);
    psi = rt_group_get(psi, 0);
  } else {
    RtPlasma pc;
    const double om = two_pi * freq, om2 = om * om, yom2 = rtf_rcp(om2);
    rtf_geo2d_lat(E, Rd, r, th, chi, pc, dip, psi, ok);
    c = rtf_stix<true>(pc, om, om2, yom2, ok);
    double mt[2], mf[2];
    RT_PAIR_LOOP
    for (int k = 0; k < 2; k++) {
      RtPlasma p;
      double d0, d1;
      rtf_geo2d_lat(E, Rd, r, k ? th + 1.0e-11 / 2.0 : th - 1.0e-11 / 2.0, chi, p, d0, d1, ok);
      const double m = rtf_stix<false>(p, om, om2, yom2, ok).mu;
      if (k == 0) mt[0] = m; else mt[1] = m;
    }
    RT_PAIR_LOOP_F
    for (int k = 0; k < 2; k++) {
      const double m = rtf_stix<false>(pc, two_pi * (k ? freq + 1.0e-5 / 2.0 : freq - 1.0e-5 / 2.0), ok).mu;
      if (k == 0) mf[0] = m; else mf[1] = m;
    }
    mu_tl = mt[0];
    mu_tr = mt[1];
    mu_fl = mf[0];
    mu_fr = mf[1];
  }
  const double dmudth = rtf_div_c0(mu_tr - mu_tl, 1.0e-11, E.y_dth, ok);
  const double dmudf = rtf_div_c0(mu_fr - mu_fl, 1.0e-5, E.y_df, ok);
  double sc, cc;
  rtf_sincos(chi, &sc, &cc, ok);
  const double mu = c.mu, dmudpsi = c.dmudpsi;
  const double mm = mu * mu;
  const double rmm = r * mm;
  const double i1 = rtf_inv(mm, ok), i2 = rtf_inv(rmm, ok);
  du[0] = i1 * (mu * cc + dmudpsi * sc);
  du[1] = i2 * (mu * sc - dmudpsi * cc);
  du[2] = i2 * (dmudth * cc - mu * sc);  // r d(mu)/dr + mu = +0 + mu = mu (r > 0, mu > 0)
  du[3] = E.inv_c * (1.0 + rtf_div(freq, mu, ok) * dmudf);
  if (ex) {
    ex->mu = mu;
    ex->dmudpsi = dmudpsi;
    ex->dip = dip;
    ex->psi = psi;
  }
  if (COOP) ok = rt_group_all(ok);
  return ok;
}
