// rt_physics.cuh — device functions for the environment, the phase refractive index and the
// Haselgrove right-hand side.
//
// This is the product path: hand-scheduled for one-thread-per-ray FP64 execution on sm_100a.
// Results are bit-identical to the reference's arithmetic order (no contraction: the file is
// compiled with -fmad=false; fused multiply-adds exist only inside rt_math.h), but the *schedule*
// is not the reference's: per RHS the reference calls its refractive-index function 9 (2D) / 12
// (3D) times; here one RHS costs 3 geometry evaluations (centre, theta-, theta+) and 5 Stix
// evaluations (centre, theta-, theta+, f-, f+), because
//   * d(mu)/dr's two evaluation points coincide with r in FP64 (r -+ 5e-12 == r), so the quotient
//     is (mu_c - mu_c)/dr  — evaluated as written, giving 0.0 or NaN  (RayTrace_lat.jl:198-208);
//   * d(mu)/dchi is computed and discarded by the reference (RayTrace_lat.jl:262);
//   * d(mu)/dphi is (mu_c - mu_c)/dphi since mu does not depend on phi (RayTrace_3D.jl:249-259);
//   * the f- and f+ evaluations share every geometry/plasma quantity with the centre.
#pragma once
#include "../../include/raytrace_b200.h"
#include "rt_math.h"

#define RT_PI 3.141592653589793

// Environment as the kernels see it: rt_env_params plus products of constants that the reference
// re-evaluates on every call (computed on the host with the same IEEE operations).
struct RtDevEnv {
  double c, re, B0, e, me, mp;
  double re3;      // (re*re)*re
  double e2;       // e*e
  double eps_me;   // eps*me
  double eps_mp;   // eps*mp
  double Lppi, Lppo, ne_Lppi, ca_coef, ca_trough;
  double iono_n0, iono_k, iono_h;
  // for the branch-free path (rt_physics_fast.cuh): refined reciprocals of the constant divisors, produced on the
  // device by rt_env_rcp_kernel with the division's own Newton sequence, and 1/c (an IEEE quotient, host-computed)
  double y_re, y_eps_me, y_eps_mp, y_me, y_mp, y_1p5, y_10, y_0p1, y_dth, y_df;
  double inv_c;
  int density_model;  // rt_env_params.density_model
  int fd_default;     // 1: both finite-difference steps are the files' 1.0e-11 (the branch-free path's precondition)
  double h_ang, h_r;  // finite-difference steps of d(mu)/d(theta|lambda), d(mu)/dr (rt_env_params.fd_step_*)
  int deriv_mode, root_mode;  // rt_env_params: [EXT] analytic partials / root selection without abs (rt_physics_ad.cuh)
};

struct RtPlasma {  // local medium at one point (independent of wave frequency)
  double w_e2, w_p2, W_e, W_p;
  double s, c;     // sin(psi), cos(psi)
};

struct RtMu {
  double mu, dmudpsi;
};

// CA92 plasmasphere + exponential ionosphere, cm^-3: RayTrace_lat.jl:69-83, RayTrace_3D.jl:105-119.
// The three CA92 segments go through ONE exp and ONE exp10 call site (arguments selected per segment) to keep
// the instruction footprint small; each lane still evaluates exactly the expression of its own segment, so
// the bits are those of the reference's if/elseif chain.
__device__ __forceinline__ double rt_density_cm3(const RtDevEnv &E, double r, double Lshell) {
  double ne_iono = E.iono_n0 * rt_exp(E.iono_k * ((r / E.re) - E.iono_h));
  const bool seg1 = Lshell <= E.Lppi;
  const bool seg2 = E.Lppi < Lshell && Lshell <= E.Lppo;
  const bool seg3 = E.Lppo < Lshell;
  double ex = 0.0, p10 = 0.0;
  if (seg1 || seg3) ex = rt_exp(seg1 ? (2.0 - Lshell) / 1.5 : (2.0 - Lshell) / 10.0);
  if (seg1 || seg2) p10 = rt_exp10(seg1 ? (-0.3145 * Lshell + 3.9043) + E.ca_coef * ex : (E.Lppi - Lshell) / 0.1);
  double ne_plasma = 0.0;
  if (seg1) ne_plasma = p10;
  if (seg2) ne_plasma = E.ne_Lppi * p10;
  if (seg3) ne_plasma = E.ca_trough * rt_pow_m4p5(Lshell) + (1.0 - ex);
  if (E.density_model == 1) {
    // plasmasphere.jl's composition (:171): diffusive_equilibrium(r, CA92(L)) + ionosphere_eq(r), :96-117
    const double T = 2500.0, rb = 7370000.0;
    const double G = rb * (1.0 - rb / r);                       // :101
    const double q = rb / 7370.0;
    const double S = ((1.506 * T) * (q * q)) * (1.0 / 1.0);     // :102 (i = 1)
    const double ne_de = ne_plasma * RT_SQRT(1.0 * rt_exp(-G / S));  // :105
    return ne_de + 1e5 * rt_exp(-10.0 * ((r / E.re) - 1.0471));  // :115
  }
  return ne_iono + ne_plasma;
}

__device__ __forceinline__ void rt_plasma_from(const RtDevEnv &E, double n_e, double Bmag, RtPlasma &p) {
  p.w_e2 = (n_e * E.e2) / E.eps_me;   // RayTrace_lat.jl:87
  p.w_p2 = (n_e * E.e2) / E.eps_mp;   // :89 (n_p = n_e)
  p.W_e = (E.e * Bmag) / E.me;        // :92
  p.W_p = (E.e * Bmag) / E.mp;        // :94
}

// Stix parameters -> mu = sqrt|mu^2_+| (and dmu/dpsi when WITH_D): RayTrace_lat.jl:96-184
__device__ __forceinline__ RtMu rt_stix(const RtPlasma &p, double omega, bool with_d) {
  double w2 = omega * omega;
  double xe = p.w_e2 / w2, xp = p.w_p2 / w2;
  double R = 1.0 - xe * (omega / (omega - p.W_e)) - xp * (omega / (omega + p.W_p));
  double L = 1.0 - xe * (omega / (omega + p.W_e)) - xp * (omega / (omega - p.W_p));
  double P = 1.0 - xe - xp;
  double D = (R - L) / 2.0;
  double S = (R + L) / 2.0;
  double s = p.s, c = p.c;
  double s2 = s * s, c2 = c * c;
  double A = S * s2 + P * c2;
  double B = (R * L) * s2 + (P * S) * (1.0 + c2);
  double X = R * L - P * S;
  double PD = P * D;
  double PDc = PD * c;
  double F2 = (X * X) * (s2 * s2) + 4.0 * (PDc * PDc);
  double F = RT_SQRT(F2);
  double mu2 = (B + F) / (2.0 * A);
  RtMu out;
  out.mu = RT_SQRT(rt_abs(mu2));
  out.dmudpsi = 0.0;
  if (with_d) {
    double dA = ((2.0 * (S - P)) * s) * c;
    double dB = ((2.0 * X) * s) * c;
    double dF = (1.0 / (2.0 * F)) * ((((X * X) * 4.0) * (s2 * s)) * c - ((8.0 * (PD * PD)) * s) * c);
    out.dmudpsi = (1.0 / (2.0 * out.mu)) * ((dB + dF) / (2.0 * A) - ((2.0 * dA) * (B + F)) / (2.0 * (A * A)));
  }
  return out;
}

// ---- 2D geometry: dip, psi, |B|, density at (r, theta-slot, chi) --------------------------------
// PRESET: RT_PRESET_LAT2D (RayTrace_lat.jl:47-84), RT_PRESET_RAYMAIN (RayMain.jl:128-154),
//         RT_PRESET_SCRATCH (scratch.jl:66-112)
template <int PRESET>
__device__ __forceinline__ void rt_geo2d(const RtDevEnv &E, double r, double th, double chi, RtPlasma &p,
                                         double &dip, double &psi) {
  double st, ct;
  rt_sincos(th, &st, &ct);
  double tn = st / ct;  // == rt_tan(th)
  double r3 = (r * r) * r;
  double Bmag, n_e;
  if (PRESET == RT_PRESET_LAT2D) {
    dip = rt_atan(2.0 * tn);
    psi = (RT_PI / 2.0 + dip) + chi;
    Bmag = (E.B0 * (E.re3 / r3)) * RT_SQRT(1.0 + (3.0 * st) * st);
    double Lshell = r / (E.re * (ct * ct));
    n_e = rt_density_cm3(E, r, Lshell) * 1e6;
  } else {
    dip = rt_atan(2.0 * (1.0 / tn));
    if (PRESET == RT_PRESET_RAYMAIN)
      psi = (RT_PI / 2.0 + dip) + chi;
    else
      psi = ((3.0 * RT_PI) / 2.0 - dip) - chi;
    Bmag = (E.B0 * (E.re3 / r3)) * RT_SQRT(1.0 + (3.0 * ct) * ct);
    if (PRESET == RT_PRESET_RAYMAIN) {
      double rE = r / E.re;
      n_e = 1.e6 * (E.iono_n0 * rt_exp(E.iono_k * (rE - E.iono_h)));
    } else {
      double cl = rt_cos(RT_PI / 2.0 - th);
      double Lshell = r / (E.re * (cl * cl));
      n_e = rt_density_cm3(E, r, Lshell) * 1e6;
    }
  }
  rt_plasma_from(E, n_e, Bmag, p);
  rt_sincos(psi, &p.s, &p.c);
}

// phase_refractive_index(r, theta, chi, freq) -> [mu, dmudpsi]
template <int PRESET>
__device__ __forceinline__ RtMu rt_phase_refractive_index_2d(const RtDevEnv &E, double r, double th, double chi,
                                                             double freq, double *dip_out, double *psi_out) {
  RtPlasma p;
  double dip, psi;
  rt_geo2d<PRESET>(E, r, th, chi, p, dip, psi);
  if (dip_out) *dip_out = dip;
  if (psi_out) *psi_out = psi;
  return rt_stix(p, (2.0 * RT_PI) * freq, true);
}

// mu only, out of line: the rarely-taken slow path of d(mu)/dr (|r| so small that r -+ 5e-12 != r)
template <int PRESET>
__device__ __noinline__ double rt_mu2d_slow(const RtDevEnv &E, double r, double th, double chi, double freq) {
  RtPlasma p;
  double dip, psi;
  rt_geo2d<PRESET>(E, r, th, chi, p, dip, psi);
  return rt_stix(p, (2.0 * RT_PI) * freq, false).mu;
}

struct RtExtras2D {  // save_func by-products, RayTrace_lat.jl:318-324
  double mu, dmudpsi, dip, psi;
};

// The five refractive-index evaluations one RHS needs, as a ROLLED loop so that the geometry and Stix code
// is instantiated once (the kernel has to fit the instruction cache; with the evaluations inlined side by side
// the first profile showed "no instruction" as the top stall).  point 0 = centre (with dmu/dpsi), 1/2 =
// theta -/+ 5e-12, 3/4 = centre medium at f -/+ 5e-6.
// Cooperative mode (rt_physics_fast.cuh, COOP = true): a ray is owned by a group of 3 adjacent lanes that all hold
// the same state (10 rays per warp, lanes 30 and 31 idle).  The three theta points of the loop above are evaluated
// by the three lanes IN PARALLEL, then lanes 0 and 1 take the two frequency points, and the results are exchanged
// with shuffles.  Every evaluation executes the same instructions on the same inputs as in the one-thread-per-ray
// code, so the results are bit-identical; what changes is the latency of one right-hand side (geometry + 2 Stix
// evaluations instead of 3 + 5) at ~1.4x the issue cost, which is what the long dependent chains of the
// magnetospherically reflected rays need.
#define RT_COOP_LANES 3
#define RT_COOP_RAYS_PER_WARP 10
__device__ __forceinline__ int rt_group_rank() { const int l = threadIdx.x & 31; return l - 3 * ((l * 11) >> 5); }
__device__ __forceinline__ int rt_group_base() { const int l = threadIdx.x & 31; return 3 * ((l * 11) >> 5); }
__device__ __forceinline__ unsigned rt_group_mask() { return 7u << rt_group_base(); }
__device__ __forceinline__ double rt_group_get(double x, int src) {
  return __shfl_sync(rt_group_mask(), x, rt_group_base() + src);
}
// true when `pred` holds on all three lanes of the group.  Three shuffles: a __ballot_sync with a per-group mask
// compiles to a loop over the warp's distinct masks (10 VOTE iterations per call in the first cooperative profile).
__device__ __forceinline__ bool rt_group_all(bool pred) {
  const unsigned m = rt_group_mask();
  const int b = rt_group_base(), v = pred ? 1 : 0;
  return (__shfl_sync(m, v, b) & __shfl_sync(m, v, b + 1) & __shfl_sync(m, v, b + 2)) != 0;
}

struct RtFivePoint {
  double mu_c, dmudpsi, mu_tl, mu_tr, mu_fl, mu_fr;
};

// haselgrove!(du,u,p,t) 2D: RayTrace_lat.jl:251-273 (signs: RayMain.jl:341-342)
template <int PRESET>
__device__ __forceinline__ void rt_haselgrove_2d(const RtDevEnv &E, const double *u, double *du, RtExtras2D *ex) {
  const double r = u[0], th = u[1], chi = u[2], freq = u[3];
  const double two_pi = 2.0 * RT_PI;
  RtPlasma pc, p;
  double dip = 0.0, psi = 0.0;
  RtFivePoint m = {0, 0, 0, 0, 0, 0};
#pragma unroll 1
  for (int k = 0; k < 5; k++) {
    double f_k = freq;
    if (k < 3) {
      double th_k = (k == 0) ? th : ((k == 1) ? th - E.h_ang / 2.0 : th + E.h_ang / 2.0);
      double d0, d1;
      rt_geo2d<PRESET>(E, r, th_k, chi, p, d0, d1);
      if (k == 0) {
        pc = p;
        dip = d0;
        psi = d1;
      }
    } else {
      p = pc;
      f_k = (k == 3) ? freq - 1.0e-5 / 2.0 : freq + 1.0e-5 / 2.0;
    }
    RtMu v = rt_stix(p, two_pi * f_k, k == 0);
    if (k == 0) { m.mu_c = v.mu; m.dmudpsi = v.dmudpsi; }
    if (k == 1) m.mu_tl = v.mu;
    if (k == 2) m.mu_tr = v.mu;
    if (k == 3) m.mu_fl = v.mu;
    if (k == 4) m.mu_fr = v.mu;
  }
  double dmudth = (m.mu_tr - m.mu_tl) / E.h_ang;
  double dmudf = (m.mu_fr - m.mu_fl) / 1.0e-5;
  double rl = r - E.h_r / 2.0, rr = r + E.h_r / 2.0;
  double dmudr;
  if (rl == r && rr == r) {
    dmudr = (m.mu_c - m.mu_c) / E.h_r;  // 0.0, or NaN when mu is not finite
  } else {
    dmudr = (rt_mu2d_slow<PRESET>(E, rr, th, chi, freq) - rt_mu2d_slow<PRESET>(E, rl, th, chi, freq)) / E.h_r;
  }
  double sc, cc;
  rt_sincos(chi, &sc, &cc);
  const double mu = m.mu_c, dmudpsi = m.dmudpsi;
  if (PRESET != RT_PRESET_RAYMAIN) {
    du[0] = (1.0 / (mu * mu)) * (mu * cc + dmudpsi * sc);
    du[1] = (1.0 / (r * (mu * mu))) * (mu * sc - dmudpsi * cc);
  } else {
    du[0] = (1.0 / (mu * mu)) * (mu * cc - dmudpsi * sc);
    du[1] = (1.0 / (r * (mu * mu))) * (mu * sc + dmudpsi * cc);
  }
  du[2] = (1.0 / (r * (mu * mu))) * (dmudth * cc - (r * dmudr + mu) * sc);
  du[3] = (1.0 / E.c) * (1.0 + (freq / mu) * dmudf);
  if (ex) {
    ex->mu = mu;
    ex->dmudpsi = dmudpsi;
    ex->dip = dip;
    ex->psi = psi;
  }
}

// ---- 3D: RayTrace_3D.jl -----------------------------------------------------------------------
struct RtGeo3D {
  RtPlasma p;
  double Br, Bt, Bp, Bmag, psi;
};

// magnetic_field (RayTrace_3D.jl:54-66) + medium (:97-141) at (r, theta) for refractive-index vector rho
__device__ __forceinline__ void rt_geo3d(const RtDevEnv &E, double r, double th, double rho_r, double rho_t,
                                         double rho_p, RtGeo3D &g) {
  double lam = RT_PI / 2.0 - th;
  double sl, cl;
  rt_sincos(lam, &sl, &cl);
  double q = E.re / r;
  double q3 = (q * q) * q;
  g.Br = ((-2.0 * E.B0) * q3) * sl;
  double Bl = (E.B0 * q3) * cl;
  g.Bt = -1.0 * Bl;
  g.Bp = 0.0;
  g.Bmag = RT_SQRT((g.Br * g.Br + g.Bt * g.Bt) + g.Bp * g.Bp);
  double L = r / (E.re * (cl * cl));  // cos(pi/2 - theta)^2, :105
  double n_e = rt_density_cm3(E, r, L) * 1e6;
  double mumag = RT_SQRT((rho_r * rho_r + rho_t * rho_t) + rho_p * rho_p);
  double dot = (g.Br * rho_r + g.Bt * rho_t) + g.Bp * rho_p;
  double cos_psi = dot / (g.Bmag * mumag);
  g.psi = rt_acos(cos_psi);
  rt_plasma_from(E, n_e, g.Bmag, g.p);
  rt_sincos(g.psi, &g.p.s, &g.p.c);
}

__device__ __noinline__ double rt_mu3d_slow(const RtDevEnv &E, double r, double th, double rho_r, double rho_t,
                                            double rho_p, double freq) {
  RtGeo3D g;
  rt_geo3d(E, r, th, rho_r, rho_t, rho_p, g);
  return rt_stix(g.p, (2.0 * RT_PI) * freq, false).mu;
}

// Kimura-form d(mu)/d(rho_k), RayTrace_3D.jl:261-311; k selects the component
__device__ __forceinline__ double rt_ddrho(const RtGeo3D &g, const RtMu &c, double rho_k, int k) {
  double v0 = (k == 0) ? rho_k : 0.0, v1 = (k == 1) ? rho_k : 0.0, v2 = (k == 2) ? rho_k : 0.0;
  double Bmag = RT_SQRT((g.Br * g.Br + g.Bt * g.Bt) + g.Bp * g.Bp);
  double cos_a = ((g.Br * v0 + g.Bt * v1) + g.Bp * v2) / (Bmag * rt_abs(rho_k));
  return c.dmudpsi * ((rho_k * g.p.c - c.mu * cos_a) / ((c.mu * c.mu) * g.p.s));
}

// haselgrove!(du,u,p,t) 3D: RayTrace_3D.jl:327-359; u = (r,theta,phi,rho_r,rho_theta,rho_phi,T), p = [freq]
__device__ __forceinline__ void rt_haselgrove_3d(const RtDevEnv &E, const double *u, double freq, double *du) {
  const double r = u[0], th = u[1], rr_ = u[3], rt_ = u[4], rp_ = u[5];
  const double two_pi = 2.0 * RT_PI;
  RtGeo3D gc, g;
  RtFivePoint m = {0, 0, 0, 0, 0, 0};
#pragma unroll 1
  for (int k = 0; k < 5; k++) {
    double f_k = freq;
    if (k < 3) {
      double th_k = (k == 0) ? th : ((k == 1) ? th - E.h_ang / 2.0 : th + E.h_ang / 2.0);
      rt_geo3d(E, r, th_k, rr_, rt_, rp_, g);
      if (k == 0) gc = g;
    } else {
      g.p = gc.p;
      f_k = (k == 3) ? freq - 1.0e-5 / 2.0 : freq + 1.0e-5 / 2.0;
    }
    RtMu v = rt_stix(g.p, two_pi * f_k, k == 0);
    if (k == 0) { m.mu_c = v.mu; m.dmudpsi = v.dmudpsi; }
    if (k == 1) m.mu_tl = v.mu;
    if (k == 2) m.mu_tr = v.mu;
    if (k == 3) m.mu_fl = v.mu;
    if (k == 4) m.mu_fr = v.mu;
  }
  RtMu c;
  c.mu = m.mu_c;
  c.dmudpsi = m.dmudpsi;
  double dmudt = (m.mu_tr - m.mu_tl) / E.h_ang;
  double dmudf = (m.mu_fr - m.mu_fl) / 1.0e-5;
  double dmudp = (c.mu - c.mu) / 1.0e-11;  // mu independent of phi
  double rl = r - E.h_r / 2.0, rh = r + E.h_r / 2.0;
  double dmudr;
  if (rl == r && rh == r) {
    dmudr = (c.mu - c.mu) / E.h_r;
  } else {
    dmudr = (rt_mu3d_slow(E, rh, th, rr_, rt_, rp_, freq) - rt_mu3d_slow(E, rl, th, rr_, rt_, rp_, freq)) / E.h_r;
  }
  double dmr = rt_ddrho(gc, c, rr_, 0);
  double dmt = rt_ddrho(gc, c, rt_, 1);
  double dmp = rt_ddrho(gc, c, rp_, 2);
  double st, ct;
  rt_sincos(th, &st, &ct);
  const double mu = c.mu;
  du[0] = (1.0 / (mu * mu)) * (rr_ - mu * dmr);
  du[1] = (1.0 / (r * (mu * mu))) * (rt_ - mu * dmt);
  du[2] = (1.0 / ((r * (mu * mu)) * st)) * (rp_ - mu * dmp);
  du[3] = ((1.0 / mu) * dmudr + rt_ * du[1]) + (rp_ * du[2]) * st;
  du[4] = (1.0 / r) * (((1.0 / mu) * dmudt - rt_ * du[0]) + ((r * rp_) * du[2]) * ct);
  du[5] = (1.0 / (r * st)) * (((1.0 / mu) * dmudp - (rp_ * du[0]) * st) - ((r * rp_) * du[1]) * ct);
  du[6] = (1.0 / E.c) * (1.0 + (freq / mu) * dmudf);
}
