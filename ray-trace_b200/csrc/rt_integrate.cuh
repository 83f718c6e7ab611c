// rt_integrate.cuh — the fused hot kernel: embedded adaptive Tsit5 + PI step control + Earth-hit event
// + trajectory sampling around the Haselgrove RHS, one thread per ray.
//
// Replaces `solve(ODEProblem(haselgrove!,u0,tspan,p), ...; callback=cb)` at RayTrace_lat.jl:337-338 and
// RayTrace_3D.jl:394-395.  The integrator semantics (OrdinaryDiffEq's Tsit5, PI controller, initial step,
// ContinuousCallback, SavingCallback — third-party, absent from the reference tree) are restated in
// DESIGN.md §3; the CPU oracle implements the same specification independently.
//
// Execution model
//   * structure-of-arrays ray state in HBM; a launch processes the rays listed in idx_in[0..nactive);
//   * each thread advances its ray by at most `attempts` step attempts (a "slice"), then writes the
//     state back; rays still running are appended to idx_out by a warp-aggregated atomic, so the next
//     launch runs on full warps again (periodic stream compaction);
//   * the seven stage derivatives live in shared memory as k[stage][component][thread] (conflict-free,
//     one 8-byte word per lane), which lets the stage loop stay rolled: the RHS is instantiated once;
//   * warps leave the attempt loop as soon as no lane is running (__any_sync vote);
//   * the work-list length is read from device memory, so slices are enqueued back to back without a host
//     round trip; the host reads the counter once per batch of slices to shrink the grid and detect the end.
#pragma once
#include "rt_physics_fast.cuh"
#include "rt_physics_ad.cuh"

#ifndef RT_BLOCK
#define RT_BLOCK 128
#endif
#ifndef RT_COOP_MINBLOCKS
#define RT_COOP_MINBLOCKS 1
#endif
#ifndef RT_MINBLOCKS
#define RT_MINBLOCKS 3  // __launch_bounds__ second argument for the integrate kernels (register budget)
#endif

struct RtDevOpts {
  int adaptive, save_mode, max_samples, save_stride;
  int maxiters, nan_policy;
  double t1, dt, reltol, abstol, dtmax, dtmin, r_esc;
};

struct RtState {
  long long n;
  double *u;        // [DIM][n]  (also the caller's u_end)
  double *k1;       // [DIM][n]  FSAL derivative
  double *t;        // [n]       (also t_end)
  double *h;        // [n]
  double *qold;     // [n]
  double *par;      // [n] 2D: launch frequency f0; 3D: wave frequency p[1]
  int *status, *naccept, *nreject, *nfe, *nreflect;  // [n] (caller's result arrays)
  int *sgn;         // [n] sign of d(latitude)/dt at the last accepted step
  int *nsamp;       // [n] trajectory samples written
  int *trunc;       // [n] 1 when a sample was dropped because max_samples was reached
  double *traj;     // [n][max_samples][NCOL]
};

template <int MODE>
struct RtTraits {
  static constexpr int DIM = (MODE == RT_PRESET_RAY3D) ? 7 : 4;
  static constexpr int NCOL = (MODE == RT_PRESET_RAY3D) ? 8 : 9;
};

// Tsit5 tableau (Tsitouras 2011); row s holds a_{s+1,1..s}; row 5 = b (FSAL); BT = b - b_hat
__constant__ double RT_A[6][6] = {
    {0.161, 0, 0, 0, 0, 0},
    {-0.008480655492356989, 0.335480655492357, 0, 0, 0, 0},
    {2.8971530571054935, -6.359448489975075, 4.3622954328695815, 0, 0, 0},
    {5.325864828439257, -11.748883564062828, 7.4955393428898365, -0.09249506636175525, 0, 0},
    {5.86145544294642, -12.92096931784711, 8.159367898576159, -0.071584973281401, -0.028269050394068383, 0},
    {0.09646076681806523, 0.01, 0.4798896504144996, 1.379008574103742, -3.290069515436081, 2.324710524099774}};
__constant__ double RT_BT[7] = {-0.00178001105222577714, -0.0008164344596567469, 0.007880878010261995,
                                -0.1447110071732629,     0.5823571654525552,     -0.45808210592918697,
                                0.015151515151515152};

// dense-output weights b_i(theta) of the free 4th-order interpolant
__host__ __device__ __forceinline__ void rt_dense_weights(double th, double *b) {
  double th2 = th * th;
  b[0] = ((-1.0530884977290216 * th) * (th - 1.3299890189751412)) *
         ((th2 - 1.4364028541716351 * th) + 0.7139816917074209);
  b[1] = (0.1017 * th2) * ((th2 - 2.1966568338249754 * th) + 1.2949852507374631);
  b[2] = (2.490627285651252793 * th2) * ((th2 - 2.38535645472061657 * th) + 1.57803468208092486);
  b[3] = ((-16.54810288924490272 * (th - 1.21712927295533244)) * (th - 0.61620406037800089)) * th2;
  b[4] = ((47.37952196281928122 * (th - 1.203071208372362603)) * (th - 0.658047292653547382)) * th2;
  b[5] = ((-34.87065786149660974 * (th - 1.2)) * (th - 0.666666666666666667)) * th2;
  b[6] = ((2.5 * (th - 1.0)) * (th - 0.6)) * th2;
}

// ContinuousCallback's interior safety check (default interp_points = 10: Theta = i/9, i = 1..8, evaluated when the end
// point shows no sign change; DiffEqBase determine_event_occurance).  The weights b_q(i/9) are evaluated once on the
// host with rt_dense_weights (same operations as the oracle's) and uploaded by rt_create.
__constant__ double RT_BI[8][7];
struct RtScanIn { double k[7], u0, h, re, g0; };
// first i in 1..8 with a strict sign change of r(Theta_i) - re against g0, 0 if none (out of line: rare)
__device__ __noinline__ int rt_event_scan(RtScanIn in) {
  for (int ip = 0; ip < 8; ip++) {
    double s = RT_BI[ip][0] * in.k[0];
#pragma unroll
    for (int q = 1; q < 7; q++) s = s + RT_BI[ip][q] * in.k[q];
    const double gi = (in.u0 + in.h * s) - in.re;
    if ((in.g0 > 0.0 && gi < 0.0) || (in.g0 < 0.0 && gi > 0.0)) return ip + 1;
  }
  return 0;
}

// ---- the right-hand side as the kernels call it -----------------------------------------------------
// Fast path: the branch-free stream of rt_physics_fast.cuh.  When its flag comes back false (non-finite trial
// stage, pole, zero numerator, argument outside a fast domain) the lane re-evaluates with the exact, branching
// code of rt_physics.cuh, out of line.  Both produce the same bits wherever the fast path vouches for itself
// (tests/test_gpu_parity.py compares against the CPU oracle, which only has the exact code).
struct RtVec7 { double v[7]; };
struct RtVec4x { double v[4]; RtExtras2D ex; };

// Exact code, out of line.  The -DRT_DEV build also keeps a census of these calls (cold path only): [0] calls, [1] r < 2^16
// or negative, [2] non-finite input, [3] non-finite result, [4] finite in and out (a guard was merely conservative),
// [5] 2D calls, [6] r is NaN, [7] |r| < 2^16; and the first RT_DBG_SAMPLES inputs (3D: u[0..6], freq) since the last reset.
#ifdef RT_DEV
__device__ unsigned long long rt_dbg_fallback[8];
#define RT_DBG_SAMPLES 8192
__device__ double rt_dbg_sample[RT_DBG_SAMPLES][8];
#endif

__device__ __noinline__ RtVec7 rt_haselgrove_3d_exact(const RtDevEnv *E, RtVec7 y, double par) {
  RtVec7 d;
  if (E->deriv_mode | E->root_mode) rt_haselgrove_3d_ext(*E, y.v, par, d.v);  // [EXT] options, rt_physics_ad.cuh
  else rt_haselgrove_3d(*E, y.v, par, d.v);
#ifdef RT_DEV
  bool fin_in = true, fin_out = true;
#pragma unroll
  for (int c = 0; c < 7; c++) {
    fin_in = fin_in && rt_isfinite(y.v[c]);
    fin_out = fin_out && rt_isfinite(d.v[c]);
  }
  const unsigned long long slot = atomicAdd(&rt_dbg_fallback[0], 1ULL);
  if (slot < RT_DBG_SAMPLES) {
#pragma unroll
    for (int c = 0; c < 7; c++) rt_dbg_sample[slot][c] = y.v[c];
    rt_dbg_sample[slot][7] = par;
  }
  if (y.v[0] != y.v[0]) atomicAdd(&rt_dbg_fallback[6], 1ULL);
  else if (rt_abs(y.v[0]) < 65536.0) atomicAdd(&rt_dbg_fallback[7], 1ULL);
  if (!(y.v[0] >= 65536.0)) atomicAdd(&rt_dbg_fallback[1], 1ULL);
  else if (!fin_in) atomicAdd(&rt_dbg_fallback[2], 1ULL);
  else if (!fin_out) atomicAdd(&rt_dbg_fallback[3], 1ULL);
  else atomicAdd(&rt_dbg_fallback[4], 1ULL);
#endif
  return d;
}
template <int PRESET>
__device__ __noinline__ RtVec4x rt_haselgrove_2d_exact(const RtDevEnv *E, RtVec4x y) {
  RtVec4x d;
  if (PRESET == RT_PRESET_LAT2D && (E->deriv_mode | E->root_mode)) rt_haselgrove_2d_lat_ext(*E, y.v, d.v, &d.ex);
  else rt_haselgrove_2d<PRESET>(*E, y.v, d.v, &d.ex);
#ifdef RT_DEV
  atomicAdd(&rt_dbg_fallback[5], 1ULL);
#endif
  return d;
}

#ifndef RT_FORCE_EXACT
#define RT_FORCE_EXACT 0
#endif

template <int MODE, bool COOP = false>
__device__ __forceinline__ void rt_rhs(const RtDevEnv &E, const double *y, double par, double *dy, RtExtras2D *ex) {
  if (MODE == RT_PRESET_RAY3D) {
    if (RT_FORCE_EXACT || !rtf_haselgrove_3d<COOP>(E, y, par, dy)) {
      RtVec7 in;
#pragma unroll
      for (int c = 0; c < 7; c++) in.v[c] = y[c];
      RtVec7 out = rt_haselgrove_3d_exact(&E, in, par);  // (cooperative mode: every lane of the group, redundantly)
#pragma unroll
      for (int c = 0; c < 7; c++) dy[c] = out.v[c];
    }
  } else if (MODE == RT_PRESET_LAT2D) {
    if (RT_FORCE_EXACT || !rtf_haselgrove_2d_lat<COOP>(E, y, dy, ex)) {
      RtVec4x in;
#pragma unroll
      for (int c = 0; c < 4; c++) in.v[c] = y[c];
      RtVec4x out = rt_haselgrove_2d_exact<RT_PRESET_LAT2D>(&E, in);
#pragma unroll
      for (int c = 0; c < 4; c++) dy[c] = out.v[c];
      if (ex) *ex = out.ex;
    }
  } else {
    rt_haselgrove_2d<MODE>(E, y, dy, ex);  // RayMain / scratch variants: exact code only (replicated when cooperative)
  }
}

// ---- trajectory output (sol.t / sol[i,:] / saved_mu: RayTrace_lat.jl:318-329,345-372) --------------------------------
// A ray's samples are rows (t, u[0..DIM)[, mu, dmudpsi, dip, psi]) of its own region traj[ray][max_samples][NCOL], so
// consecutive samples of one ray are consecutive in memory.  The kernel does not store a row when it is produced (32
// lanes x NCOL scattered 8-byte stores, one sector each): every thread stages up to RT_STAGE rows in shared memory and
// the WARP writes a full stage out together, one lane's rows at a time: RT_STAGE * NCOL consecutive doubles per store
// instruction (144 B in 2D, 128 B in 3D).  Capacity: rows 0 .. max_samples-2 take the regular samples (t0 and every
// save_stride-th accepted step); the last row is reserved for the final sample (the event point, the step that
// reaches t1 or escapes), so a trajectory that does not fit is cut in the middle, never at its end, and is flagged
// in trunc[] (rt_traj_truncated).
#ifndef RT_STAGE
#define RT_STAGE 2
#endif

// one row, straight to global memory (rt_init_kernel's sample 0)
template <int MODE>
__device__ __forceinline__ void rt_traj_row(double *row, double t, const double *u, const RtExtras2D &ex) {
  constexpr int DIM = RtTraits<MODE>::DIM;
  row[0] = t;
#pragma unroll
  for (int c = 0; c < DIM; c++) row[1 + c] = u[c];
  if (MODE != RT_PRESET_RAY3D) {
    row[5] = ex.mu;
    row[6] = ex.dmudpsi;
    row[7] = ex.dip;
    row[8] = ex.psi;
  }
}

// stage one sample in this thread's shared-memory slot (cnt rows are in global memory, staged in the slot)
template <int MODE>
__device__ __forceinline__ void rt_traj_push(const RtDevOpts &O, double *slot, int cnt, int &staged, bool &truncated,
                                             bool final_sample, double t, const double *u, const RtExtras2D &ex,
                                             bool writer) {
  constexpr int NCOL = RtTraits<MODE>::NCOL;
  if (!final_sample && cnt + staged >= O.max_samples - 1) {
    truncated = true;
    return;
  }
  if (writer) rt_traj_row<MODE>(slot + staged * NCOL, t, u, ex);
  staged++;
}

// warp-cooperative flush: every lane with `want` set gets its staged rows written by the whole warp as one contiguous
// burst.  Must be reached by all 32 lanes.  `stage_warp` = the shared-memory slots of this warp's 32 threads.
template <int MODE>
__device__ __forceinline__ void rt_traj_flush(const RtState &S, const RtDevOpts &O, const double *stage_warp, int i,
                                              int &cnt, int &staged, bool want, bool writer) {
  constexpr int NCOL = RtTraits<MODE>::NCOL;
  const int lane = threadIdx.x & 31;
  const bool mine = want && staged > 0;
  unsigned m = __ballot_sync(0xffffffffu, mine && writer);
  __syncwarp();  // the owners' staging stores are visible to the warp
  while (m) {
    const int L = __ffs(m) - 1;
    m &= m - 1;
    const int iL = __shfl_sync(0xffffffffu, i, L);
    const int cL = __shfl_sync(0xffffffffu, cnt, L);
    const int nL = __shfl_sync(0xffffffffu, staged, L) * NCOL;
    const double *src = stage_warp + L * (RT_STAGE * NCOL);
    double *dst = S.traj + ((size_t)iL * (size_t)O.max_samples + (size_t)cL) * NCOL;
    for (int e = lane; e < nL; e += 32) dst[e] = src[e];
  }
  __syncwarp();  // the slots may be overwritten from here on
  if (mine) {
    cnt += staged;
    staged = 0;
  }
}

// x^p through exp(p log x), out of line: three call sites per attempt would otherwise inline exp+log each
__device__ __noinline__ double rt_pow_dev(double x, double p) { return rt_pow_pos(x, p); }

// save_func at an event point (the only sample that is not a stage point), out of line
template <int PRESET>
__device__ __noinline__ void rt_extras_at(const RtDevEnv &E, const double *u, RtExtras2D *ex) {
  if (PRESET == RT_PRESET_LAT2D && E.root_mode) {
    RtMuAD o;
    rd_mu2d_lat(E, u[0], u[1], u[2], u[3], o);
    *ex = RtExtras2D{o.mu.v, o.dmudpsi, o.dip, o.psi};
    return;
  }
  RtMu m = rt_phase_refractive_index_2d<PRESET>(E, u[0], u[1], u[2], u[3], &ex->dip, &ex->psi);
  ex->mu = m.mu;
  ex->dmudpsi = m.dmudpsi;
}

__device__ __forceinline__ int rt_sign(double x) { return (x > 0.0) - (x < 0.0); }

template <int DIM>
__device__ __forceinline__ double rt_rms(const double *v, const double *sk) {
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < DIM; c++) {
    double q = v[c] / sk[c];
    s = s + q * q;
  }
  return RT_SQRT(s / (double)DIM);
}

// EEst = sqrt(mean((e_i / sk_i)^2)) with IEEE division and sqrt (cold path of the kernel's error norm)
template <int DIM>
struct RtErrIn {
  double e[DIM], sk[DIM];
};
template <int DIM>
__device__ __noinline__ double rt_eest_exact(RtErrIn<DIM> in) {
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < DIM; c++) {
    double q = in.e[c] / in.sk[c];
    s = s + q * q;
  }
  return RT_SQRT(s / (double)DIM);
}

// Append ray i to the next launch's work list (warp-aggregated atomic).
__device__ __forceinline__ void rt_enqueue(bool pred, int i, int *idx_out, int *count_out) {
  unsigned m = __ballot_sync(0xffffffffu, pred);
  if (m == 0) return;
  int lane = threadIdx.x & 31;
  int leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(count_out, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (pred) idx_out[base + __popc(m & ((1u << lane) - 1u))] = i;
}

// ---- init: ODEProblem(haselgrove!, u0, tspan, p) + first RHS + automatic initial step -----------
// launch columns (device pointers): 2D  L[0..4] = r, theta, chi, f, t0
//                                   3D  L[0..8] = r, theta, phi, rho_r, rho_theta, rho_phi, T0, f, t0
struct RtLaunch {
  const double *col[9];
};

template <int MODE>
// Rays are appended to one of two work lists.  List 1 ("priority") takes the rays that start upward
// (dr/dt > 0 at launch): those are the ones that can travel to the conjugate hemisphere and need 10-40x the
// steps of a ray that goes straight down, so the host starts them first, on their own stream, and the long
// dependent chains overlap the bulk of the sweep instead of trailing it.  Scheduling only: results do not
// depend on which list a ray is in.
__global__ void __launch_bounds__(RT_BLOCK) rt_init_kernel(const __grid_constant__ RtDevEnv E, RtDevOpts O, RtState S, RtLaunch L,
                                                           int *idx_out, int *count_out, int *idx_prio,
                                                           int *count_prio) {
  constexpr int DIM = RtTraits<MODE>::DIM;
  long long i = (long long)blockIdx.x * RT_BLOCK + threadIdx.x;
  bool valid = i < S.n;
  bool running = false, prio = false;
  if (valid) {
    double u[DIM], k1[DIM], y[DIM], f1[DIM], sk[DIM];
#pragma unroll
    for (int c = 0; c < DIM; c++) u[c] = L.col[c][i];
    double par = (MODE == RT_PRESET_RAY3D) ? L.col[7][i] : u[3];
    double t0 = L.col[(MODE == RT_PRESET_RAY3D) ? 8 : 4][i];
    RtExtras2D ex = {0, 0, 0, 0}, ex1;
    int nfe = 0, status = RT_RUNNING;
    const bool auto_dt = O.adaptive && !(O.dt > 0.0);
    double h = O.dt, d1 = 0.0, dt0 = 0.0;
    // pass 0: k1 = f(u0); pass 1 (automatic initial step only): f(u0 + dt0 k1).  Rolled: one RHS instance.
#pragma unroll 1
    for (int pass = 0; pass < (auto_dt ? 2 : 1); pass++) {
      if (pass == 0) {
#pragma unroll
        for (int c = 0; c < DIM; c++) y[c] = u[c];
      } else {
#pragma unroll
        for (int c = 0; c < DIM; c++) sk[c] = O.abstol + rt_abs(u[c]) * O.reltol;
        double d0 = rt_rms<DIM>(u, sk);
        d1 = rt_rms<DIM>(k1, sk);
        dt0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * (d0 / d1);
        if (dt0 > O.dtmax) dt0 = O.dtmax;
        if (dt0 > O.t1 - t0) dt0 = O.t1 - t0;
#pragma unroll
        for (int c = 0; c < DIM; c++) y[c] = u[c] + dt0 * k1[c];
      }
      rt_rhs<MODE>(E, y, par, f1, &ex1);
      nfe++;
      if (pass == 0) {
        ex = ex1;
#pragma unroll
        for (int c = 0; c < DIM; c++) k1[c] = f1[c];
      }
    }
    if (auto_dt) {
#pragma unroll
      for (int c = 0; c < DIM; c++) f1[c] = f1[c] - k1[c];
      double d2 = rt_rms<DIM>(f1, sk) / dt0;
      double md = (d1 > d2) ? d1 : d2;
      double dt1;
      if (md <= 1e-15) {
        double a = dt0 * 1e-3;
        dt1 = (1e-6 > a) ? 1e-6 : a;
      } else {
        dt1 = rt_exp10(-(2.0 + rt_log10(md)) / 5.0);
      }
      double c100 = 100.0 * dt0;
      h = (c100 < dt1) ? c100 : dt1;
      if (h > O.dtmax) h = O.dtmax;
      if (!rt_isfinite(h) || d1 != d1 || d2 != d2) {
        status = RT_DT_NAN;
        h = rt_nan();
      }
    }
#pragma unroll
    for (int c = 0; c < DIM; c++) {
      S.u[(size_t)c * S.n + i] = u[c];
      S.k1[(size_t)c * S.n + i] = k1[c];
    }
    S.t[i] = t0;
    S.h[i] = h;
    S.qold[i] = 1e-4;
    S.par[i] = par;
    S.status[i] = status;
    S.naccept[i] = 0;
    S.nreject[i] = 0;
    S.nfe[i] = nfe;
    S.nreflect[i] = 0;
    S.sgn[i] = rt_sign(k1[1]);
    int cnt = 0;
    if (O.save_mode) {
      const bool room = O.max_samples >= 2;  // row 0 is a regular sample; the last row is reserved for the final one
      if (room) rt_traj_row<MODE>(S.traj + (size_t)i * (size_t)O.max_samples * RtTraits<MODE>::NCOL, t0, u, ex);
      cnt = room ? 1 : 0;
      S.trunc[i] = room ? 0 : 1;
    }
    S.nsamp[i] = cnt;
    running = (status == RT_RUNNING);
    prio = (idx_prio != nullptr) && (k1[0] > 0.0);
  }
  rt_enqueue(running && !prio, (int)i, idx_out, count_out);
  if (idx_prio != nullptr) rt_enqueue(running && prio, (int)i, idx_prio, count_prio);
}

// ---- the slice kernel -----------------------------------------------------------------------------
template <int MODE, bool COOP, bool SAVE>
__global__ void __launch_bounds__(RT_BLOCK, COOP ? RT_COOP_MINBLOCKS : RT_MINBLOCKS)  // COOP runs a few warps per SM: no register cap
    rt_integrate_kernel(const __grid_constant__ RtDevEnv E, RtDevOpts O, RtState S, const int *__restrict__ idx_in,
                        const int *__restrict__ count_in, int *idx_out, int *count_out, int *count_clear,
                        int attempts) {
  constexpr int DIM = RtTraits<MODE>::DIM;
  extern __shared__ double rt_smem[];
#define RT_K(s, c) rt_smem[((s) * DIM + (c)) * RT_BLOCK + threadIdx.x]
  // The work-list length lives in device memory (written by the previous launch), so the host can enqueue a
  // batch of slices without synchronising; blocks beyond the list exit at once.  Three counters rotate:
  // this launch reads count_in, appends to count_out and clears the one the next launch will append to.
  const int nactive = *count_in;
  if (blockIdx.x == 0 && threadIdx.x == 0) *count_clear = 0;
  // COOP: 3 lanes share one ray, 10 rays per warp, lanes 30 and 31 idle (rt_physics.cuh)
  constexpr int RAYS_PER_BLOCK = COOP ? (RT_BLOCK / 32) * RT_COOP_RAYS_PER_WARP : RT_BLOCK;
  if (blockIdx.x * RAYS_PER_BLOCK >= nactive) return;
  const int lane = threadIdx.x & 31;
  const int j = blockIdx.x * RAYS_PER_BLOCK +
                (COOP ? (threadIdx.x >> 5) * RT_COOP_RAYS_PER_WARP + ((lane * 11) >> 5) : threadIdx.x);
  const bool writer = !COOP || rt_group_rank() == 0;
  const bool valid = j < nactive && (!COOP || lane < 3 * RT_COOP_RAYS_PER_WARP);
  const long long i = valid ? idx_in[j] : 0;
  const size_t n = (size_t)S.n;

  double u[DIM], un[DIM], y[DIM];
  double t = 0.0, h = 0.0, qold = 0.0, par = 0.0;
  int naccept = 0, nreject = 0, nfe = 0, nreflect = 0, sgn_prev = 0, cnt = 0, status = RT_SUCCESS;
  if (valid) {
#pragma unroll
    for (int c = 0; c < DIM; c++) {
      u[c] = S.u[c * n + i];
      RT_K(0, c) = S.k1[c * n + i];
    }
    t = S.t[i];
    h = S.h[i];
    qold = S.qold[i];
    par = S.par[i];
    naccept = S.naccept[i];
    nreject = S.nreject[i];
    nfe = S.nfe[i];
    nreflect = S.nreflect[i];
    sgn_prev = S.sgn[i];
    cnt = S.nsamp[i];
    status = S.status[i];
  }
  bool running = valid && status == RT_RUNNING, hit = false;
  double hit_hi = 1.0;  // upper end of the event bracket in Theta
  // trajectory staging (save_mode): RT_STAGE rows per thread behind the stage derivatives in shared memory
  constexpr int NCOL = RtTraits<MODE>::NCOL;
  double *const stage_warp = rt_smem + 7 * DIM * RT_BLOCK + (threadIdx.x & ~31) * (RT_STAGE * NCOL);
  double *const slot = stage_warp + lane * (RT_STAGE * NCOL);
  int staged = 0;
  bool truncated = false;
  RtExtras2D ex = {0, 0, 0, 0};

  for (int a = 0; a < attempts; a++) {
    if (!__any_sync(0xffffffffu, running)) break;
    if (SAVE) rt_traj_flush<MODE>(S, O, stage_warp, (int)i, cnt, staged, staged == RT_STAGE, writer);
    if (!running) continue;
    if (t >= O.t1) { status = RT_SUCCESS; running = false; continue; }
    if (naccept + nreject >= O.maxiters) { status = RT_MAX_ITERS; running = false; continue; }
    if (h > O.dtmax) h = O.dtmax;
    bool clipped = false;
    {  // tstop handling: a step landing within 100 eps of t1 goes to t1 exactly
      double tt = t + h, at = rt_abs(t), a1 = rt_abs(O.t1);
      double tol = (100.0 * 2.220446049250313e-16) * ((at > a1) ? at : a1);
      if (tt >= O.t1 || O.t1 - tt <= tol) { h = O.t1 - t; clipped = true; }
    }

    // stages 2..7: y = u + h * sum_j a_sj k_j (left to right), k_s = f(y)
    for (int s = 0; s < 6; s++) {
      // one specialised, fully unrolled sum per stage row (all shared-memory loads up front, DIM independent
      // chains) instead of a rolled loop over the row: same left-to-right order, hence the same bits
#define RT_STAGE_ROW(S)                                                      \
  case S: {                                                                  \
    _Pragma("unroll") for (int c = 0; c < DIM; c++) {                        \
      double acc = RT_A[S][0] * RT_K(0, c);                                  \
      _Pragma("unroll") for (int q = 1; q <= S; q++) acc = acc + RT_A[S][q] * RT_K(q, c); \
      y[c] = u[c] + h * acc;                                                 \
    }                                                                        \
  } break;
      switch (s) {
        RT_STAGE_ROW(0)
        RT_STAGE_ROW(1)
        RT_STAGE_ROW(2)
        RT_STAGE_ROW(3)
        RT_STAGE_ROW(4)
        RT_STAGE_ROW(5)
      }
#undef RT_STAGE_ROW
      double dy[DIM];
      rt_rhs<MODE, COOP>(E, y, par, dy, &ex);
#pragma unroll
      for (int c = 0; c < DIM; c++) RT_K(s + 1, c) = dy[c];
    }
#pragma unroll
    for (int c = 0; c < DIM; c++) un[c] = y[c];
    nfe += 6;

    double hnew = h;
    if (O.adaptive) {
      // error norm: branch-free divisions / sqrt first; the exact ones, out of line, when their flag says no
      double ev[DIM], skv[DIM];
      double ssum = 0.0;
      bool okE = true;
#pragma unroll
      for (int c = 0; c < DIM; c++) {
        double acc = RT_BT[0] * RT_K(0, c);
#pragma unroll
        for (int q = 1; q < 7; q++) acc = acc + RT_BT[q] * RT_K(q, c);
        double e = h * acc;
        double a0 = rt_abs(u[c]), a1 = rt_abs(un[c]);
        double sk = O.abstol + ((a0 > a1) ? a0 : a1) * O.reltol;
        ev[c] = e;
        skv[c] = sk;
        bool okq = true;
        double qq = rtf_div(e, sk, okq);
        const bool z = (e == 0.0) && (sk > 0.0);  // +-0 / positive = +-0
        qq = z ? e : qq;
        okE = okE && (okq || z);
        ssum = ssum + qq * qq;
      }
      double EEst = rtf_sqrt(rtf_div(ssum, (double)DIM, okE), okE);
      if (!okE) {
        RtErrIn<DIM> in;
#pragma unroll
        for (int c = 0; c < DIM; c++) {
          in.e[c] = ev[c];
          in.sk[c] = skv[c];
        }
        EEst = rt_eest_exact<DIM>(in);
      }
      if (!(EEst <= 1.0)) {  // reject (NaN included)
        nreject++;
        if (EEst != EEst) {
          if (O.nan_policy == 1) { status = RT_DT_NAN; running = false; continue; }  // OrdinaryDiffEq: dt = NaN -> DtNaN
          h = h * 0.2;
        } else {
          double qq = rt_pow_dev(EEst, 0.14) / 0.9;
          h = h / ((5.0 < qq) ? 5.0 : qq);
        }
        if (h <= O.dtmin) { status = RT_DT_LESS_THAN_MIN; running = false; }
        continue;
      }
      double q = rt_pow_dev(EEst, 0.14) / rt_pow_dev(qold, 0.08);
      q = q / 0.9;
      if (q > 5.0) q = 5.0;
      if (q < 0.1) q = 0.1;
      if (q >= 1.0 && q <= 1.2) q = 1.0;  // qsteady_min <= q <= qsteady_max: the accepted step keeps dt
      qold = (EEst > 1e-4) ? EEst : 1e-4;
      hnew = h / q;
    }
    const double tn = clipped ? O.t1 : t + h;
    naccept++;
    bool finite = true;
#pragma unroll
    for (int c = 0; c < DIM; c++) finite = finite && rt_isfinite(un[c]);
    if (!finite) {
#pragma unroll
      for (int c = 0; c < DIM; c++) u[c] = un[c];
      t = tn;
      status = RT_UNSTABLE;
      running = false;
      continue;
    }
    {  // re_cb: ContinuousCallback on u[1] - re
      double g0 = u[0] - E.re, g1 = un[0] - E.re;
      if ((g0 > 0.0 && g1 <= 0.0) || (g0 < 0.0 && g1 >= 0.0)) {
        // the ray stops here; the root is located after the loop, once per slice for all the lanes of the warp
        // that hit the ground in it (inside the loop, lanes hitting in different attempts would serialise)
        hit = true;
        running = false;
        continue;
      }
      if (g0 != 0.0) {
        // interior safety check, filtered: sum_q b_q(Theta) = Theta, so r(Theta) - re = g0 + h Theta k_1 + E with
        // |E| <= 7.59 |h| max_q |k_q - k_1| (max over Theta of sum_q |b_q| = 7.582); no sign change is possible while
        // |g0| exceeds |h| (|k_1| + 8 sum_q |k_q - k_1|) by more than the evaluation's rounding (1 mm + 1e-12 |r|).
        // Otherwise the eight points are evaluated exactly as the oracle does (which always evaluates them).
        const double k10 = RT_K(0, 0);
        double dev = rt_abs(RT_K(1, 0) - k10);
#pragma unroll
        for (int q = 2; q < 7; q++) dev = dev + rt_abs(RT_K(q, 0) - k10);
        const double reach = rt_abs(h) * (rt_abs(k10) + 8.0 * dev) + (1.0e-3 + 1.0e-12 * rt_abs(u[0]));
        if (!(rt_abs(g0) > reach)) {
          RtScanIn in;
#pragma unroll
          for (int q = 0; q < 7; q++) in.k[q] = RT_K(q, 0);
          in.u0 = u[0]; in.h = h; in.re = E.re; in.g0 = g0;
          const int ip = rt_event_scan(in);
          if (ip > 0) {
            hit = true;
            hit_hi = (double)ip / 9.0;
            running = false;
            continue;
          }
        }
      }
    }
    t = tn;
#pragma unroll
    for (int c = 0; c < DIM; c++) {
      u[c] = un[c];
      RT_K(0, c) = RT_K(6, c);
    }
    int sgn = rt_sign(RT_K(0, 1));
    if (sgn != 0 && sgn_prev != 0 && sgn != sgn_prev) nreflect++;
    if (sgn != 0) sgn_prev = sgn;
    const bool escaped = (O.r_esc > 0.0 && u[0] > O.r_esc);
    if (SAVE) {
      bool last = (t >= O.t1) || escaped;
      if ((naccept % O.save_stride) == 0 || last) rt_traj_push<MODE>(O, slot, cnt, staged, truncated, last, t, u, ex, writer);
    }
    if (escaped) { status = RT_ESCAPED; running = false; continue; }
    h = hnew;
  }

  if (SAVE) rt_traj_flush<MODE>(S, O, stage_warp, (int)i, cnt, staged, staged == RT_STAGE, writer);  // room for the event row
  if (hit) {  // re_cb's root: bisection on the step's dense output (u, h and the seven stages are those of the step)
    const double g0 = u[0] - E.re;
    double lo = 0.0, hi = hit_hi, b[7];
    for (int it = 0; it < 52; it++) {
      double mid = 0.5 * (lo + hi);
      rt_dense_weights(mid, b);
      double sacc = b[0] * RT_K(0, 0);
#pragma unroll
      for (int q = 1; q < 7; q++) sacc = sacc + b[q] * RT_K(q, 0);
      double gm = (u[0] + h * sacc) - E.re;
      bool same = (g0 > 0.0) ? (gm > 0.0) : (gm < 0.0);
      if (same) lo = mid; else hi = mid;
    }
    rt_dense_weights(lo, b);
#pragma unroll
    for (int c = 0; c < DIM; c++) {
      double sacc = b[0] * RT_K(0, c);
#pragma unroll
      for (int q = 1; q < 7; q++) sacc = sacc + b[q] * RT_K(q, c);
      y[c] = u[c] + h * sacc;
    }
    t = t + lo * h;
#pragma unroll
    for (int c = 0; c < DIM; c++) u[c] = y[c];
    status = RT_TERMINATED_EARTH;
    running = false;
    if (SAVE) {
      if (MODE != RT_PRESET_RAY3D) {
        rt_extras_at<(MODE == RT_PRESET_RAY3D) ? RT_PRESET_LAT2D : MODE>(E, u, &ex);
      }
      rt_traj_push<MODE>(O, slot, cnt, staged, truncated, true, t, u, ex, writer);
    }
  }
  if (SAVE) rt_traj_flush<MODE>(S, O, stage_warp, (int)i, cnt, staged, true, writer);

  if (valid && writer) {
#pragma unroll
    for (int c = 0; c < DIM; c++) {
      S.u[c * n + i] = u[c];
      S.k1[c * n + i] = RT_K(0, c);
    }
    S.t[i] = t;
    S.h[i] = h;
    S.qold[i] = qold;
    S.naccept[i] = naccept;
    S.nreject[i] = nreject;
    S.nfe[i] = nfe;
    S.nreflect[i] = nreflect;
    S.sgn[i] = sgn_prev;
    S.nsamp[i] = cnt;
    if (truncated) S.trunc[i] = 1;
    S.status[i] = status;
  }
  rt_enqueue(running && writer, (int)i, idx_out, count_out);
#undef RT_K
}
