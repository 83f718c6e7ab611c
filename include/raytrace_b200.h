/* raytrace_b200.h — C ABI of the B200-native batched whistler ray tracer.
 *
 * Drop-in boundary for the hot path of andersontodds/ray-trace.  The reference has
 * no FFI of its own: its boundary is the set of Julia functions and the SciML solve convention
 * cited beside each entry point below (file:line into the reference tree).  A Julia host binds
 * these symbols with `ccall` (see INTEGRATION.md and julia/RayTraceB200.jl); the Python mirror in
 * ray-trace_b200/api.py binds the same symbols with ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; every array the caller receives back is caller-allocated;
 *   - all functions return 0 on success, a negative rt_error otherwise; rt_last_error() gives text;
 *   - per-ray numerical failures are DATA (status[]), never API errors;
 *   - a context owns one CUDA device, one stream and its device buffers; it is not thread-safe;
 *   - there is NO CPU fallback: without a usable CUDA device rt_create fails with RT_ERR_NO_DEVICE.
 */
#ifndef RAYTRACE_B200_H
#define RAYTRACE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_API __attribute__((visibility("default")))

typedef enum rt_error {
  RT_OK = 0,
  RT_ERR_INVALID = -1,   /* bad argument */
  RT_ERR_NO_DEVICE = -2, /* no CUDA device / driver */
  RT_ERR_CUDA = -3,      /* CUDA runtime error (text in rt_last_error) */
  RT_ERR_NOMEM = -4,
  RT_ERR_STATE = -5      /* call order (e.g. trajectory copy before a solve) */
} rt_error;

/* Variant presets: which reference file's constants, geometry and signs are used. */
typedef enum rt_preset {
  RT_PRESET_LAT2D = 0,   /* RayTrace_lat.jl: theta slot = latitude, CA92 plasmasphere            */
  RT_PRESET_RAY3D = 1,   /* RayTrace_3D.jl: colatitude, refractive-index vector rho, group delay */
  RT_PRESET_RAYMAIN = 2, /* RayMain.jl: colatitude, ionosphere-only density, opposite signs      */
  RT_PRESET_SCRATCH = 3  /* scratch.jl:63-331: colatitude, psi = 3pi/2 - dip - chi, CA92          */
} rt_preset;

/* Per-ray termination class (second table).  Values are part of the ABI. */
typedef enum rt_status {
  RT_RUNNING = 0,          /* never returned to the caller */
  RT_SUCCESS = 1,          /* reached t1 (SciML retcode Success)                                */
  RT_TERMINATED_EARTH = 2, /* u[1]-re crossed zero: re_cb, RayTrace_lat.jl:303-315, _3D.jl:364-372 */
  RT_ESCAPED = 3,          /* [EXT] r > r_esc after an accepted step (disabled when r_esc <= 0)   */
  RT_MAX_ITERS = 4,        /* step attempts exceeded maxiters (retcode MaxIters)                  */
  RT_DT_LESS_THAN_MIN = 5, /* rejected step shrank to <= dtmin (retcode DtLessThanMin)            */
  RT_UNSTABLE = 6,         /* non-finite accepted state (retcode Unstable)                        */
  RT_DT_NAN = 7            /* non-finite initial step (retcode DtNaN), e.g. rho_phi = 0 in 3D     */
} rt_status;

/* Environment: constants RayTrace_lat.jl:13-31 / RayTrace_3D.jl:38-51 and the one-off
 * plasmasphere pre-solve RayTrace_lat.jl:21-41.  Fill with rt_env_preset(). */
typedef struct rt_env_params {
  int32_t preset;        /* rt_preset */
  int32_t density_model; /* 0 = the ray-tracing files' density: CA92(L) + exponential ionosphere (RayTrace_lat.jl:69-83);
                            1 = plasmasphere.jl's composition (:171): diffusive_equilibrium(r, CA92(L)) + ionosphere_eq(r) */
  double c, re, B0, e, me, mp, eps; /* physical constants */
  double Lppi;           /* 5.6 - 0.46 Kp_max */
  double Lppo;           /* argmin |ne2 - ne3| over the L grid */
  double ne_Lppi;        /* plasmasphere density at Lppi, cm^-3 */
  double ca_coef;        /* 0.15(cos..-0.5cos..) + 0.00127 Rbar - 0.0635 (file-specific parenthesisation) */
  double ca_trough;      /* 5800 + 300 mlt */
  double iono_n0;        /* 1.8e5 cm^-3 */
  double iono_k;         /* -4.183119 */
  double iono_h;         /* 1.0471 */
  /* [EXT] finite-difference steps of d(mu)/d(lam|theta) and d(mu)/dr (RayTrace_lat.jl:198-222, RayTrace_3D.jl:223-247).
   * 0 = the files' 1.0e-11: the parity path, on which r -+ 5e-12 == r and d(mu)/dr is identically zero.  > 0 = a step
   * FP64 can resolve (radians / metres, e.g. 1e-6 / 10): physically meaningful gradients, never bit-comparable with the
   * reference; the integrate kernel then runs its exact right-hand side (7 refractive indices instead of 5) */
  double fd_step_angle, fd_step_r;
  /* [EXT] the physically-correct option set (TODOs at RayMain.jl:97-107, RayTrace_3D.jl:22-27,
   * "abs() is not physical" at RayTrace_lat.jl:157,174).  Never on the parity path; LAT2D and RAY3D presets only.
   * deriv_mode: 0 = central differences (the steps above), 1 = d(mu)/dr and d(mu)/d(theta|lambda) analytically (forward-
   *   mode differentiation of the refractive-index code: no step, no cancellation); d(mu)/df stays the 1e-5 difference.
   * root_mode: 0 = mu = sqrt(abs(mu^2_+)) and the d(mu)/d(psi) formula as written; 1 = the propagating root without abs
   *   (mu^2_+ if positive, else mu^2_- if positive, else NaN) with the correct quotient rule for d(mu)/d(psi). */
  int32_t deriv_mode, root_mode;
} rt_env_params;

/* Solver options: the keyword arguments of solve(), RayTrace_lat.jl:338 / RayTrace_3D.jl:395. */
typedef struct rt_solver_opts {
  int32_t adaptive;      /* 1 = Tsit5 + PI controller; 0 = Tsit5(), adaptive=false, dt=dt */
  int32_t save_mode;     /* 0 = terminal state only; 1 = every accepted step (sol.t / sol[i,:]) */
  int32_t max_samples;   /* per-ray row capacity when save_mode = 1 (>= 2; see rt_traj_truncated) */
  int32_t save_stride;   /* keep every save_stride-th accepted step (>= 1); end point always kept */
  int64_t maxiters;      /* step attempts (accepted + rejected); SciML default 100000 */
  double t1;             /* tspan[2]; tspan[1] comes per ray from the launch list */
  double dt;             /* fixed step, or initial step for adaptive (<= 0: automatic) */
  double reltol, abstol; /* 1e-7 / 1e-6 (2D call), 1e-3 / 1e-6 (3D defaults) */
  double dtmax, dtmin;   /* 1e6 / 1e-8 (2D call); <= 0 means t1 - t0 / 0 */
  double r_esc;          /* [EXT] escape radius in metres, <= 0 disables */
  int32_t slice_attempts;/* step attempts per kernel launch between compactions (0 = default 4) */
  int32_t coop_threshold;/* live-ray count at or below which 3 lanes share one ray (0 = default 10000, < 0 = never) */
  int32_t single_chain;  /* 0 = rays that start upward run as a separate, prioritised chain of launches; 1 = one chain */
  int32_t nan_policy;    /* non-finite error estimate of a step attempt: 0 = reject, shrink the step by qmin = 1/5 and retry
                            ([EXT]; ends with DT_LESS_THAN_MIN); 1 = end the ray with RT_DT_NAN at once, which is
                            what OrdinaryDiffEq's reject path does (dt / min(5, NaN) = NaN -> retcode DtNaN) */
} rt_solver_opts;

/* Caller-allocated structure-of-arrays result.  dim = 4 (2D) or 7 (3D). */
typedef struct rt_result {
  double *u_end;         /* [dim][n] terminal state, component-major */
  double *t_end;         /* [n] */
  double *group_delay;   /* [n] 2D: u4_end - f0 (RayTrace_lat.jl:257,273); 3D: u7 (RayTrace_3D.jl:356) */
  int32_t *status;       /* [n] rt_status */
  int32_t *naccept;      /* [n] accepted steps */
  int32_t *nreject;      /* [n] rejected steps */
  int32_t *nfe;          /* [n] haselgrove! evaluations */
  int32_t *nreflect;     /* [n] [EXT] sign changes of d(latitude)/dt between accepted steps */
} rt_result;

typedef struct rt_ctx rt_ctx;

/* ---- context ------------------------------------------------------------------------------ */
RT_API int rt_create(rt_ctx **ctx, int device);
RT_API void rt_destroy(rt_ctx *ctx);
RT_API const char *rt_last_error(const rt_ctx *ctx); /* ctx may be NULL: last global error */
RT_API int rt_set_env(rt_ctx *ctx, const rt_env_params *env);
RT_API int rt_set_solver(rt_ctx *ctx, const rt_solver_opts *opts);
/* use an existing CUDA stream (cudaStream_t as void*; NULL = the context's own stream) */
RT_API int rt_set_stream(rt_ctx *ctx, void *cuda_stream);
RT_API int rt_version(void);

/* ---- environment (host side, one-off) ------------------------------------------------------- */
/* script-top constants + pre-solve of the named file: RayTrace_lat.jl:13-41, RayTrace_3D.jl:38-52 */
RT_API int rt_env_preset(int preset, rt_env_params *env);
RT_API int rt_solver_preset(int preset, rt_solver_opts *opts);
/* initialize_plasmasphere(Lppi, d, Rbar, mlt) -> Lppo, plasmasphere.jl:48-71; also returns ne_Lppi */
RT_API int rt_initialize_plasmasphere(double Lppi, double d, double Rbar, double mlt, double re,
                                      double *Lppo, double *ne_Lppi);

/* ---- batch pointwise evaluators (host arrays, run on the GPU) ------------------------------- */
/* phase_refractive_index(r, lam, chi, freq) -> [mu, dmudpsi], RayTrace_lat.jl:44-194 */
RT_API int rt_phase_refractive_index_2d(rt_ctx *ctx, int64_t n, const double *r, const double *lam,
                                        const double *chi, const double *freq, double *mu,
                                        double *dmudpsi);
/* refractive_index(r,th,ph,rho_r,rho_th,rho_ph,freq) -> [mu, dmudpsi, psi, Bvec], RayTrace_3D.jl:93-219.
 * bvec is [3][n]. */
RT_API int rt_refractive_index_3d(rt_ctx *ctx, int64_t n, const double *r, const double *th,
                                  const double *ph, const double *rho_r, const double *rho_th,
                                  const double *rho_ph, const double *freq, double *mu,
                                  double *dmudpsi, double *psi, double *bvec);
/* haselgrove!(du,u,p,t), RayTrace_lat.jl:251-296: u, du are [4][n] */
RT_API int rt_haselgrove_2d(rt_ctx *ctx, int64_t n, const double *u, double *du);
/* haselgrove!(du,u,p,t), RayTrace_3D.jl:327-359: u, du are [7][n], freq is p[1] per ray */
RT_API int rt_haselgrove_3d(rt_ctx *ctx, int64_t n, const double *u, const double *freq, double *du);

/* plasmasphere.jl as a batch evaluator: L_rlam :122, plasmasphere_eq :73-94, diffusive_equilibrium :96-106,
 * ionosphere_eq :108-117 and ne_total = diffusive_equilibrium(r, plasmasphere_eq(L)) + ionosphere_eq(r) :171.
 * out5 is [5][n]: L, plasmasphere_eq, diffusive_equilibrium, ionosphere_eq, ne_total (cm^-3).  Lppi, Lppo, ne_Lppi,
 * the seasonal bracket and mlt come from the environment (2D presets only). */
RT_API int rt_plasmasphere_density(rt_ctx *ctx, int64_t n, const double *r, const double *lam, double *out5);
/* meridian-plane maps on the grid x[nx] x y[ny] (metres), plasmasphere.jl:157-175 and RayTrace_3D.jl:504-517:
 * out4 = [L_xy, ne_total (diffusive_equilibrium(r, plasmasphere_eq(L)) + ionosphere_eq(r), cm^-3), log10 ne_total,
 * |B| (B0 (re/r)^3 (1 + 3 cos^2(theta)), as the file writes it)][nx*ny], index ix*ny + iy */
RT_API int rt_meridian_maps(rt_ctx *ctx, int64_t nx, int64_t ny, const double *x, const double *y, double *out4);

/* ---- the solve: ODEProblem + solve over a launch list --------------------------------------- */
/* 2D launch list columns (r, theta, chi, f, t) as in README.md:11; RayTrace_lat.jl:333-338 */
RT_API int rt_solve_2d(rt_ctx *ctx, int64_t n, const double *r, const double *theta,
                       const double *chi, const double *f, const double *t0, rt_result *res);
/* 3D launch list (r, theta, phi, rho_r, rho_theta, rho_phi, T0, f, t0); RayTrace_3D.jl:390-395 */
RT_API int rt_solve_3d(rt_ctx *ctx, int64_t n, const double *r, const double *theta,
                       const double *phi, const double *rho_r, const double *rho_theta,
                       const double *rho_phi, const double *T0, const double *f, const double *t0,
                       rt_result *res);
/* Same, with every pointer (launch list and rt_result members) in DEVICE memory; work is enqueued
 * on the context's stream and the call returns after the stream has drained. */
RT_API int rt_solve_2d_device(rt_ctx *ctx, int64_t n, const double *r, const double *theta,
                              const double *chi, const double *f, const double *t0, rt_result *res);
RT_API int rt_solve_3d_device(rt_ctx *ctx, int64_t n, const double *r, const double *theta,
                              const double *phi, const double *rho_r, const double *rho_theta,
                              const double *rho_phi, const double *T0, const double *f,
                              const double *t0, rt_result *res);

/* rt_solve_3d_device_static: rt_solve_3d_device for callers that cannot allocate device memory.
 * No counterpart in the reference: added for callers that may
 * not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments (every pointer in DEVICE memory), same results bit for bit, but the integrator's
 * working set (FSAL derivative, step sizes, work lists, counters: 100 bytes per ray) lives in a
 * file-scope device array of the library instead of cudaMalloc'ed buffers, and the call itself makes
 * no cudaMalloc / cudaFree / allocating library call.  Limits: n <= RT_STATIC_MAX_RAYS (larger n is
 * refused with RT_ERR_INVALID and a message naming the limit); save_mode must be 0 (trajectory rows
 * need the growable buffer of the allocating entry points: RT_ERR_STATE).  Every device that loads the
 * library holds the array (RT_STATIC_SCRATCH_BYTES, ~1.6 MB) until the process exits.  ALL contexts of
 * a device share it and it keeps the previous call's contents (the call clears the counters it needs
 * at zero): two such calls that overlap on one GPU — two contexts, two host threads — race; serialise
 * them.  The call runs one chain of slices on the context's stream and returns after it drained. */
#define RT_STATIC_MAX_RAYS 16384
#define RT_STATIC_SCRATCH_BYTES (256 + 100 * RT_STATIC_MAX_RAYS)
RT_API int rt_solve_3d_device_static(rt_ctx *ctx, int64_t n, const double *r, const double *theta,
                                     const double *phi, const double *rho_r, const double *rho_theta,
                                     const double *rho_phi, const double *T0, const double *f,
                                     const double *t0, rt_result *res);

/* counters and timings of the last solve of a context (rt_get_stats, rt_multi_get_stats) */
typedef struct rt_stats {
  int64_t kernel_launches;   /* kernels launched by the last solve */
  int64_t slices;            /* integrate-kernel launches of the last solve */
  double integrate_ms;       /* CUDA-event time of the integrate kernels of the last solve */
  double total_ms;           /* CUDA-event time of the whole last solve on the stream */
  int64_t total_nfe;         /* reserved, always 0: the evaluation counts are per ray, rt_result.nfe */
  int64_t prio_rays;         /* rays the last solve put on the priority chain */
} rt_stats;

/* ---- the same solve over several GPUs of one box, behind the same boundary ------------------- */
/* One process (the Julia session), one context and one host thread per device.  Rays are independent
 * (RayTrace_3D.jl:327-359 shares nothing between rays), so the launch list is partitioned inside the library —
 * block-cyclically: blocks of `block` consecutive rays go round the devices, which spreads the expensive rays of a
 * sweep (they cluster in launch-parameter space) and lets every column travel as ONE strided copy per device,
 * straight between the caller's arrays and device memory.  Results arrive in the caller's arrays in launch order;
 * there is no device-to-device exchange.  All pointers are HOST pointers (pinned arrays allow the early result copy,
 * see rt_solve_3d).  devices = NULL means 0 .. ngpu-1; ngpu <= 0 means every visible device. */
typedef struct rt_multi rt_multi;
RT_API int rt_multi_create(rt_multi **m, int ngpu, const int *devices);
RT_API void rt_multi_destroy(rt_multi *m);
RT_API int rt_multi_count(const rt_multi *m);
RT_API const char *rt_multi_last_error(const rt_multi *m);
RT_API int rt_multi_set_env(rt_multi *m, const rt_env_params *env);       /* every device */
RT_API int rt_multi_set_solver(rt_multi *m, const rt_solver_opts *opts);  /* every device */
RT_API int rt_multi_set_block(rt_multi *m, int64_t block);                /* 0 = automatic (<= 4096 rays) */
/* ODEProblem + solve over a launch list, RayTrace_lat.jl:333-338 / RayTrace_3D.jl:390-395, on all devices of m */
RT_API int rt_solve_2d_multi(rt_multi *m, int64_t n, const double *r, const double *theta, const double *chi,
                             const double *f, const double *t0, rt_result *res);
RT_API int rt_solve_3d_multi(rt_multi *m, int64_t n, const double *r, const double *theta, const double *phi,
                             const double *rho_r, const double *rho_theta, const double *rho_phi, const double *T0,
                             const double *f, const double *t0, rt_result *res);
/* the partition a solve of n rays uses: block length and rays per device (counts[rt_multi_count]); the launch-list
 * index of device gpu's local ray; the per-device context (for rt_traj_* and rt_get_stats after a solve: trajectories
 * stay on the device that traced them and are indexed by local ray) */
RT_API int rt_multi_partition(rt_multi *m, int64_t n, int64_t *block, int64_t *counts);
RT_API int64_t rt_multi_global_index(rt_multi *m, int64_t n, int gpu, int64_t local);
RT_API rt_ctx *rt_multi_ctx(rt_multi *m, int gpu);
RT_API int rt_multi_get_stats(const rt_multi *m, int gpu, rt_stats *stats);

/* ---- trajectory output of the last solve (save_mode = 1) ------------------------------------- */
/* columns: 2D  t, r, lam, chi, u4, mu, dmudpsi, dip, psi  (RayTrace_lat.jl:318-324,345-349) = 9
 *          3D  t, r, th, ph, rho_r, rho_th, rho_ph, T      (RayTrace_3D.jl:333)             = 8 */
RT_API int rt_traj_columns(const rt_ctx *ctx);
RT_API int rt_traj_size(rt_ctx *ctx, int64_t *total_samples);
/* flags[n]: 1 where a trajectory did not fit max_samples.  Rows 0 .. max_samples-2 take the start and every
 * save_stride-th accepted step; the last row is reserved for the final sample (event point / tspan end / escape),
 * so a trajectory that does not fit loses samples before its end, never the end itself */
RT_API int rt_traj_truncated(rt_ctx *ctx, int32_t *flags);
/* offsets[n+1] (CSR), samples[total][columns] row-major */
RT_API int rt_traj_copy(rt_ctx *ctx, int64_t *offsets, double *samples);
/* the same packed rows and offsets left in DEVICE memory (owned by the context, valid until its next solve or
 * trajectory call): for consumers that stay on the GPU */
RT_API int rt_traj_pack_device(rt_ctx *ctx, const int64_t **d_offsets, const double **d_samples, int64_t *total_samples);
/* what the reference derives from a trajectory after the solve, computed on the device and packed with the same
 * offsets (as returned by rt_traj_copy): 2D  x, y, dt  with x = r cos(lam), y = r sin(lam) (RayTrace_lat.jl:351-352;
 * colatitude presets x = r sin(th), y = r cos(th): RayMain.jl:400-401, scratch.jl:407-408) and
 * dt[i] = t[i+1] - t[i], last 0 (RayTrace_lat.jl:369-372) = 3 columns;
 * 3D  x, y, z, dt  with the spherical-to-Cartesian map of (r, th, ph) = 4 columns */
RT_API int rt_traj_derived_columns(rt_ctx *ctx);
RT_API int rt_traj_derived(rt_ctx *ctx, const int64_t *offsets, double *out);

/* ---- instrumentation ------------------------------------------------------------------------- */
RT_API int rt_get_stats(const rt_ctx *ctx, rt_stats *stats);
/* dependent-free DFMA throughput of this device in FLOP/s (roofline denominator) */
RT_API int rt_measure_fp64_peak(rt_ctx *ctx, double *flops_per_s);


#ifdef __cplusplus
}
#endif
#endif /* RAYTRACE_B200_H */
