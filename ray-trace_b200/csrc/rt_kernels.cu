// rt_kernels.cu — C-ABI implementation (include/raytrace_b200.h) over the sm_100a kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared -Xcompiler -fPIC
// There is no CPU fallback anywhere in this file: every compute entry point launches a kernel.
#include <cuda_runtime.h>
#include <cub/device/device_scan.cuh>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

#include "rt_integrate.cuh"
#ifdef RT_DEV
#include "../../include/raytrace_b200_dev.h"
#endif

#define RT_VERSION_NUM 100

static thread_local std::string g_last_error;

// Developer switches (scheduling experiments, timelines): environment variables, read ONLY by the -DRT_DEV build
// (libraytrace_b200_dev.so, what the diagnostics tests and profiles/ use).  The product library reads no environment
// variable: every switch is its measured default.
#ifdef RT_DEV
static long long dev_env(const char *name, long long dflt) { const char *v = getenv(name); return v ? atoll(v) : dflt; }
static bool dev_flag(const char *name) { return getenv(name) != nullptr; }
#else
static inline long long dev_env(const char *, long long dflt) { return dflt; }
static inline bool dev_flag(const char *) { return false; }
#endif

#define RT_LATE_CAP 65536  // rays that may still be running when the early result copy starts

// Where a context's rays sit in the caller's HOST arrays.  Single-GPU calls: local ray j is element j.  Multi-GPU calls
// (rt_solve_*_multi) partition the launch list block-cyclically: blocks of B consecutive rays go round the N devices, so
// device d's local ray j is element (j / B) (N B) + d B + j % B of the caller's arrays, and every column moves with ONE
// 2D copy (width B elements, pitch N B elements) straight between the caller's arrays and device memory.
struct RtHostLayout {
  long long B;  // block length in rays
  int N, d;     // devices, this device
};
static inline long long rt_to_global(const RtHostLayout *L, long long j) {
  return L ? (j / L->B) * (L->B * (long long)L->N) + (long long)L->d * L->B + (j % L->B) : j;
}
static long long rt_local_count(long long n, long long B, int N, int d) {
  const long long cycle = B * (long long)N, full = n / cycle, rem = n % cycle - (long long)d * B;
  return full * B + (rem < 0 ? 0 : (rem > B ? B : rem));
}
// one column (n_local elements of esz bytes, local order on the device) <-> the caller's host array (global order)
static cudaError_t rt_copy_col(const RtHostLayout *L, void *dev, void *host, size_t esz, long long n_local, bool to_host,
                               cudaStream_t st) {
  const cudaMemcpyKind kind = to_host ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  if (n_local <= 0) return cudaSuccess;
  if (!L) return to_host ? cudaMemcpyAsync(host, dev, (size_t)n_local * esz, kind, st) : cudaMemcpyAsync(dev, host, (size_t)n_local * esz, kind, st);
  const long long full = n_local / L->B, rem = n_local % L->B;
  char *h0 = (char *)host + (size_t)L->d * L->B * esz;
  const size_t w = (size_t)L->B * esz, hp = w * (size_t)L->N;
  cudaError_t e = cudaSuccess;
  if (full > 0)
    e = to_host ? cudaMemcpy2DAsync(h0, hp, dev, w, w, (size_t)full, kind, st) : cudaMemcpy2DAsync(dev, w, h0, hp, w, (size_t)full, kind, st);
  if (e == cudaSuccess && rem > 0) {
    char *hh = h0 + (size_t)full * hp, *dd = (char *)dev + (size_t)full * w;
    e = to_host ? cudaMemcpyAsync(hh, dd, (size_t)rem * esz, kind, st) : cudaMemcpyAsync(dd, hh, (size_t)rem * esz, kind, st);
  }
  return e;
}

struct rt_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;      // stream in use
  cudaStream_t own_stream = nullptr;  // created by rt_create
  rt_env_params env{};
  rt_solver_opts opts{};
  RtDevEnv denv{};
  std::string err;
  // work buffers (device), grown on demand
  size_t cap_rays = 0;
  int cap_dim = 0;
  double *k1 = nullptr, *h = nullptr, *qold = nullptr, *par = nullptr;
  int *sgn = nullptr, *nsamp = nullptr, *trunc = nullptr, *idx[2] = {nullptr, nullptr}, *count = nullptr;
  // trajectory packing (rt_traj_copy / rt_traj_derived): CSR offsets, scan scratch and the packed rows, grown on demand
  long long *pack_off = nullptr;
  size_t pack_off_cap = 0;
  void *scan_tmp = nullptr;
  size_t scan_tmp_bytes = 0;
  double *pack_out = nullptr;
  size_t pack_out_bytes = 0;
  int *idx_prio[2] = {nullptr, nullptr};  // work lists of the priority chains (chain c's counters: count + 4 c .. + 2)
  cudaStream_t prio_stream = nullptr;     // high-priority stream of the priority chain
  cudaStream_t aux_stream[2] = {nullptr, nullptr};  // second halves of the normal / priority list (chains 2 / 3)
  cudaEvent_t ev_chain[4] = {nullptr, nullptr, nullptr, nullptr}, ev_fork = nullptr;
  cudaEvent_t ev_join[3] = {nullptr, nullptr, nullptr};
  // early result copy of the host-pointer entry points (see run_solve): results of the rays already finished go to
  // the caller's pinned arrays while the long rays are still running; those are patched at the end
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_snap[4] = {nullptr, nullptr, nullptr, nullptr}, ev_copy = nullptr;
  int *late_idx = nullptr, *late_i = nullptr, *late_hi = nullptr;  // device list, device / pinned packed int records
  double *late_d = nullptr, *late_hd = nullptr;                    // device / pinned packed double records
  const rt_result *early_host = nullptr;                           // non-null during a host-pointer solve with pinned results
  const RtHostLayout *layout = nullptr;                            // non-null during a multi-GPU solve (block-cyclic host arrays)
  long long n_global = 0;                                          // length of the caller's host arrays during a host-pointer solve
  bool early_done = false;
  long long late_n = 0;
  double *traj = nullptr;
  size_t traj_bytes = 0;
  int *h_count = nullptr;  // pinned
  // staging buffers for the host-pointer API
  double *stage_d = nullptr;
  size_t stage_bytes = 0;
  int *stage_i = nullptr;
  size_t stage_i_bytes = 0;
  // last solve
  long long last_n = 0;
  int last_mode = -1, last_max_samples = 0;
  rt_stats stats{};
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  unsigned char *static_scratch = nullptr;  // this device's address of rt_static_scratch (rt_solve_3d_device_static)
};

static int set_err(rt_ctx *ctx, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  if (ctx) ctx->err = buf;
  return code;
}

#define RT_CUDA(ctx, call)                                                                        \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return set_err(ctx, RT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                     __LINE__);                                                                   \
  } while (0)

// ---------------------------------------------------------------------------------------------------
// Host-side environment: script-top constants and the Lppo pre-solve (RayTrace_lat.jl:13-41).
// Uses the host build of rt_math.h so that the constants are the same bits the kernels would compute.
// ---------------------------------------------------------------------------------------------------
static double host_ca_coef(int preset, double d, double Rbar) {
  if (preset == RT_PRESET_RAY3D) {  // RayTrace_3D.jl:109 parenthesisation
    double c1 = rt_cos((2.0 * RT_PI) * (d + 9.0)) / 365.0;
    double c2 = (0.5 * rt_cos((4.0 * RT_PI) * (d + 9.0))) / 365.0;
    return ((0.15 * (c1 - c2)) + 0.00127 * Rbar) - 0.0635;
  }
  double c1 = rt_cos(((2.0 * RT_PI) * (d + 9.0)) / 365.0);  // RayTrace_lat.jl:73
  double c2 = rt_cos(((4.0 * RT_PI) * (d + 9.0)) / 365.0);
  return ((0.15 * (c1 - 0.5 * c2)) + 0.00127 * Rbar) - 0.0635;
}

extern "C" int rt_initialize_plasmasphere(double Lppi, double d, double Rbar, double mlt, double re, double *Lppo,
                                          double *ne_Lppi) {
  if (!(re > 0.0)) return set_err(nullptr, RT_ERR_INVALID, "rt_initialize_plasmasphere: re must be > 0");
  double coef = host_ca_coef(RT_PRESET_LAT2D, d, Rbar);
  double neL = rt_exp10((-0.3145 * Lppi + 3.9043) + coef * rt_exp((2.0 - Lppi) / 1.5));  // :33
  double span = (10.0 * re - re) / 1000.0;
  long npts = (long)span + 1;  // re:1000:(10*re)
  double best = 0.0, Lbest = 0.0;
  for (long k = 0; k < npts; k++) {
    double r = re + 1000.0 * (double)k;
    double L = r / re;                                                    // cos(0)^2 = 1, :24
    double ne2 = neL * rt_exp10((-1.0 * (L - Lppi)) / 0.1);               // :37
    double ne3 = (5800.0 + 300.0 * mlt) * rt_pow_m4p5(L) + (1.0 - rt_exp((2.0 - L) / 10.0));  // :38
    double a = rt_abs(ne2 - ne3);
    if (k == 0 || a < best) {  // findmin keeps the first minimum, :40
      best = a;
      Lbest = L;
    }
  }
  if (Lppo) *Lppo = Lbest;
  if (ne_Lppi) *ne_Lppi = neL;
  return RT_OK;
}

extern "C" int rt_env_preset(int preset, rt_env_params *env) {
  if (!env || preset < 0 || preset > 3) return set_err(nullptr, RT_ERR_INVALID, "rt_env_preset: bad preset %d", preset);
  memset(env, 0, sizeof(*env));
  env->preset = preset;
  env->c = 2.99792458e8;
  env->re = 6.3712e6;
  env->B0 = (preset == RT_PRESET_RAY3D) ? 3.12e-5 : 3.0696381e-5;  // RayTrace_3D.jl:40 vs RayTrace_lat.jl:15
  env->e = 1.602e-19;
  env->me = 9.1093e-31;
  env->mp = 1.6726219e-27;
  env->eps = 8.854e-12;
  const double Kp_max = 3.0, d = 0.0, Rbar = 90.0, mlt = 2.0;  // RayTrace_lat.jl:26-31
  env->Lppi = 5.6 - 0.46 * Kp_max;
  rt_initialize_plasmasphere(env->Lppi, d, Rbar, mlt, env->re, &env->Lppo, &env->ne_Lppi);
  env->ca_coef = host_ca_coef(preset, d, Rbar);
  env->ca_trough = 5800.0 + 300.0 * mlt;
  env->iono_n0 = 1.8e5;  // RayTrace_lat.jl:70
  env->iono_k = -4.183119;
  env->iono_h = 1.0471;
  return RT_OK;
}

extern "C" int rt_solver_preset(int preset, rt_solver_opts *o) {
  if (!o || preset < 0 || preset > 3) return set_err(nullptr, RT_ERR_INVALID, "rt_solver_preset: bad preset %d", preset);
  memset(o, 0, sizeof(*o));
  o->adaptive = 1;
  o->save_mode = 0;
  o->max_samples = 0;
  o->save_stride = 1;
  o->maxiters = 100000;  // SciML default
  o->t1 = (preset == RT_PRESET_SCRATCH) ? 5.0e10 : 5.0e9;  // scratch.jl:390 / RayTrace_lat.jl:335
  o->dt = 0.0;
  o->abstol = 1e-6;
  if (preset == RT_PRESET_RAY3D) {  // solve(prob, callback=cb): library defaults, RayTrace_3D.jl:395
    o->reltol = 1e-3;
    o->dtmax = 0.0;
    o->dtmin = 0.0;
  } else if (preset == RT_PRESET_LAT2D) {  // RayTrace_lat.jl:338
    o->reltol = 1e-7;
    o->dtmax = 1e6;
    o->dtmin = 1e-8;
  } else {  // RayMain.jl:387, scratch.jl:393: reltol only
    o->reltol = 1e-7;
    o->dtmax = 0.0;
    o->dtmin = 0.0;
  }
  o->r_esc = 0.0;
  o->slice_attempts = 0;
  return RT_OK;
}

// refined reciprocals of the constant divisors, with the division's own Newton sequence (rt_fast.cuh)
__global__ void rt_env_rcp_kernel(RtDevEnv E, double *out) {
  out[0] = rtf_rcp(E.re);
  out[1] = rtf_rcp(E.eps_me);
  out[2] = rtf_rcp(E.eps_mp);
  out[3] = rtf_rcp(E.me);
  out[4] = rtf_rcp(E.mp);
  out[5] = rtf_rcp(1.5);
  out[6] = rtf_rcp(10.0);
  out[7] = rtf_rcp(0.1);
  out[8] = rtf_rcp(1.0e-11);
  out[9] = rtf_rcp(1.0e-5);
}

static void make_dev_env(const rt_env_params &e, RtDevEnv &d) {
  d.c = e.c;
  d.re = e.re;
  d.B0 = e.B0;
  d.e = e.e;
  d.me = e.me;
  d.mp = e.mp;
  d.re3 = (e.re * e.re) * e.re;
  d.e2 = e.e * e.e;
  d.eps_me = e.eps * e.me;
  d.eps_mp = e.eps * e.mp;
  d.Lppi = e.Lppi;
  d.Lppo = e.Lppo;
  d.ne_Lppi = e.ne_Lppi;
  d.ca_coef = e.ca_coef;
  d.ca_trough = e.ca_trough;
  d.iono_n0 = e.iono_n0;
  d.iono_k = e.iono_k;
  d.iono_h = e.iono_h;
  d.inv_c = 1.0 / e.c;
  d.density_model = e.density_model;
  d.h_ang = e.fd_step_angle > 0.0 ? e.fd_step_angle : 1.0e-11;
  d.h_r = e.fd_step_r > 0.0 ? e.fd_step_r : 1.0e-11;
  d.deriv_mode = e.deriv_mode;
  d.root_mode = e.root_mode;
  // the branch-free right-hand side covers the files' own arithmetic only
  d.fd_default = !(e.fd_step_angle > 0.0) && !(e.fd_step_r > 0.0) && e.deriv_mode == 0 && e.root_mode == 0;
}

static int refresh_rcp(rt_ctx *ctx) {
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  double *d = nullptr, h[10];
  RT_CUDA(ctx, cudaMalloc(&d, sizeof(h)));
  rt_env_rcp_kernel<<<1, 1, 0, ctx->stream>>>(ctx->denv, d);
  cudaError_t e = cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d);
  if (e != cudaSuccess) return set_err(ctx, RT_ERR_CUDA, "refresh_rcp: %s", cudaGetErrorString(e));
  RtDevEnv &v = ctx->denv;
  v.y_re = h[0]; v.y_eps_me = h[1]; v.y_eps_mp = h[2]; v.y_me = h[3]; v.y_mp = h[4];
  v.y_1p5 = h[5]; v.y_10 = h[6]; v.y_0p1 = h[7]; v.y_dth = h[8]; v.y_df = h[9];
  return RT_OK;
}

static RtDevOpts make_dev_opts(const rt_solver_opts &o) {
  RtDevOpts d;
  d.adaptive = o.adaptive;
  d.save_mode = o.save_mode && o.max_samples > 0;
  d.max_samples = o.max_samples;
  d.save_stride = o.save_stride > 0 ? o.save_stride : 1;
  long long mi = o.maxiters > 0 ? o.maxiters : 100000;
  d.maxiters = mi > 2000000000LL ? 2000000000 : (int)mi;
  d.nan_policy = o.nan_policy;
  d.t1 = o.t1;
  d.dt = o.dt;
  d.reltol = o.reltol;
  d.abstol = o.abstol;
  d.dtmax = o.dtmax > 0.0 ? o.dtmax : std::numeric_limits<double>::infinity();
  d.dtmin = o.dtmin > 0.0 ? o.dtmin : 0.0;
  d.r_esc = o.r_esc;
  return d;
}

// ---------------------------------------------------------------------------------------------------
// Context
// ---------------------------------------------------------------------------------------------------
extern "C" int rt_version(void) { return RT_VERSION_NUM; }

extern "C" const char *rt_last_error(const rt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

extern "C" int rt_create(rt_ctx **out, int device) {
  if (!out) return set_err(nullptr, RT_ERR_INVALID, "rt_create: null output pointer");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return set_err(nullptr, RT_ERR_NO_DEVICE,
                   "rt_create: no CUDA device available (%s); this library has no CPU fallback",
                   e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0 || device >= ndev) return set_err(nullptr, RT_ERR_INVALID, "rt_create: device %d of %d", device, ndev);
  rt_ctx *ctx = new rt_ctx();
  ctx->device = device;
  RT_CUDA(ctx, cudaSetDevice(device));
  RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
  ctx->stream = ctx->own_stream;
  for (int k = 0; k < 4; k++) RT_CUDA(ctx, cudaEventCreate(&ctx->ev[k]));
  RT_CUDA(ctx, cudaMallocHost(&ctx->h_count, 16 * sizeof(int)));
  RT_CUDA(ctx, cudaMalloc(&ctx->count, 16 * sizeof(int)));
  {
    int lo = 0, hi = 0;
    RT_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    RT_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->prio_stream, cudaStreamNonBlocking, hi));
    RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->aux_stream[0], cudaStreamNonBlocking));
    RT_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->aux_stream[1], cudaStreamNonBlocking, hi));
    for (int k = 0; k < 4; k++) RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_chain[k], cudaEventDisableTiming));
    RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    for (int k = 0; k < 3; k++) RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_join[k], cudaEventDisableTiming));
    RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int k = 0; k < 4; k++) RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_snap[k], cudaEventDisableTiming));
    RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_copy, cudaEventDisableTiming));
    RT_CUDA(ctx, cudaMalloc(&ctx->late_idx, RT_LATE_CAP * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->late_i, RT_LATE_CAP * 5 * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->late_d, RT_LATE_CAP * 9 * sizeof(double)));
    RT_CUDA(ctx, cudaMallocHost(&ctx->late_hi, RT_LATE_CAP * 5 * sizeof(int)));
    RT_CUDA(ctx, cudaMallocHost(&ctx->late_hd, RT_LATE_CAP * 9 * sizeof(double)));
  }
  {  // interior event-check weights b_q(i/9), host-evaluated (rt_integrate.cuh: RT_BI)
    double bi[8][7];
    for (int ip = 0; ip < 8; ip++) rt_dense_weights((double)(ip + 1) / 9.0, bi[ip]);
    RT_CUDA(ctx, cudaMemcpyToSymbol(RT_BI, bi, sizeof(bi)));
  }
  rt_env_preset(RT_PRESET_LAT2D, &ctx->env);
  rt_solver_preset(RT_PRESET_LAT2D, &ctx->opts);
  make_dev_env(ctx->env, ctx->denv);
  if (int rc = refresh_rcp(ctx)) {
    g_last_error = ctx->err;
    rt_destroy(ctx);
    return rc;
  }
  *out = ctx;
  return RT_OK;
}

static void free_work(rt_ctx *ctx) {
  cudaFree(ctx->k1);
  cudaFree(ctx->h);
  cudaFree(ctx->qold);
  cudaFree(ctx->par);
  cudaFree(ctx->sgn);
  cudaFree(ctx->nsamp);
  cudaFree(ctx->trunc);
  ctx->trunc = nullptr;
  cudaFree(ctx->idx[0]);
  cudaFree(ctx->idx[1]);
  cudaFree(ctx->idx_prio[0]);
  cudaFree(ctx->idx_prio[1]);
  ctx->idx_prio[0] = ctx->idx_prio[1] = nullptr;
  ctx->k1 = ctx->h = ctx->qold = ctx->par = nullptr;
  ctx->sgn = ctx->nsamp = ctx->idx[0] = ctx->idx[1] = nullptr;
  ctx->cap_rays = 0;
  ctx->cap_dim = 0;
}

extern "C" void rt_destroy(rt_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  free_work(ctx);
  cudaFree(ctx->traj);
  cudaFree(ctx->pack_off);
  cudaFree(ctx->scan_tmp);
  cudaFree(ctx->pack_out);
  cudaFree(ctx->count);
  cudaFree(ctx->stage_d);
  cudaFree(ctx->stage_i);
  cudaFreeHost(ctx->h_count);
  for (int k = 0; k < 4; k++)
    if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
  for (int k = 0; k < 4; k++)
    if (ctx->ev_chain[k]) cudaEventDestroy(ctx->ev_chain[k]);
  if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
  for (int k = 0; k < 3; k++)
    if (ctx->ev_join[k]) cudaEventDestroy(ctx->ev_join[k]);
  if (ctx->prio_stream) cudaStreamDestroy(ctx->prio_stream);
  for (int k = 0; k < 2; k++)
    if (ctx->aux_stream[k]) cudaStreamDestroy(ctx->aux_stream[k]);
  for (int k = 0; k < 4; k++)
    if (ctx->ev_snap[k]) cudaEventDestroy(ctx->ev_snap[k]);
  if (ctx->ev_copy) cudaEventDestroy(ctx->ev_copy);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  cudaFree(ctx->late_idx);
  cudaFree(ctx->late_i);
  cudaFree(ctx->late_d);
  cudaFreeHost(ctx->late_hi);
  cudaFreeHost(ctx->late_hd);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

extern "C" int rt_set_env(rt_ctx *ctx, const rt_env_params *env) {
  if (!ctx || !env) return set_err(ctx, RT_ERR_INVALID, "rt_set_env: null argument");
  if (env->preset < 0 || env->preset > 3) return set_err(ctx, RT_ERR_INVALID, "rt_set_env: bad preset %d", env->preset);
  if (env->density_model < 0 || env->density_model > 1)
    return set_err(ctx, RT_ERR_INVALID, "rt_set_env: bad density_model %d", env->density_model);
  if (!(env->fd_step_angle >= 0.0) || !(env->fd_step_r >= 0.0))
    return set_err(ctx, RT_ERR_INVALID, "rt_set_env: finite-difference steps must be >= 0 (0 = the reference's 1e-11)");
  if (env->deriv_mode < 0 || env->deriv_mode > 1 || env->root_mode < 0 || env->root_mode > 1)
    return set_err(ctx, RT_ERR_INVALID, "rt_set_env: deriv_mode and root_mode must be 0 or 1");
  if ((env->deriv_mode || env->root_mode) && env->preset != RT_PRESET_LAT2D && env->preset != RT_PRESET_RAY3D)
    return set_err(ctx, RT_ERR_INVALID, "rt_set_env: deriv_mode / root_mode are implemented for the LAT2D and RAY3D presets");
  ctx->env = *env;
  make_dev_env(ctx->env, ctx->denv);
  return refresh_rcp(ctx);
}

extern "C" int rt_set_solver(rt_ctx *ctx, const rt_solver_opts *o) {
  if (!ctx || !o) return set_err(ctx, RT_ERR_INVALID, "rt_set_solver: null argument");
  if (!o->adaptive && !(o->dt > 0.0)) return set_err(ctx, RT_ERR_INVALID, "rt_set_solver: fixed-step mode needs dt > 0");
  if (o->adaptive && !(o->reltol > 0.0 || o->abstol > 0.0))
    return set_err(ctx, RT_ERR_INVALID, "rt_set_solver: adaptive mode needs a tolerance");
  if (o->nan_policy < 0 || o->nan_policy > 1) return set_err(ctx, RT_ERR_INVALID, "rt_set_solver: nan_policy must be 0 or 1");
  if (o->save_mode && o->max_samples < 2)
    return set_err(ctx, RT_ERR_INVALID, "rt_set_solver: save_mode=1 needs max_samples >= 2 (the start and the final sample)");
  ctx->opts = *o;
  return RT_OK;
}

extern "C" int rt_set_stream(rt_ctx *ctx, void *s) {
  if (!ctx) return set_err(ctx, RT_ERR_INVALID, "rt_set_stream: null context");
  ctx->stream = s ? (cudaStream_t)s : ctx->own_stream;
  return RT_OK;
}

extern "C" int rt_get_stats(const rt_ctx *ctx, rt_stats *st) {
  if (!ctx || !st) return set_err(nullptr, RT_ERR_INVALID, "rt_get_stats: null argument");
  *st = ctx->stats;
  return RT_OK;
}

static int ensure_stage(rt_ctx *ctx, size_t dbytes, size_t ibytes) {
  if (dbytes > ctx->stage_bytes) {
    cudaFree(ctx->stage_d);
    ctx->stage_d = nullptr;
    ctx->stage_bytes = 0;
    RT_CUDA(ctx, cudaMalloc(&ctx->stage_d, dbytes));
    ctx->stage_bytes = dbytes;
  }
  if (ibytes > ctx->stage_i_bytes) {
    cudaFree(ctx->stage_i);
    ctx->stage_i = nullptr;
    ctx->stage_i_bytes = 0;
    RT_CUDA(ctx, cudaMalloc(&ctx->stage_i, ibytes));
    ctx->stage_i_bytes = ibytes;
  }
  return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// Batch pointwise evaluators
// ---------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void rt_mu2d_kernel(const __grid_constant__ RtDevEnv E, long long n, const double *r, const double *th, const double *chi,
                               const double *f, double *mu, double *dmu) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (MODE == RT_PRESET_LAT2D && E.root_mode) {  // [EXT] root selection without abs (rt_physics_ad.cuh)
    RtMuAD o;
    rd_mu2d_lat(E, r[i], th[i], chi[i], f[i], o);
    mu[i] = o.mu.v;
    dmu[i] = o.dmudpsi;
    return;
  }
  RtMu m = rt_phase_refractive_index_2d<MODE>(E, r[i], th[i], chi[i], f[i], nullptr, nullptr);
  mu[i] = m.mu;
  dmu[i] = m.dmudpsi;
}

__global__ void rt_mu3d_kernel(const __grid_constant__ RtDevEnv E, long long n, const double *in /*[7][n]*/, double *out /*[6][n]*/) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  RtGeo3D g;
  rt_geo3d(E, in[i], in[n + i], in[3 * n + i], in[4 * n + i], in[5 * n + i], g);
  RtMu m = rt_stix(g.p, (2.0 * RT_PI) * in[6 * n + i], true);
  if (E.root_mode) {  // [EXT] root selection without abs (rt_physics_ad.cuh)
    RtMuAD o;
    rd_mu3d(E, in[i], in[n + i], in[3 * n + i], in[4 * n + i], in[5 * n + i], in[6 * n + i], o);
    m.mu = o.mu.v;
    m.dmudpsi = o.dmudpsi;
  }
  out[i] = m.mu;
  out[n + i] = m.dmudpsi;
  out[2 * n + i] = g.psi;
  out[3 * n + i] = g.Br;
  out[4 * n + i] = g.Bt;
  out[5 * n + i] = g.Bp;
}

template <int MODE>
__global__ void rt_rhs_kernel(const __grid_constant__ RtDevEnv E, long long n, const double *u, const double *freq, double *du) {
  constexpr int DIM = RtTraits<MODE>::DIM;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double y[DIM], dy[DIM];
#pragma unroll
  for (int c = 0; c < DIM; c++) y[c] = u[(size_t)c * n + i];
  double par = (MODE == RT_PRESET_RAY3D) ? freq[i] : y[3];
  rt_rhs<MODE>(E, y, par, dy, nullptr);
#pragma unroll
  for (int c = 0; c < DIM; c++) du[(size_t)c * n + i] = dy[c];
}

static inline unsigned grid_for(long long n, int block) { return (unsigned)((n + block - 1) / block); }

extern "C" int rt_phase_refractive_index_2d(rt_ctx *ctx, int64_t n, const double *r, const double *lam,
                                            const double *chi, const double *freq, double *mu, double *dmudpsi) {
  if (!ctx || n < 0 || (n > 0 && (!r || !lam || !chi || !freq || !mu || !dmudpsi)))
    return set_err(ctx, RT_ERR_INVALID, "rt_phase_refractive_index_2d: bad argument");
  if (ctx->env.preset == RT_PRESET_RAY3D)
    return set_err(ctx, RT_ERR_INVALID, "rt_phase_refractive_index_2d: environment preset is 3D");
  if (n == 0) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  size_t nb = (size_t)n * sizeof(double);
  if (int rc = ensure_stage(ctx, 6 * nb, 0)) return rc;
  double *d = ctx->stage_d;
  const double *src[4] = {r, lam, chi, freq};
  for (int k = 0; k < 4; k++) RT_CUDA(ctx, cudaMemcpyAsync(d + k * n, src[k], nb, cudaMemcpyHostToDevice, ctx->stream));
  unsigned g = grid_for(n, 128);
  switch (ctx->env.preset) {
    case RT_PRESET_LAT2D:
      rt_mu2d_kernel<RT_PRESET_LAT2D><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, d + n, d + 2 * n, d + 3 * n, d + 4 * n, d + 5 * n);
      break;
    case RT_PRESET_RAYMAIN:
      rt_mu2d_kernel<RT_PRESET_RAYMAIN><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, d + n, d + 2 * n, d + 3 * n, d + 4 * n, d + 5 * n);
      break;
    default:
      rt_mu2d_kernel<RT_PRESET_SCRATCH><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, d + n, d + 2 * n, d + 3 * n, d + 4 * n, d + 5 * n);
      break;
  }
  RT_CUDA(ctx, cudaGetLastError());
  RT_CUDA(ctx, cudaMemcpyAsync(mu, d + 4 * n, nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaMemcpyAsync(dmudpsi, d + 5 * n, nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

extern "C" int rt_refractive_index_3d(rt_ctx *ctx, int64_t n, const double *r, const double *th, const double *ph,
                                      const double *rho_r, const double *rho_th, const double *rho_ph,
                                      const double *freq, double *mu, double *dmudpsi, double *psi, double *bvec) {
  if (!ctx || n < 0 || (n > 0 && (!r || !th || !ph || !rho_r || !rho_th || !rho_ph || !freq || !mu || !dmudpsi || !psi || !bvec)))
    return set_err(ctx, RT_ERR_INVALID, "rt_refractive_index_3d: bad argument");
  if (ctx->env.preset != RT_PRESET_RAY3D)
    return set_err(ctx, RT_ERR_INVALID, "rt_refractive_index_3d: environment preset is not RT_PRESET_RAY3D");
  if (n == 0) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  size_t nb = (size_t)n * sizeof(double);
  if (int rc = ensure_stage(ctx, 13 * nb, 0)) return rc;
  double *d = ctx->stage_d;
  const double *src[7] = {r, th, ph, rho_r, rho_th, rho_ph, freq};
  for (int k = 0; k < 7; k++) RT_CUDA(ctx, cudaMemcpyAsync(d + k * n, src[k], nb, cudaMemcpyHostToDevice, ctx->stream));
  rt_mu3d_kernel<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->denv, n, d, d + 7 * n);
  RT_CUDA(ctx, cudaGetLastError());
  double *o = d + 7 * n;
  RT_CUDA(ctx, cudaMemcpyAsync(mu, o, nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaMemcpyAsync(dmudpsi, o + n, nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaMemcpyAsync(psi, o + 2 * n, nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaMemcpyAsync(bvec, o + 3 * n, 3 * nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

// plasmasphere.jl as a batch evaluator: L_rlam :122, plasmasphere_eq :73-94,
// diffusive_equilibrium :96-106 (i = 1, T = 2500 K, alpha = 1), ionosphere_eq :108-117 and the composition
// ne_total = diffusive_equilibrium(r, plasmasphere_eq(L)) + ionosphere_eq(r) :171.  The file's undefined globals are
// taken from the environment: ne_Lppi (:88) = the :61 formula, `t` (:90) = mlt (both inside rt_env_params).
// out = [L, plasmasphere_eq, diffusive_equilibrium, ionosphere_eq, ne_total][n], cm^-3.
__global__ void rt_plasmasphere_kernel(const __grid_constant__ RtDevEnv E, long long n, const double *r, const double *lam,
                                       double *out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double ri = r[i];
  const double c = rt_cos(lam[i]);
  const double L = ri / (E.re * (c * c));
  double ne_eq = 0.0;
  if (L <= E.Lppi) {
    ne_eq = rt_exp10((-0.3145 * L + 3.9043) + E.ca_coef * rt_exp((2.0 - L) / 1.5));
  } else if (E.Lppi < L && L <= E.Lppo) {
    ne_eq = E.ne_Lppi * rt_exp10((E.Lppi - L) / 0.1);
  } else if (E.Lppo < L) {
    ne_eq = E.ca_trough * rt_pow_m4p5(L) + (1.0 - rt_exp((2.0 - L) / 10.0));
  }
  const double T = 2500.0, rb = 7370000.0;
  const double G = rb * (1.0 - rb / ri);
  const double q = rb / 7370.0;
  const double S = ((1.506 * T) * (q * q)) * (1.0 / 1.0);
  const double ne_de = ne_eq * RT_SQRT(1.0 * rt_exp(-G / S));
  const double ne_iono = 1e5 * rt_exp(-10.0 * ((ri / E.re) - 1.0471));
  out[i] = L;
  out[n + i] = ne_eq;
  out[2 * n + i] = ne_de;
  out[3 * n + i] = ne_iono;
  out[4 * n + i] = ne_de + ne_iono;
}

extern "C" int rt_plasmasphere_density(rt_ctx *ctx, int64_t n, const double *r, const double *lam, double *out5) {
  if (!ctx || n < 0 || (n > 0 && (!r || !lam || !out5)))
    return set_err(ctx, RT_ERR_INVALID, "rt_plasmasphere_density: bad argument");
  if (ctx->env.preset == RT_PRESET_RAY3D)
    return set_err(ctx, RT_ERR_INVALID,
                   "rt_plasmasphere_density: plasmasphere.jl uses the RayTrace_lat.jl seasonal term; the environment "
                   "preset is RT_PRESET_RAY3D");
  if (n == 0) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t nb = (size_t)n * sizeof(double);
  if (int rc = ensure_stage(ctx, 7 * nb, 0)) return rc;
  double *d = ctx->stage_d;
  RT_CUDA(ctx, cudaMemcpyAsync(d, r, nb, cudaMemcpyHostToDevice, ctx->stream));
  RT_CUDA(ctx, cudaMemcpyAsync(d + n, lam, nb, cudaMemcpyHostToDevice, ctx->stream));
  rt_plasmasphere_kernel<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->denv, n, d, d + n, d + 2 * n);
  RT_CUDA(ctx, cudaGetLastError());
  RT_CUDA(ctx, cudaMemcpyAsync(out5, d + 2 * n, 5 * nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

// Meridian-plane maps: plasmasphere.jl:157-175 on a grid x[nx] x y[ny] in metres — r_xy :160,
// theta_xy :161, L_xy :164, ne_total = diffusive_equilibrium(r, plasmasphere_eq(L)) + ionosphere_eq(r) :171,
// log10 :174 — and the |B| map of RayTrace_3D.jl:514 (B0 (1/r)^3 (1 + 3 cos cos) as written, r in Earth radii).
// out = [L, ne_total, log10 ne_total, Bmag][nx*ny], index ix*ny + iy.
__global__ void rt_meridian_maps_kernel(const __grid_constant__ RtDevEnv E, long long nx, long long ny, const double *x,
                                        const double *y, double *out) {
  const long long n = nx * ny;
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double xx = x[k / ny], yy = y[k % ny];
  const double r = RT_SQRT(xx * xx + yy * yy);
  const double th = rt_atan(yy / xx);
  const double c = rt_cos(th);
  const double L = (r / E.re) / (c * c);
  double ne_eq = 0.0;
  if (L <= E.Lppi) {
    ne_eq = rt_exp10((-0.3145 * L + 3.9043) + E.ca_coef * rt_exp((2.0 - L) / 1.5));
  } else if (E.Lppi < L && L <= E.Lppo) {
    ne_eq = E.ne_Lppi * rt_exp10((E.Lppi - L) / 0.1);
  } else if (E.Lppo < L) {
    ne_eq = E.ca_trough * rt_pow_m4p5(L) + (1.0 - rt_exp((2.0 - L) / 10.0));
  }
  const double T = 2500.0, rb = 7370000.0;
  const double G = rb * (1.0 - rb / r);
  const double qb = rb / 7370.0;
  const double S = ((1.506 * T) * (qb * qb)) * (1.0 / 1.0);
  const double ne = ne_eq * RT_SQRT(1.0 * rt_exp(-G / S)) + 1e5 * rt_exp(-10.0 * ((r / E.re) - 1.0471));
  const double q = 1.0 / (r / E.re);
  out[k] = L;
  out[n + k] = ne;
  out[2 * n + k] = rt_log10(ne);
  out[3 * n + k] = (E.B0 * ((q * q) * q)) * (1.0 + (3.0 * c) * c);
}

extern "C" int rt_meridian_maps(rt_ctx *ctx, int64_t nx, int64_t ny, const double *x, const double *y, double *out4) {
  if (!ctx || nx < 0 || ny < 0 || (nx * ny > 0 && (!x || !y || !out4)))
    return set_err(ctx, RT_ERR_INVALID, "rt_meridian_maps: bad argument");
  if (ctx->env.preset == RT_PRESET_RAY3D)
    return set_err(ctx, RT_ERR_INVALID,
                   "rt_meridian_maps: plasmasphere.jl uses the RayTrace_lat.jl seasonal term; the environment preset is "
                   "RT_PRESET_RAY3D");
  const int64_t n = nx * ny;
  if (n == 0) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  if (int rc = ensure_stage(ctx, (size_t)(nx + ny + 4 * n) * sizeof(double), 0)) return rc;
  double *d = ctx->stage_d;
  RT_CUDA(ctx, cudaMemcpyAsync(d, x, (size_t)nx * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  RT_CUDA(ctx, cudaMemcpyAsync(d + nx, y, (size_t)ny * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  rt_meridian_maps_kernel<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->denv, nx, ny, d, d + nx, d + nx + ny);
  RT_CUDA(ctx, cudaGetLastError());
  RT_CUDA(ctx, cudaMemcpyAsync(out4, d + nx + ny, (size_t)4 * n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

static int haselgrove_batch(rt_ctx *ctx, int64_t n, const double *u, const double *freq, double *du, bool want3d) {
  if (!ctx || n < 0 || (n > 0 && (!u || !du || (want3d && !freq))))
    return set_err(ctx, RT_ERR_INVALID, "rt_haselgrove: bad argument");
  const int preset = ctx->env.preset;
  if ((preset == RT_PRESET_RAY3D) != want3d)
    return set_err(ctx, RT_ERR_INVALID, "rt_haselgrove: environment preset %d does not match the %s call", preset,
                   want3d ? "3D" : "2D");
  if (n == 0) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  const int dim = want3d ? 7 : 4;
  size_t nb = (size_t)n * sizeof(double);
  if (int rc = ensure_stage(ctx, (2 * dim + 1) * nb, 0)) return rc;
  double *d = ctx->stage_d;
  RT_CUDA(ctx, cudaMemcpyAsync(d, u, dim * nb, cudaMemcpyHostToDevice, ctx->stream));
  if (want3d) RT_CUDA(ctx, cudaMemcpyAsync(d + dim * n, freq, nb, cudaMemcpyHostToDevice, ctx->stream));
  double *dfreq = d + dim * n, *ddu = d + (dim + 1) * n;
  unsigned g = grid_for(n, 128);
  switch (preset) {
    case RT_PRESET_LAT2D: rt_rhs_kernel<RT_PRESET_LAT2D><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, dfreq, ddu); break;
    case RT_PRESET_RAY3D: rt_rhs_kernel<RT_PRESET_RAY3D><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, dfreq, ddu); break;
    case RT_PRESET_RAYMAIN: rt_rhs_kernel<RT_PRESET_RAYMAIN><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, dfreq, ddu); break;
    default: rt_rhs_kernel<RT_PRESET_SCRATCH><<<g, 128, 0, ctx->stream>>>(ctx->denv, n, d, dfreq, ddu); break;
  }
  RT_CUDA(ctx, cudaGetLastError());
  RT_CUDA(ctx, cudaMemcpyAsync(du, ddu, dim * nb, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

extern "C" int rt_haselgrove_2d(rt_ctx *ctx, int64_t n, const double *u, double *du) {
  return haselgrove_batch(ctx, n, u, nullptr, du, false);
}
extern "C" int rt_haselgrove_3d(rt_ctx *ctx, int64_t n, const double *u, const double *freq, double *du) {
  return haselgrove_batch(ctx, n, u, freq, du, true);
}

// ---------------------------------------------------------------------------------------------------
// The solve
// ---------------------------------------------------------------------------------------------------
static int ensure_work(rt_ctx *ctx, size_t n, int dim, const RtDevOpts &O, int ncol) {
  if (n > ctx->cap_rays || dim > ctx->cap_dim) {
    free_work(ctx);
    size_t cap = n;
    RT_CUDA(ctx, cudaMalloc(&ctx->k1, cap * dim * sizeof(double)));
    RT_CUDA(ctx, cudaMalloc(&ctx->h, cap * sizeof(double)));
    RT_CUDA(ctx, cudaMalloc(&ctx->qold, cap * sizeof(double)));
    RT_CUDA(ctx, cudaMalloc(&ctx->par, cap * sizeof(double)));
    RT_CUDA(ctx, cudaMalloc(&ctx->sgn, cap * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->nsamp, cap * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->trunc, cap * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->idx[0], cap * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->idx[1], cap * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->idx_prio[0], cap * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&ctx->idx_prio[1], cap * sizeof(int)));
    ctx->cap_rays = cap;
    ctx->cap_dim = dim;
  }
  size_t tb = O.save_mode ? n * (size_t)O.max_samples * ncol * sizeof(double) : 0;
  if (tb > ctx->traj_bytes) {
    cudaFree(ctx->traj);
    ctx->traj = nullptr;
    ctx->traj_bytes = 0;
    RT_CUDA(ctx, cudaMalloc(&ctx->traj, tb));
    ctx->traj_bytes = tb;
  }
  return RT_OK;
}

__global__ void rt_group_delay_kernel(long long n, const double *u_last, const double *par, int is3d, double *gd) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  gd[i] = is3d ? u_last[i] : (u_last[i] - par[i]);  // RayTrace_3D.jl:356 / RayTrace_lat.jl:257,273
}

// terminal records of the rays listed in late_idx, packed component-major: od[(dim + 2)][m], oi[5][m]
__global__ void rt_late_gather_kernel(int m, int dim, long long n, const int *late_idx, rt_result R, double *od, int *oi) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const long long i = late_idx[j];
  for (int c = 0; c < dim; c++) od[(size_t)c * m + j] = R.u_end[(size_t)c * n + i];
  od[(size_t)dim * m + j] = R.t_end[i];
  od[(size_t)(dim + 1) * m + j] = R.group_delay[i];
  oi[j] = R.status[i];
  oi[(size_t)m + j] = R.naccept[i];
  oi[(size_t)2 * m + j] = R.nreject[i];
  oi[(size_t)3 * m + j] = R.nfe[i];
  oi[(size_t)4 * m + j] = R.nreflect[i];
}

// every result array of the (device) record R to the caller's host arrays H, on stream st
static int copy_results_to_host(rt_ctx *ctx, const rt_result &H, const rt_result &R, long long n, int dim, cudaStream_t st) {
  const RtHostLayout *Y = ctx->layout;
  const size_t ng = (size_t)ctx->n_global;
  for (int c = 0; c < dim; c++)
    RT_CUDA(ctx, rt_copy_col(Y, R.u_end + (size_t)c * n, H.u_end + (size_t)c * ng, sizeof(double), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.t_end, H.t_end, sizeof(double), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.group_delay, H.group_delay, sizeof(double), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.status, H.status, sizeof(int), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.naccept, H.naccept, sizeof(int), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.nreject, H.nreject, sizeof(int), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.nfe, H.nfe, sizeof(int), n, true, st));
  RT_CUDA(ctx, rt_copy_col(Y, R.nreflect, H.nreflect, sizeof(int), n, true, st));
  return RT_OK;
}

template <int MODE>
static int run_solve(rt_ctx *ctx, long long n, const RtLaunch &L, const rt_result &R) {
  constexpr int DIM = RtTraits<MODE>::DIM, NCOL = RtTraits<MODE>::NCOL;
  RtDevOpts O = make_dev_opts(ctx->opts);
  if (int rc = ensure_work(ctx, (size_t)n, DIM, O, NCOL)) return rc;
  RtState S;
  S.n = n;
  S.u = R.u_end;
  S.k1 = ctx->k1;
  S.t = R.t_end;
  S.h = ctx->h;
  S.qold = ctx->qold;
  S.par = ctx->par;
  S.status = R.status;
  S.naccept = R.naccept;
  S.nreject = R.nreject;
  S.nfe = R.nfe;
  S.nreflect = R.nreflect;
  S.sgn = ctx->sgn;
  S.nsamp = ctx->nsamp;
  S.trunc = ctx->trunc;
  S.traj = ctx->traj;
  cudaStream_t st = ctx->stream;
  // stage derivatives k[7][DIM][thread] and, when trajectories are saved, RT_STAGE staged rows per thread
  const size_t smem = (size_t)(7 * DIM + (O.save_mode ? RT_STAGE * NCOL : 0)) * RT_BLOCK * sizeof(double);
  // (kernel variants: one thread / 3 lanes per ray, without / with trajectory staging)
  // developer switch: the long rays' one-thread-per-ray launches ask for more shared memory than they use, so that an
  // SM hosting one of their blocks takes one bulk block beside it instead of two (130 kB) or none (180 kB)
  static const long long prio_smem_kb = dev_env("RT_PRIO_SMEM_KB", 0);
  static const long long prio_smem_rays = dev_env("RT_PRIO_SMEM_RAYS", 19000);
  const size_t smem_prio = (size_t)prio_smem_kb * 1024 > smem ? (size_t)prio_smem_kb * 1024 : smem;
  auto k1 = O.save_mode ? rt_integrate_kernel<MODE, false, true> : rt_integrate_kernel<MODE, false, false>;
  auto k3 = O.save_mode ? rt_integrate_kernel<MODE, true, true> : rt_integrate_kernel<MODE, true, false>;
  RT_CUDA(ctx, cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_prio));
  RT_CUDA(ctx, cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // below this many live rays a chain cannot fill the GPU with one thread per ray: switch to 3 lanes per ray
  const long long coop_threshold = ctx->opts.coop_threshold == 0 ? 10000 : ctx->opts.coop_threshold;
  const int attempts = ctx->opts.slice_attempts > 0 ? ctx->opts.slice_attempts : 4;
  const bool two_chains = ctx->opts.single_chain == 0;
  ctx->stats = rt_stats{};
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
  RT_CUDA(ctx, cudaMemsetAsync(ctx->count, 0, 16 * sizeof(int), st));
  rt_init_kernel<MODE><<<grid_for(n, RT_BLOCK), RT_BLOCK, 0, st>>>(ctx->denv, O, S, L, ctx->idx[0], ctx->count,
                                                                  two_chains ? ctx->idx_prio[0] : nullptr,
                                                                  ctx->count + 4);
  RT_CUDA(ctx, cudaGetLastError());
  ctx->stats.kernel_launches++;
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
  RT_CUDA(ctx, cudaMemcpyAsync(ctx->h_count, ctx->count, sizeof(int), cudaMemcpyDeviceToHost, st));
  RT_CUDA(ctx, cudaMemcpyAsync(ctx->h_count + 1, ctx->count + 4, sizeof(int), cudaMemcpyDeviceToHost, st));
  RT_CUDA(ctx, cudaStreamSynchronize(st));
  // Independent chains of slices.  Chain 0: the normal list, on the caller's stream; chain 1: the priority list
  // (rays that start upward, see rt_init_kernel), on a high-priority stream.  A list of more than a wave or so is
  // cut in two halves that run as separate chains (2: second half of the normal list, 3: of the priority list) so
  // that the partial last wave of one launch is filled by the other chain's blocks (100 k rays are 782 blocks on
  // 444 resident ones: 1.76 waves per launch by themselves).  Within a chain, launch k reads counter k%3, appends
  // survivors to counter (k+1)%3 and clears (k+2)%3, so a batch of slices needs no host round trip; after each
  // batch the host reads the live count (an upper bound for the next batch's grid: a list only shrinks).  The
  // host polls the chains' events and feeds whichever is idle.
  struct Chain {
    cudaStream_t st;
    int *idx[2];
    int *cnt;
    int *h;
    cudaEvent_t ev;
    long long upper;
    int cur, k;
    bool pending;
    int att;            // attempts per launch of the next batch (adaptive, see below)
    long long nominal;  // attempts enqueued so far (diagnostics)
  } ch[4];
  static const long long split_min = dev_env("RT_SPLIT_MIN", 32768);
  {
    cudaStream_t streams[4] = {st, ctx->prio_stream, ctx->aux_stream[0], ctx->aux_stream[1]};
    int *h_init = ctx->h_count + 8;  // first counters of the four chains, uploaded below when a list is cut
    bool cut = false;
    for (int c = 0; c < 2; c++) {
      int *const *base = c ? ctx->idx_prio : ctx->idx;
      const long long total = ctx->h_count[c];
      long long half = total;
      if (two_chains && total >= split_min) half = ((total / 2 + RT_BLOCK - 1) / RT_BLOCK) * RT_BLOCK;
      cut = cut || half < total;
      for (int part = 0; part < 2; part++) {
        const int id = c + 2 * part;
        const long long off = part ? half : 0;
        ch[id] = Chain{streams[id], {base[0] ? base[0] + off : nullptr, base[1] ? base[1] + off : nullptr},
                       ctx->count + 4 * id, ctx->h_count + 4 + id, ctx->ev_chain[id],
                       part ? total - half : half, 0, 0, false, attempts, 0};
        h_init[id] = (int)ch[id].upper;
      }
    }
    if (cut)
      for (int id = 0; id < 4; id++)
        RT_CUDA(ctx, cudaMemcpyAsync(ctx->count + 4 * id, h_init + id, sizeof(int), cudaMemcpyHostToDevice, st));
  }
  RT_CUDA(ctx, cudaEventRecord(ctx->ev_fork, st));
  // Slice length adapts when the caller did not fix it: a batch that lost fewer than 3 % of its rays doubles the
  // next batch's attempts per launch (up to 64) — lists of long-lived rays (iteration-capped 2D sweeps, the
  // magnetospheric paths) need no compaction, and fewer launches mean fewer waits for SM slots; any batch that
  // does lose rays falls back to the base length.
  const bool adapt = ctx->opts.slice_attempts <= 0;
  ctx->stats.prio_rays = ch[1].upper + ch[3].upper;
  for (int c = 1; c < 4; c++)
    if (ch[c].upper > 0) RT_CUDA(ctx, cudaStreamWaitEvent(ch[c].st, ctx->ev_fork, 0));
  const bool dbg = dev_flag("RT_DEBUG");
  const int order[4] = {1, 3, 0, 2};  // priority chains first
  // Early result copy (host-pointer entry points with pinned result arrays, large lists): once few rays are left,
  // each chain — at its next batch boundary — saves its live list and records an event; after the four events the
  // copy stream sends every result array to the caller while the long rays go on; their records are gathered and
  // patched in after the end (solve_common).  Rays finished before a chain's event are final in memory by then.
  const long long early_min = dev_env("RT_EARLY_MIN", 262144);
  const bool early = ctx->early_host != nullptr && n >= early_min;
  // (late rays are patched in by a host scatter of 14 cache lines each, 0.13 us per ray: the snapshot waits until
  // n / 128 rays are left — on cfg3 that is ~4 ms before the end, enough for its 92 MB; measured e2e 56.5 ms with
  // n / 512 or no early copy at all, 55.5 ms with n / 128, 56.0 with n / 32: profiles/r2_early_copy_matrix.log)
  static const long long early_div = dev_env("RT_EARLY_DIV", 128);
  const long long early_thr = (n / early_div < 1024) ? 1024 : ((n / early_div > RT_LATE_CAP) ? RT_LATE_CAP : n / early_div);
  bool early_trigger = false, snap[4] = {false, false, false, false};
  int nsnap = 0;
  ctx->early_done = false;
  ctx->late_n = 0;
  for (;;) {
    int npending = 0, nactive = 0, last_pending = -1;
    for (int oi = 0; oi < 4; oi++) {
      const int c = order[oi];
      Chain &C = ch[c];
      if (C.pending) {
        cudaError_t q = cudaEventQuery(C.ev);
        if (q == cudaSuccess) {
          C.pending = false;
          const long long before = C.upper;
          C.upper = *C.h;
          if (adapt) C.att = (C.upper * 100 >= before * 97) ? (C.att * 2 > 64 ? 64 : C.att * 2) : attempts;
          // developer switch: a priority chain past its first `prio_keep` attempts keeps long slices while its rays die
          // off (few blocks: their idle lanes cost little, every launch of theirs waits for SM slots)
          static const long long prio_keep = dev_env("RT_PRIO_KEEP_ATT", 0);
          if (adapt && prio_keep > 0 && (c & 1) && C.nominal >= prio_keep && C.att < 64) C.att = 64;
          if (dbg) {
            float ms = 0;
            cudaEventRecord(ctx->ev[2], C.st);
            cudaEventSynchronize(ctx->ev[2]);
            cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]);
            fprintf(stderr, "[rt] chain %d after %d slices (%lld attempts): active=%lld t=%.2f ms\n", c, C.k, C.nominal,
                    C.upper, ms);
          }
        } else if (q != cudaErrorNotReady) {
          return set_err(ctx, RT_ERR_CUDA, "integrate chain %d: %s", c, cudaGetErrorString(q));
        }
      }
      if (early && !ctx->early_done && !C.pending) {
        if (!early_trigger && ch[0].upper + ch[1].upper + ch[2].upper + ch[3].upper <= early_thr) early_trigger = true;
        if (early_trigger && !snap[c]) {
          if (C.upper > 0)
            RT_CUDA(ctx, cudaMemcpyAsync(ctx->late_idx + ctx->late_n, C.idx[C.cur], (size_t)C.upper * sizeof(int),
                                         cudaMemcpyDeviceToDevice, C.st));
          ctx->late_n += C.upper;
          RT_CUDA(ctx, cudaEventRecord(ctx->ev_snap[c], C.st));
          snap[c] = true;
          if (++nsnap == 4) {
            cudaStream_t cs = ctx->copy_stream;
            const rt_result &H = *ctx->early_host;
            for (int q = 0; q < 4; q++) RT_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_snap[q], 0));
            rt_group_delay_kernel<<<grid_for(n, 256), 256, 0, cs>>>(n, R.u_end + (size_t)(DIM - 1) * n, ctx->par,
                                                                    MODE == RT_PRESET_RAY3D, R.group_delay);
            ctx->stats.kernel_launches++;
            if (int rc = copy_results_to_host(ctx, H, R, n, DIM, cs)) return rc;
            RT_CUDA(ctx, cudaEventRecord(ctx->ev_copy, cs));
            ctx->early_done = true;
            if (dbg) fprintf(stderr, "[rt] early result copy enqueued, %lld rays still running\n", ctx->late_n);
          }
        }
      }
      if (!C.pending && C.upper > 0) {
        // 3 lanes per ray (rt_physics.cuh) buys latency (19 us per attempt instead of 32 with the GPU to itself)
        // with throughput: a cooperative block holds 40 rays in the registers of 128, and next to full lists it was
        // measured to cost 2.7x the SM time per attempt of one thread per ray (profiles/r1_schedule_experiments.md),
        // which is exactly what it saves at the end -- so it is used once a chain is down to its long rays AND all
        // chains together leave SMs idle (<= 35 k rays, i.e. 270 of the 444 resident blocks).
        // RT_COOP_POLICY=0 (developer switch) drops the second condition.
        const long long all = ch[0].upper + ch[1].upper + ch[2].upper + ch[3].upper;
        static const int policy = (int)dev_env("RT_COOP_POLICY", 1);
        static const long long coop_sum = dev_env("RT_COOP_SUM", 25000);
        const bool coop = C.upper <= coop_threshold && (policy == 0 || all <= coop_threshold + coop_sum);
        // a launch of the priority chain has to wait for blocks of the other chain to retire before it gets SM
        // slots (stream priority does not preempt), so its slices are long: fewer waits per attempt
        static const int coop_att = (int)dev_env("RT_COOP_ATT", 32);
        const int att = coop ? (C.att > coop_att ? C.att : coop_att) : C.att;
        // a short list still on one thread per ray is waiting for the other chains to drain: keep its batches short
        // (<= 64 attempts) so that the switch to 3 lanes per ray is not delayed by a batch already enqueued
        static const int wait_att = (int)dev_env("RT_WAIT_ATT", 64);
        const int batch = coop ? 4 : ((C.upper <= coop_threshold && att * 8 > wait_att) ? (wait_att / att > 1 ? wait_att / att : 1) : 8);
        C.nominal += (long long)att * batch;
        for (int b = 0; b < batch; b++, C.k++) {
          const int k = C.k;
          if (coop)
            k3<<<grid_for(C.upper, (RT_BLOCK / 32) * RT_COOP_RAYS_PER_WARP), RT_BLOCK, smem, C.st>>>(
                ctx->denv, O, S, C.idx[C.cur], C.cnt + (k % 3), C.idx[C.cur ^ 1], C.cnt + ((k + 1) % 3),
                C.cnt + ((k + 2) % 3), att);
          else
            k1<<<grid_for(C.upper, RT_BLOCK), RT_BLOCK,
                 ((c & 1) && ch[1].upper + ch[3].upper <= prio_smem_rays) ? smem_prio : smem, C.st>>>(
                ctx->denv, O, S, C.idx[C.cur], C.cnt + (k % 3), C.idx[C.cur ^ 1], C.cnt + ((k + 1) % 3),
                C.cnt + ((k + 2) % 3), att);
          C.cur ^= 1;
          ctx->stats.kernel_launches++;
          ctx->stats.slices++;
        }
        RT_CUDA(ctx, cudaGetLastError());
        RT_CUDA(ctx, cudaMemcpyAsync(C.h, C.cnt + (C.k % 3), sizeof(int), cudaMemcpyDeviceToHost, C.st));
        RT_CUDA(ctx, cudaEventRecord(C.ev, C.st));
        C.pending = true;
      }
      if (C.pending) {
        npending++;
        last_pending = c;
      }
      if (C.pending || C.upper > 0) nactive++;
    }
    if (npending == 0) break;
    if (nactive == 1)
      cudaEventSynchronize(ch[last_pending].ev);  // one chain left: block on it
    else
      std::this_thread::yield();  // several busy: poll
  }
  for (int c = 1; c < 4; c++)
    if (ch[c].k > 0) {  // join the other streams back into the caller's stream
      RT_CUDA(ctx, cudaEventRecord(ctx->ev_join[c - 1], ch[c].st));
      RT_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_join[c - 1], 0));
    }
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[2], st));
  if (ctx->early_done) RT_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_copy, 0));  // the early copy reads group_delay
  rt_group_delay_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, R.u_end + (size_t)(DIM - 1) * n, ctx->par,
                                                          MODE == RT_PRESET_RAY3D, R.group_delay);
  RT_CUDA(ctx, cudaGetLastError());
  ctx->stats.kernel_launches++;
  if (ctx->early_done && ctx->late_n > 0) {
    const int m = (int)ctx->late_n;
    rt_late_gather_kernel<<<grid_for(m, 128), 128, 0, st>>>(m, DIM, n, ctx->late_idx, R, ctx->late_d, ctx->late_i);
    RT_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches++;
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->late_hd, ctx->late_d, (size_t)m * (DIM + 2) * sizeof(double), cudaMemcpyDeviceToHost, st));
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->late_hi, ctx->late_i, (size_t)m * 5 * sizeof(int), cudaMemcpyDeviceToHost, st));
  }
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[3], st));
  RT_CUDA(ctx, cudaStreamSynchronize(st));
  float ms_int = 0, ms_tot = 0;
  RT_CUDA(ctx, cudaEventElapsedTime(&ms_int, ctx->ev[1], ctx->ev[2]));
  RT_CUDA(ctx, cudaEventElapsedTime(&ms_tot, ctx->ev[0], ctx->ev[3]));
  ctx->stats.integrate_ms = ms_int;
  ctx->stats.total_ms = ms_tot;
  ctx->last_n = n;
  ctx->last_mode = MODE;
  ctx->last_max_samples = O.save_mode ? O.max_samples : 0;
  return RT_OK;
}

static int dispatch_solve(rt_ctx *ctx, long long n, const RtLaunch &L, const rt_result &R) {
  switch (ctx->env.preset) {
    case RT_PRESET_LAT2D: return run_solve<RT_PRESET_LAT2D>(ctx, n, L, R);
    case RT_PRESET_RAY3D: return run_solve<RT_PRESET_RAY3D>(ctx, n, L, R);
    case RT_PRESET_RAYMAIN: return run_solve<RT_PRESET_RAYMAIN>(ctx, n, L, R);
    default: return run_solve<RT_PRESET_SCRATCH>(ctx, n, L, R);
  }
}

static bool result_ok(const rt_result *r) {
  return r && r->u_end && r->t_end && r->group_delay && r->status && r->naccept && r->nreject && r->nfe && r->nreflect;
}

// n = this context's rays.  Host-pointer calls: cols / res are the caller's arrays of n_global elements and `layout`
// says which of them are this context's (null: all, n_global == n).
static int solve_common(rt_ctx *ctx, int64_t n, const double *const *cols, int ncols, rt_result *res, bool want3d,
                        bool on_device, const RtHostLayout *layout = nullptr, int64_t n_global = -1) {
  if (n_global < 0) n_global = n;
  if (!ctx) return set_err(ctx, RT_ERR_INVALID, "rt_solve: null context");
  if (n < 0 || n > 2000000000LL) return set_err(ctx, RT_ERR_INVALID, "rt_solve: n out of range");
  if ((ctx->env.preset == RT_PRESET_RAY3D) != want3d)
    return set_err(ctx, RT_ERR_INVALID, "rt_solve: environment preset %d does not match the %s call", ctx->env.preset,
                   want3d ? "3D" : "2D");
  if (n == 0) {
    ctx->last_n = 0;
    ctx->stats = rt_stats{};
    return RT_OK;
  }
  if (!result_ok(res)) return set_err(ctx, RT_ERR_INVALID, "rt_solve: rt_result has null members");
  for (int k = 0; k < ncols; k++)
    if (!cols[k]) return set_err(ctx, RT_ERR_INVALID, "rt_solve: launch-list column %d is null", k);
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  const int dim = want3d ? 7 : 4;
  RtLaunch L{};
  ctx->layout = nullptr;
  ctx->n_global = n;
  if (on_device) {
    for (int k = 0; k < ncols; k++) L.col[k] = cols[k];
    return dispatch_solve(ctx, n, L, *res);
  }
  // host buffers: stage launch list and results in device memory
  const size_t nb = (size_t)n * sizeof(double);
  if (int rc = ensure_stage(ctx, (size_t)(ncols + dim + 2) * nb, (size_t)5 * n * sizeof(int))) return rc;
  double *d = ctx->stage_d;
  ctx->layout = layout;
  ctx->n_global = n_global;
  for (int k = 0; k < ncols; k++) {
    RT_CUDA(ctx, rt_copy_col(layout, d + (size_t)k * n, const_cast<double *>(cols[k]), sizeof(double), n, false, ctx->stream));
    L.col[k] = d + (size_t)k * n;
  }
  rt_result D;
  D.u_end = d + (size_t)ncols * n;
  D.t_end = D.u_end + (size_t)dim * n;
  D.group_delay = D.t_end + n;
  D.status = ctx->stage_i;
  D.naccept = D.status + n;
  D.nreject = D.naccept + n;
  D.nfe = D.nreject + n;
  D.nreflect = D.nfe + n;
  // pinned result arrays allow the early asynchronous copy (pageable ones would block the scheduling loop)
  const void *hp[8] = {res->u_end, res->t_end, res->group_delay, res->status, res->naccept, res->nreject, res->nfe, res->nreflect};
  bool pinned = true;
  for (int k = 0; k < 8 && pinned; k++) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, hp[k]) != cudaSuccess) {
      cudaGetLastError();
      pinned = false;
    } else {
      pinned = a.type == cudaMemoryTypeHost;
    }
  }
  ctx->early_host = pinned ? res : nullptr;
  const int rc_solve = dispatch_solve(ctx, n, L, D);
  ctx->early_host = nullptr;
  if (rc_solve) return rc_solve;
  cudaStream_t st = ctx->stream;
  if (ctx->early_done) {
    // everything but the rays that were still running at the snapshot is already in the caller's arrays
    RT_CUDA(ctx, cudaEventSynchronize(ctx->ev_copy));
    const size_t m = (size_t)ctx->late_n;
    std::vector<int> li(m);
    if (m) RT_CUDA(ctx, cudaMemcpy(li.data(), ctx->late_idx, m * sizeof(int), cudaMemcpyDeviceToHost));
    int *ints[5] = {res->status, res->naccept, res->nreject, res->nfe, res->nreflect};
    const size_t ng = (size_t)n_global;
    for (size_t j = 0; j < m; j++) {
      const size_t i = (size_t)rt_to_global(layout, li[j]);
      for (int c = 0; c < dim; c++) res->u_end[c * ng + i] = ctx->late_hd[(size_t)c * m + j];
      res->t_end[i] = ctx->late_hd[(size_t)dim * m + j];
      res->group_delay[i] = ctx->late_hd[(size_t)(dim + 1) * m + j];
      for (int q = 0; q < 5; q++) ints[q][i] = ctx->late_hi[(size_t)q * m + j];
    }
    return RT_OK;
  }
  if (int rc = copy_results_to_host(ctx, *res, D, n, dim, st)) return rc;
  RT_CUDA(ctx, cudaStreamSynchronize(st));
  return RT_OK;
}

extern "C" int rt_solve_2d(rt_ctx *ctx, int64_t n, const double *r, const double *theta, const double *chi,
                           const double *f, const double *t0, rt_result *res) {
  const double *cols[5] = {r, theta, chi, f, t0};
  return solve_common(ctx, n, cols, 5, res, false, false);
}
extern "C" int rt_solve_2d_device(rt_ctx *ctx, int64_t n, const double *r, const double *theta, const double *chi,
                                  const double *f, const double *t0, rt_result *res) {
  const double *cols[5] = {r, theta, chi, f, t0};
  return solve_common(ctx, n, cols, 5, res, false, true);
}
extern "C" int rt_solve_3d(rt_ctx *ctx, int64_t n, const double *r, const double *theta, const double *phi,
                           const double *rho_r, const double *rho_theta, const double *rho_phi, const double *T0,
                           const double *f, const double *t0, rt_result *res) {
  const double *cols[9] = {r, theta, phi, rho_r, rho_theta, rho_phi, T0, f, t0};
  return solve_common(ctx, n, cols, 9, res, true, false);
}
extern "C" int rt_solve_3d_device(rt_ctx *ctx, int64_t n, const double *r, const double *theta, const double *phi,
                                  const double *rho_r, const double *rho_theta, const double *rho_phi,
                                  const double *T0, const double *f, const double *t0, rt_result *res) {
  const double *cols[9] = {r, theta, phi, rho_r, rho_theta, rho_phi, T0, f, t0};
  return solve_common(ctx, n, cols, 9, res, true, true);
}

// ---------------------------------------------------------------------------------------------------
// The solve without allocations (rt_solve_3d_device_static; no counterpart in the reference, see the header): the
// working set that ensure_work() cudaMallocs lives in this file-scope device array instead.  Layout for
// N = RT_STATIC_MAX_RAYS (every part starts on a 256-byte boundary):
//   [0, 64)        work-list counters (3 rotate, see run_solve), cleared by every call
//   256            k1[7][n] FSAL derivative (56 N bytes reserved)
//   256 + 56 N     h[N], qold[N], par[N]
//   256 + 80 N     sgn[N], nsamp[N], trunc[N], work lists idx0[N], idx1[N]           total 256 + 100 N
// Everything but the counters is written by rt_init_kernel before it is read, so stale contents of an earlier call are
// harmless.  The array is per DEVICE and shared by all contexts of that device: overlapping calls race (header).
// ---------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char rt_static_scratch[RT_STATIC_SCRATCH_BYTES];

// Set-up, once per context, with the context's device current.  In host code the array's name is only a stand-in for
// the symbol (not a device address): kernels get the address cudaGetSymbolAddress returns for the current device.
static int static_scratch_ready(rt_ctx *ctx) {
  if (ctx->static_scratch) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  void *p = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&p, rt_static_scratch);
  if (e != cudaSuccess || p == nullptr)
    return set_err(ctx, RT_ERR_CUDA, "rt_solve_3d_device_static: cudaGetSymbolAddress(rt_static_scratch) on device %d: %s",
                   ctx->device, cudaGetErrorString(e));
  ctx->static_scratch = (unsigned char *)p;
  return RT_OK;
}

extern "C" int rt_solve_3d_device_static(rt_ctx *ctx, int64_t n, const double *r, const double *theta, const double *phi,
                                         const double *rho_r, const double *rho_theta, const double *rho_phi,
                                         const double *T0, const double *f, const double *t0, rt_result *res) {
  constexpr int MODE = RT_PRESET_RAY3D, DIM = RtTraits<MODE>::DIM;
  static_assert(RT_STATIC_SCRATCH_BYTES == 256 + (size_t)RT_STATIC_MAX_RAYS * (DIM * 8 + 3 * 8 + 5 * 4), "scratch layout");
  static_assert(RT_STATIC_MAX_RAYS % 64 == 0, "256-byte alignment of the parts");
  if (!ctx) return set_err(ctx, RT_ERR_INVALID, "rt_solve_3d_device_static: null context");
  if (n < 0) return set_err(ctx, RT_ERR_INVALID, "rt_solve_3d_device_static: n out of range");
  if (n > RT_STATIC_MAX_RAYS)
    return set_err(ctx, RT_ERR_INVALID,
                   "rt_solve_3d_device_static: n = %lld exceeds RT_STATIC_MAX_RAYS = %d, the capacity of the static "
                   "scratch array (use rt_solve_3d_device, which allocates, or split the launch list)",
                   (long long)n, (int)RT_STATIC_MAX_RAYS);
  if (ctx->env.preset != RT_PRESET_RAY3D)
    return set_err(ctx, RT_ERR_INVALID, "rt_solve: environment preset %d does not match the 3D call", ctx->env.preset);
  const RtDevOpts O = make_dev_opts(ctx->opts);
  if (O.save_mode)
    return set_err(ctx, RT_ERR_STATE, "rt_solve_3d_device_static: save_mode must be 0 (trajectory rows need the growable "
                                      "buffer of rt_solve_3d_device)");
  if (n == 0) {
    ctx->last_n = 0;
    ctx->stats = rt_stats{};
    return RT_OK;
  }
  if (!result_ok(res)) return set_err(ctx, RT_ERR_INVALID, "rt_solve: rt_result has null members");
  const double *cols[9] = {r, theta, phi, rho_r, rho_theta, rho_phi, T0, f, t0};
  RtLaunch L{};
  for (int k = 0; k < 9; k++) {
    if (!cols[k]) return set_err(ctx, RT_ERR_INVALID, "rt_solve: launch-list column %d is null", k);
    L.col[k] = cols[k];
  }
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  if (int rc = static_scratch_ready(ctx)) return rc;
  const rt_result &R = *res;
  constexpr size_t N = RT_STATIC_MAX_RAYS;
  int *cnt = (int *)ctx->static_scratch;
  RtState S;
  S.n = n;
  S.u = R.u_end;
  S.k1 = (double *)(ctx->static_scratch + 256);
  S.t = R.t_end;
  S.h = S.k1 + DIM * N;
  S.qold = S.h + N;
  S.par = S.qold + N;
  S.status = R.status;
  S.naccept = R.naccept;
  S.nreject = R.nreject;
  S.nfe = R.nfe;
  S.nreflect = R.nreflect;
  S.sgn = (int *)(S.par + N);
  S.nsamp = S.sgn + N;
  S.trunc = S.nsamp + N;
  S.traj = nullptr;
  int *idx[2] = {S.trunc + N, S.trunc + 2 * N};
  cudaStream_t st = ctx->stream;
  const size_t smem = (size_t)(7 * DIM) * RT_BLOCK * sizeof(double);
  auto k1 = rt_integrate_kernel<MODE, false, false>;
  auto k3 = rt_integrate_kernel<MODE, true, false>;
  RT_CUDA(ctx, cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  RT_CUDA(ctx, cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long coop_threshold = ctx->opts.coop_threshold == 0 ? 10000 : ctx->opts.coop_threshold;
  const int attempts = ctx->opts.slice_attempts > 0 ? ctx->opts.slice_attempts : 4;
  const bool adapt = ctx->opts.slice_attempts <= 0;
  ctx->stats = rt_stats{};
  ctx->last_mode = -1;  // no trajectories to fetch after this call
  ctx->last_max_samples = 0;
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
  RT_CUDA(ctx, cudaMemsetAsync(cnt, 0, 16 * sizeof(int), st));  // the array keeps the previous call's counters
  rt_init_kernel<MODE><<<grid_for(n, RT_BLOCK), RT_BLOCK, 0, st>>>(ctx->denv, O, S, L, idx[0], cnt, nullptr, cnt + 4);
  RT_CUDA(ctx, cudaGetLastError());
  ctx->stats.kernel_launches++;
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
  // one chain of slices (run_solve's rotation of three counters); n bounds the list until the first count is read
  long long upper = n;
  int cur = 0, k = 0, att = attempts;
  while (upper > 0) {
    const bool coop = upper <= coop_threshold;  // 3 lanes per ray once the list cannot fill the GPU
    const int a = coop ? (att > 32 ? att : 32) : att;
    const int batch = coop ? 4 : 8;
    for (int b = 0; b < batch; b++, k++) {
      if (coop)
        k3<<<grid_for(upper, (RT_BLOCK / 32) * RT_COOP_RAYS_PER_WARP), RT_BLOCK, smem, st>>>(
            ctx->denv, O, S, idx[cur], cnt + (k % 3), idx[cur ^ 1], cnt + ((k + 1) % 3), cnt + ((k + 2) % 3), a);
      else
        k1<<<grid_for(upper, RT_BLOCK), RT_BLOCK, smem, st>>>(ctx->denv, O, S, idx[cur], cnt + (k % 3), idx[cur ^ 1],
                                                             cnt + ((k + 1) % 3), cnt + ((k + 2) % 3), a);
      cur ^= 1;
      ctx->stats.kernel_launches++;
      ctx->stats.slices++;
    }
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->h_count, cnt + (k % 3), sizeof(int), cudaMemcpyDeviceToHost, st));  // pinned (rt_create)
    RT_CUDA(ctx, cudaStreamSynchronize(st));
    const long long before = upper;
    upper = ctx->h_count[0];
    if (adapt) att = (upper * 100 >= before * 97) ? (att * 2 > 64 ? 64 : att * 2) : attempts;
  }
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[2], st));
  rt_group_delay_kernel<<<grid_for(n, 256), 256, 0, st>>>(n, R.u_end + (size_t)(DIM - 1) * n, S.par, 1, R.group_delay);
  RT_CUDA(ctx, cudaGetLastError());
  ctx->stats.kernel_launches++;
  RT_CUDA(ctx, cudaEventRecord(ctx->ev[3], st));
  RT_CUDA(ctx, cudaStreamSynchronize(st));
  float ms_int = 0, ms_tot = 0;
  RT_CUDA(ctx, cudaEventElapsedTime(&ms_int, ctx->ev[1], ctx->ev[2]));
  RT_CUDA(ctx, cudaEventElapsedTime(&ms_tot, ctx->ev[0], ctx->ev[3]));
  ctx->stats.integrate_ms = ms_int;
  ctx->stats.total_ms = ms_tot;
  ctx->last_n = n;
  return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// Multi-GPU behind the boundary: one process, one context and one host thread per device, the launch
// list partitioned block-cyclically inside the library, terminal records delivered straight into the caller's arrays.
// Rays are independent (RayTrace_3D.jl:327-359 shares nothing between rays): no device-to-device exchange at all.
// ---------------------------------------------------------------------------------------------------
struct rt_multi {
  std::vector<rt_ctx *> ctx;
  std::vector<RtHostLayout> lay;
  std::vector<long long> nloc;
  std::string err;
  long long block = 0;  // 0 = automatic
};

extern "C" int rt_multi_create(rt_multi **out, int ngpu, const int *devices) {
  if (!out) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_create: null output pointer");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return set_err(nullptr, RT_ERR_NO_DEVICE, "rt_multi_create: no CUDA device available (%s); this library has no CPU fallback",
                   e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (ngpu <= 0) ngpu = ndev;
  if (ngpu > ndev && !devices) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_create: %d devices asked, %d visible", ngpu, ndev);
  rt_multi *m = new rt_multi();
  for (int k = 0; k < ngpu; k++) {
    rt_ctx *c = nullptr;
    const int rc = rt_create(&c, devices ? devices[k] : k);  // (the same device may be listed twice: tests on one GPU)
    if (rc != RT_OK) {
      for (rt_ctx *q : m->ctx) rt_destroy(q);
      delete m;
      return rc;
    }
    m->ctx.push_back(c);
  }
  m->lay.resize(ngpu);
  m->nloc.resize(ngpu);
  *out = m;
  return RT_OK;
}

extern "C" void rt_multi_destroy(rt_multi *m) {
  if (!m) return;
  for (rt_ctx *c : m->ctx) rt_destroy(c);
  delete m;
}

extern "C" int rt_multi_count(const rt_multi *m) { return m ? (int)m->ctx.size() : RT_ERR_INVALID; }
extern "C" rt_ctx *rt_multi_ctx(rt_multi *m, int k) { return (m && k >= 0 && k < (int)m->ctx.size()) ? m->ctx[k] : nullptr; }
extern "C" const char *rt_multi_last_error(const rt_multi *m) { return m ? m->err.c_str() : g_last_error.c_str(); }

extern "C" int rt_multi_set_env(rt_multi *m, const rt_env_params *env) {
  if (!m) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_set_env: null context");
  for (rt_ctx *c : m->ctx)
    if (int rc = rt_set_env(c, env)) { m->err = c->err; return rc; }
  return RT_OK;
}
extern "C" int rt_multi_set_solver(rt_multi *m, const rt_solver_opts *o) {
  if (!m) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_set_solver: null context");
  for (rt_ctx *c : m->ctx)
    if (int rc = rt_set_solver(c, o)) { m->err = c->err; return rc; }
  return RT_OK;
}
extern "C" int rt_multi_set_block(rt_multi *m, int64_t block) {
  if (!m || block < 0) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_set_block: bad argument");
  m->block = block;
  return RT_OK;
}

// the partition the next solve of n rays would use: block length, and per device the local count
static long long multi_block(const rt_multi *m, long long n) {
  if (m->block > 0) return m->block;
  // rows of 32 kB keep the 2D copies at PCIe speed; small enough that every device sees >= 32 blocks spread over the
  // whole launch list (the expensive rays of a sweep cluster in launch-parameter space: on the 8 x 10^6-ray cfg3 sweep
  // 4096-ray blocks leave the devices' work within 0.5 % of each other, 16384-ray blocks within 1 %)
  const long long N = (long long)m->ctx.size();
  long long B = 4096;
  while (B > 128 && n / (B * N) < 32) B /= 2;
  return B;
}
extern "C" int rt_multi_partition(rt_multi *m, int64_t n, int64_t *block, int64_t *counts) {
  if (!m || n < 0) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_partition: bad argument");
  const long long B = multi_block(m, n);
  if (block) *block = B;
  if (counts)
    for (int d = 0; d < (int)m->ctx.size(); d++) counts[d] = rt_local_count(n, B, (int)m->ctx.size(), d);
  return RT_OK;
}
extern "C" int64_t rt_multi_global_index(rt_multi *m, int64_t n, int gpu, int64_t local) {
  if (!m || gpu < 0 || gpu >= (int)m->ctx.size()) return -1;
  RtHostLayout L{multi_block(m, n), (int)m->ctx.size(), gpu};
  return rt_to_global(&L, local);
}

static int solve_multi(rt_multi *m, int64_t n, const double *const *cols, int ncols, rt_result *res, bool want3d) {
  if (!m) return set_err(nullptr, RT_ERR_INVALID, "rt_solve_multi: null context");
  const int N = (int)m->ctx.size();
  if (n < 0) { m->err = "rt_solve_multi: n out of range"; return RT_ERR_INVALID; }
  if (n > 0 && !result_ok(res)) { m->err = "rt_solve_multi: rt_result has null members"; return RT_ERR_INVALID; }
  const long long B = multi_block(m, n);
  std::vector<int> rc(N, RT_OK);
  std::vector<std::thread> th;
  for (int d = 0; d < N; d++) {
    m->lay[d] = RtHostLayout{B, N, d};
    m->nloc[d] = rt_local_count(n, B, N, d);
    th.emplace_back([=, &rc]() {
      rc[d] = solve_common(m->ctx[d], m->nloc[d], cols, ncols, res, want3d, false, &m->lay[d], n);
    });
  }
  for (auto &t : th) t.join();
  for (int d = 0; d < N; d++)
    if (rc[d] != RT_OK) {
      m->err = "device " + std::to_string(d) + ": " + m->ctx[d]->err;
      g_last_error = m->err;
      return rc[d];
    }
  return RT_OK;
}

extern "C" int rt_solve_2d_multi(rt_multi *m, int64_t n, const double *r, const double *theta, const double *chi,
                                 const double *f, const double *t0, rt_result *res) {
  const double *cols[5] = {r, theta, chi, f, t0};
  return solve_multi(m, n, cols, 5, res, false);
}
extern "C" int rt_solve_3d_multi(rt_multi *m, int64_t n, const double *r, const double *theta, const double *phi,
                                 const double *rho_r, const double *rho_theta, const double *rho_phi, const double *T0,
                                 const double *f, const double *t0, rt_result *res) {
  const double *cols[9] = {r, theta, phi, rho_r, rho_theta, rho_phi, T0, f, t0};
  return solve_multi(m, n, cols, 9, res, true);
}
extern "C" int rt_multi_get_stats(const rt_multi *m, int gpu, rt_stats *st) {
  if (!m || gpu < 0 || gpu >= (int)m->ctx.size()) return set_err(nullptr, RT_ERR_INVALID, "rt_multi_get_stats: bad argument");
  return rt_get_stats(m->ctx[gpu], st);
}

// ---------------------------------------------------------------------------------------------------
// Trajectory output (sol.t / sol[i,:] / saved_mu: RayTrace_lat.jl:345-372)
// ---------------------------------------------------------------------------------------------------
extern "C" int rt_traj_columns(const rt_ctx *ctx) {
  if (!ctx) return RT_ERR_INVALID;
  return ctx->env.preset == RT_PRESET_RAY3D ? 8 : 9;
}

static int traj_ready(rt_ctx *ctx, const char *who) {
  if (!ctx) return set_err(ctx, RT_ERR_INVALID, "%s: null context", who);
  if (ctx->last_mode < 0 || ctx->last_max_samples <= 0)
    return set_err(ctx, RT_ERR_STATE, "%s: the last solve did not save trajectories (save_mode = 0)", who);
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  return RT_OK;
}

// CSR offsets of the last solve's trajectories on the device: pack_off[0..n] = exclusive prefix sum of nsamp (cub scan)
static int traj_offsets_device(rt_ctx *ctx) {
  const long long n = ctx->last_n;
  if ((size_t)(n + 1) > ctx->pack_off_cap) {
    cudaFree(ctx->pack_off);
    ctx->pack_off = nullptr;
    ctx->pack_off_cap = 0;
    RT_CUDA(ctx, cudaMalloc(&ctx->pack_off, (size_t)(n + 1) * sizeof(long long)));
    ctx->pack_off_cap = (size_t)(n + 1);
  }
  RT_CUDA(ctx, cudaMemsetAsync(ctx->pack_off, 0, sizeof(long long), ctx->stream));
  if (n == 0) return RT_OK;
  size_t need = 0;
  RT_CUDA(ctx, cub::DeviceScan::InclusiveSum(nullptr, need, ctx->nsamp, ctx->pack_off + 1, (int)n, ctx->stream));
  if (need > ctx->scan_tmp_bytes) {
    cudaFree(ctx->scan_tmp);
    ctx->scan_tmp = nullptr;
    ctx->scan_tmp_bytes = 0;
    RT_CUDA(ctx, cudaMalloc(&ctx->scan_tmp, need));
    ctx->scan_tmp_bytes = need;
  }
  RT_CUDA(ctx, cub::DeviceScan::InclusiveSum(ctx->scan_tmp, need, ctx->nsamp, ctx->pack_off + 1, (int)n, ctx->stream));
  return RT_OK;
}

static int ensure_pack_out(rt_ctx *ctx, size_t bytes) {
  if (bytes > ctx->pack_out_bytes) {
    cudaFree(ctx->pack_out);
    ctx->pack_out = nullptr;
    ctx->pack_out_bytes = 0;
    cudaError_t e = cudaMalloc(&ctx->pack_out, bytes);
    if (e != cudaSuccess) return set_err(ctx, RT_ERR_NOMEM, "trajectory packing: %s (%zu bytes)", cudaGetErrorString(e), bytes);
    ctx->pack_out_bytes = bytes;
  }
  return RT_OK;
}

extern "C" int rt_traj_size(rt_ctx *ctx, int64_t *total) {
  if (!total) return set_err(ctx, RT_ERR_INVALID, "rt_traj_size: null argument");
  if (int rc = traj_ready(ctx, "rt_traj_size")) return rc;
  if (int rc = traj_offsets_device(ctx)) return rc;
  long long t = 0;
  RT_CUDA(ctx, cudaMemcpyAsync(&t, ctx->pack_off + ctx->last_n, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *total = t;
  return RT_OK;
}

// 1 where a trajectory did not fit max_samples (samples were dropped before its final one), flags[n]
extern "C" int rt_traj_truncated(rt_ctx *ctx, int32_t *flags) {
  if (!flags) return set_err(ctx, RT_ERR_INVALID, "rt_traj_truncated: null argument");
  if (int rc = traj_ready(ctx, "rt_traj_truncated")) return rc;
  if (ctx->last_n == 0) return RT_OK;
  RT_CUDA(ctx, cudaMemcpyAsync(flags, ctx->trunc, (size_t)ctx->last_n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

// rows [n][cap][ncol] -> packed [total][ncol] by the CSR offsets: a ray's rows are contiguous on both sides, so each
// block streams one ray's cnt * ncol doubles (coalesced loads and stores; the grid strides over the rays)
__global__ void __launch_bounds__(256) rt_traj_pack_kernel(long long n, int cap, int ncol, const double *__restrict__ traj,
                                                           const long long *__restrict__ offsets, double *__restrict__ out) {
  for (long long ray = blockIdx.x; ray < n; ray += gridDim.x) {
    const long long lo = offsets[ray], len = (offsets[ray + 1] - lo) * ncol;
    const double *src = traj + (size_t)ray * cap * ncol;
    double *dst = out + (size_t)lo * ncol;
    for (long long k = threadIdx.x; k < len; k += blockDim.x) dst[k] = src[k];
  }
}

static unsigned pack_grid(long long n) { return (unsigned)(n < 148LL * 64 ? (n > 0 ? n : 1) : 148LL * 64); }

// packed rows and offsets in DEVICE memory (owned by the context until the next solve / packing call)
extern "C" int rt_traj_pack_device(rt_ctx *ctx, const int64_t **d_offsets, const double **d_samples, int64_t *total) {
  if (int rc = traj_ready(ctx, "rt_traj_pack_device")) return rc;
  if (int rc = traj_offsets_device(ctx)) return rc;
  const long long n = ctx->last_n;
  const int ncol = ctx->last_mode == RT_PRESET_RAY3D ? 8 : 9;
  long long t = 0;
  RT_CUDA(ctx, cudaMemcpyAsync(&t, ctx->pack_off + n, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (int rc = ensure_pack_out(ctx, (size_t)(t > 0 ? t : 1) * ncol * sizeof(double))) return rc;
  if (t > 0) {
    rt_traj_pack_kernel<<<pack_grid(n), 256, 0, ctx->stream>>>(n, ctx->last_max_samples, ncol, ctx->traj, ctx->pack_off, ctx->pack_out);
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
  if (d_offsets) *d_offsets = (const int64_t *)ctx->pack_off;
  if (d_samples) *d_samples = ctx->pack_out;
  if (total) *total = t;
  return RT_OK;
}

extern "C" int rt_traj_copy(rt_ctx *ctx, int64_t *offsets, double *samples) {
  if (!offsets || !samples) return set_err(ctx, RT_ERR_INVALID, "rt_traj_copy: null argument");
  int64_t total = 0;
  if (int rc = rt_traj_pack_device(ctx, nullptr, nullptr, &total)) return rc;
  const long long n = ctx->last_n;
  const int ncol = ctx->last_mode == RT_PRESET_RAY3D ? 8 : 9;
  RT_CUDA(ctx, cudaMemcpyAsync(offsets, ctx->pack_off, (size_t)(n + 1) * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
  if (total > 0)
    RT_CUDA(ctx, cudaMemcpyAsync(samples, ctx->pack_out, (size_t)total * ncol * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

// Post-processing of the saved trajectories on the device: the Cartesian track and the step-size
// series the reference derives after a solve — x = r cos(lambda), y = r sin(lambda) (RayTrace_lat.jl:351-352;
// colatitude presets: x = r sin(theta), y = r cos(theta), RayMain.jl:400-401, scratch.jl:407-408),
// dt[i] = t[i+1] - t[i] with a trailing 0 (RayTrace_lat.jl:369-372).  3D [EXT]: x = r sin(theta) cos(phi),
// y = r sin(theta) sin(phi), z = r cos(theta).  Packed like rt_traj_copy: out[offsets[ray] + k][3 or 4].
__global__ void __launch_bounds__(128) rt_traj_derived_kernel(long long n, int cap, int ncol, int mode, const double *traj,
                                                              const long long *offsets, double *out) {
  const int nd = (mode == RT_PRESET_RAY3D) ? 4 : 3;
  for (long long ray = blockIdx.x; ray < n; ray += gridDim.x) {
    const long long lo = offsets[ray], cnt = offsets[ray + 1] - lo;
    const double *src = traj + (size_t)ray * cap * ncol;
    for (long long k = threadIdx.x; k < cnt; k += blockDim.x) {
      const double *row = src + (size_t)k * ncol;
      const double t = row[0], r = row[1], a = row[2];
      double *o = out + (size_t)(lo + k) * nd;
      double sa, ca;
      rt_sincos(a, &sa, &ca);
      if (mode == RT_PRESET_RAY3D) {
        double sp, cp;
        rt_sincos(row[3], &sp, &cp);
        o[0] = (r * sa) * cp;
        o[1] = (r * sa) * sp;
        o[2] = r * ca;
      } else if (mode == RT_PRESET_LAT2D) {
        o[0] = r * ca;
        o[1] = r * sa;
      } else {
        o[0] = r * sa;
        o[1] = r * ca;
      }
      o[nd - 1] = (k + 1 < cnt) ? src[(size_t)(k + 1) * ncol] - t : 0.0;
    }
  }
}

extern "C" int rt_traj_derived_columns(rt_ctx *ctx) {
  if (!ctx) return set_err(ctx, RT_ERR_INVALID, "rt_traj_derived_columns: null context");
  return ctx->env.preset == RT_PRESET_RAY3D ? 4 : 3;
}

// `offsets` (as returned by rt_traj_copy) is accepted for the caller's convenience; the device-side offsets of the last
// solve are what the kernel uses
extern "C" int rt_traj_derived(rt_ctx *ctx, const int64_t *offsets, double *out) {
  if (!offsets || !out) return set_err(ctx, RT_ERR_INVALID, "rt_traj_derived: null argument");
  if (int rc = traj_ready(ctx, "rt_traj_derived")) return rc;
  if (int rc = traj_offsets_device(ctx)) return rc;
  const long long n = ctx->last_n;
  const int ncol = ctx->last_mode == RT_PRESET_RAY3D ? 8 : 9, nd = ctx->last_mode == RT_PRESET_RAY3D ? 4 : 3;
  const int64_t total = offsets[n];
  if (total == 0) return RT_OK;
  if (int rc = ensure_pack_out(ctx, (size_t)total * nd * sizeof(double))) return rc;
  rt_traj_derived_kernel<<<pack_grid(n), 128, 0, ctx->stream>>>(n, ctx->last_max_samples, ncol, ctx->last_mode, ctx->traj,
                                                                ctx->pack_off, ctx->pack_out);
  RT_CUDA(ctx, cudaGetLastError());
  RT_CUDA(ctx, cudaMemcpyAsync(out, ctx->pack_out, (size_t)total * nd * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

#ifdef RT_DEV
// census of exact-path fallbacks since the last reset (diagnostics)
extern "C" int rt_debug_fallback_counts(rt_ctx *ctx, unsigned long long *out8, int reset) {
  if (!ctx || !out8) return set_err(ctx, RT_ERR_INVALID, "rt_debug_fallback_counts: null argument");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  RT_CUDA(ctx, cudaMemcpyFromSymbol(out8, rt_dbg_fallback, 8 * sizeof(unsigned long long)));
  if (reset) {
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    RT_CUDA(ctx, cudaMemcpyToSymbol(rt_dbg_fallback, z, sizeof(z)));
  }
  return RT_OK;
}

// Both versions of the 3D right-hand side on caller-chosen inputs (tests: wherever the branch-free one says its
// result is exact, it has to have the exact one's bits, or NaN where that is NaN).  in[n][8] = u[0..6], freq.
__global__ void rt_rhs3d_probe_kernel(const __grid_constant__ RtDevEnv E, int n, const double *in, double *du_fast,
                                      int *ok, double *du_exact) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double u[7], df[7], de[7];
  for (int c = 0; c < 7; c++) u[c] = in[(size_t)i * 8 + c];
  const double f = in[(size_t)i * 8 + 7];
  for (int c = 0; c < 7; c++) df[c] = 0.0;
  ok[i] = rtf_haselgrove_3d<false>(E, u, f, df) ? 1 : 0;
  rt_haselgrove_3d(E, u, f, de);
  for (int c = 0; c < 7; c++) {
    du_fast[(size_t)i * 7 + c] = df[c];
    du_exact[(size_t)i * 7 + c] = de[c];
  }
}

extern "C" int rt_debug_rhs3d(rt_ctx *ctx, int n, const double *in, double *du_fast, int *ok, double *du_exact) {
  if (!ctx || n < 0 || !in || !du_fast || !ok || !du_exact) return set_err(ctx, RT_ERR_INVALID, "rt_debug_rhs3d: bad argument");
  if (n == 0) return RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  double *d_in = nullptr, *d_f = nullptr, *d_e = nullptr;
  int *d_ok = nullptr;
  cudaError_t e = cudaMalloc(&d_in, (size_t)n * 8 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&d_f, (size_t)n * 7 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&d_e, (size_t)n * 7 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&d_ok, (size_t)n * sizeof(int));
  if (e == cudaSuccess) e = cudaMemcpy(d_in, in, (size_t)n * 8 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    rt_rhs3d_probe_kernel<<<(n + 63) / 64, 64>>>(ctx->denv, n, d_in, d_f, d_ok, d_e);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(du_fast, d_f, (size_t)n * 7 * sizeof(double), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(du_exact, d_e, (size_t)n * 7 * sizeof(double), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(ok, d_ok, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost);
  cudaFree(d_in);
  cudaFree(d_f);
  cudaFree(d_e);
  cudaFree(d_ok);
  if (e != cudaSuccess) return set_err(ctx, RT_ERR_CUDA, "rt_debug_rhs3d: %s", cudaGetErrorString(e));
  return RT_OK;
}

// the inputs of the first exact-path calls since the last counter reset: out[max_rows][8] = u[0..6], freq (3D)
extern "C" int rt_debug_fallback_samples(rt_ctx *ctx, double *out, int max_rows) {
  if (!ctx || !out || max_rows < 0) return set_err(ctx, RT_ERR_INVALID, "rt_debug_fallback_samples: bad argument");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  const int rows = max_rows < RT_DBG_SAMPLES ? max_rows : RT_DBG_SAMPLES;
  RT_CUDA(ctx, cudaMemcpyFromSymbol(out, rt_dbg_sample, (size_t)rows * 8 * sizeof(double)));
  return rows;
}

#endif  // RT_DEV

// ---------------------------------------------------------------------------------------------------
// FP64 roofline denominator: NCH independent DFMA chains per thread, every SM full.  Several shapes are timed and the
// best rate is returned (chains x resident warps: the pipe issues one FP64 warp instruction per 2 cycles per
// scheduler and a dependent DFMA returns after 8, so >= 4 chains in flight per scheduler saturate it; the shapes
// differ only in how much slack they leave).  profiles/r2_dfma_peak_ncu.md has the pipe-utilisation proof.
// ---------------------------------------------------------------------------------------------------
template <int NCH>
__global__ void __launch_bounds__(256) rt_dfma_peak_kernel(double *out, int iters, double a, double b) {
  double x[NCH];
#pragma unroll
  for (int c = 0; c < NCH; c++) x[c] = (double)threadIdx.x + c;
#pragma unroll 4
  for (int k = 0; k < iters; k++) {
#pragma unroll
    for (int c = 0; c < NCH; c++) x[c] = __fma_rn(x[c], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < NCH; c++) s += x[c];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" int rt_measure_fp64_peak(rt_ctx *ctx, double *flops) {
  if (!ctx || !flops) return set_err(ctx, RT_ERR_INVALID, "rt_measure_fp64_peak: null argument");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaDeviceProp prop;
  RT_CUDA(ctx, cudaGetDeviceProperties(&prop, ctx->device));
  const int max_blocks = prop.multiProcessorCount * 8;
  double *d = nullptr;
  RT_CUDA(ctx, cudaMalloc(&d, (size_t)max_blocks * 256 * sizeof(double)));
  double best = 0.0;
  // shape = (chains per thread, blocks of 256 threads per SM); each launch is ~2 ms of DFMAs
  const int shapes[4][2] = {{8, 8}, {8, 4}, {16, 4}, {4, 8}};
  for (int sh = 0; sh < 4; sh++) {
    const int nch = shapes[sh][0], blocks = prop.multiProcessorCount * shapes[sh][1];
    const int iters = (1 << 20) / (nch * shapes[sh][1]);
    for (int rep = 0; rep < 4; rep++) {
      cudaEventRecord(ctx->ev[0], ctx->stream);
      if (nch == 4) rt_dfma_peak_kernel<4><<<blocks, 256, 0, ctx->stream>>>(d, iters, 0.9999999, 1e-7);
      else if (nch == 8) rt_dfma_peak_kernel<8><<<blocks, 256, 0, ctx->stream>>>(d, iters, 0.9999999, 1e-7);
      else rt_dfma_peak_kernel<16><<<blocks, 256, 0, ctx->stream>>>(d, iters, 0.9999999, 1e-7);
      cudaEventRecord(ctx->ev[1], ctx->stream);
      cudaError_t e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) {
        cudaFree(d);
        return set_err(ctx, RT_ERR_CUDA, "rt_measure_fp64_peak: %s", cudaGetErrorString(e));
      }
      float ms = 0;
      cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
      const double fl = 2.0 * nch * (double)iters * (double)blocks * 256.0 / (ms * 1e-3);
      if (rep > 0 && fl > best) best = fl;
    }
  }
  cudaFree(d);
  *flops = best;
  return RT_OK;
}
