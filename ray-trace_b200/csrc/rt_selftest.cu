// rt_selftest.cu — on-device bit-equality check of the branch-free primitives (rt_fast.cuh) against the exact
// ones (IEEE `/`, sqrt, rt_math.h), plus dependent-chain latency probes.  Test/diagnostic entry points only.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/raytrace_b200_dev.h"
#include "rt_fast.cuh"

__device__ __forceinline__ uint64_t sm64(uint64_t &s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ULL);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
// uniform in [0,1)
__device__ __forceinline__ double u01(uint64_t &s) { return (double)(sm64(s) >> 11) * (1.0 / 9007199254740992.0); }
// random sign, random mantissa, exponent uniform in [elo, ehi]
__device__ __forceinline__ double rnd_exp(uint64_t &s, int elo, int ehi) {
  uint64_t m = sm64(s);
  int e = elo + (int)(sm64(s) % (uint64_t)(ehi - elo + 1));
  uint64_t bits = (m & 0x800fffffffffffffULL) | ((uint64_t)(e + 1023) << 52);
  return rt_u2d(bits);
}
__device__ __forceinline__ double special(uint64_t &s) {
  const double tab[12] = {0.0, -0.0, 1.0, -1.0, 0.5, 1.5, rt_inf(), -rt_inf(), rt_nan(), 4.9e-324, 1.7976931348623157e308, 2.2250738585072014e-308};
  return tab[sm64(s) % 12];
}
__device__ __forceinline__ bool same(double a, double b) { return rt_d2u(a) == rt_d2u(b) || (a != a && b != b); }

// counters: [f][0] = compared (ok), [f][1] = mismatches among them, [f][2] = flagged (not ok)
#define NF 10
__global__ void rt_selftest_kernel(long long per_thread, uint64_t seed, unsigned long long *cnt) {
  uint64_t s = seed ^ ((uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0xD1342543DE82EF95ULL);
  unsigned long long c[NF][3];
  for (int f = 0; f < NF; f++) c[f][0] = c[f][1] = c[f][2] = 0;
#define TALLY(f, okv, fast, exact) do { if (okv) { c[f][0]++; if (!same(fast, exact)) c[f][1]++; } else c[f][2]++; } while (0)
  for (long long it = 0; it < per_thread; it++) {
    const bool sp = (it % 64) == 63;
    {  // 0: division, wide exponents
      double a = sp ? special(s) : rnd_exp(s, -60, 60), b = sp ? special(s) : rnd_exp(s, -60, 60);
      if ((it % 16) == 5) { a = rnd_exp(s, -1070, 1023); b = rnd_exp(s, -1070, 1023); }
      bool ok = true;
      double q = rtf_div(a, b, ok);
      TALLY(0, ok, q, a / b);
      // 1: shared reciprocal, constant-divisor flavour
      double a2 = rnd_exp(s, -30, 30);
      bool ok2 = true;
      double y = rtf_rcp(b);
      double q2 = rtf_div_y(a2, b, y, ok2);
      TALLY(1, ok2, q2, a2 / b);
      bool ok3 = rtf_midrange(b);
      double q3 = rtf_div_c(a2, b, y, ok3);
      TALLY(2, ok3, q3, a2 / b);
    }
    {  // 3: sqrt
      double x = sp ? special(s) : rt_abs(rnd_exp(s, -80, 80));
      if ((it % 16) == 7) x = rnd_exp(s, -1070, 1023);
      bool ok = true;
      double r = rtf_sqrt(x, ok);
      TALLY(3, ok, r, __dsqrt_rn(x));
    }
    {  // 4: exp, 5: exp10
      double x = sp ? special(s) : (u01(s) - 0.5) * 1500.0;
      if ((it % 4) == 1) x = (u01(s) - 0.5) * 20.0;
      bool ok = true;
      double r = rtf_exp(x, ok);
      TALLY(4, ok, r, rt_exp(x));
      double x10 = x * 0.43;
      bool ok2 = true;
      double r2 = rtf_exp10(x10, ok2);
      TALLY(5, ok2, r2, rt_exp10(x10));
    }
    {  // 6: sincos
      double x = sp ? special(s) : (u01(s) - 0.5) * 20.0;
      if ((it % 8) == 3) x = rnd_exp(s, -30, 50);
      bool ok = true;
      double sn, cs, se, ce;
      rtf_sincos(x, &sn, &cs, ok);
      rt_sincos(x, &se, &ce);
      if (ok) { c[6][0]++; if (!same(sn, se) || !same(cs, ce)) c[6][1]++; } else c[6][2]++;
    }
    {  // 7: acos
      double x = sp ? special(s) : (u01(s) - 0.5) * 2.0;
      if ((it % 8) == 2) x = (u01(s) < 0.5 ? -1.0 : 1.0) * (1.0 - rt_abs(rnd_exp(s, -50, -1)));
      if ((it % 8) == 6) x = rnd_exp(s, -400, 1);
      bool ok = true;
      double r = rtf_acos(x, ok);
      TALLY(7, ok, r, rt_acos(x));
    }
    {  // 8: atan
      double x = sp ? special(s) : rnd_exp(s, -20, 20);
      if ((it % 4) == 0) x = (u01(s) - 0.5) * 6.0;
      if ((it % 32) == 9) x = (sm64(s) & 1) ? 0.5 : ((sm64(s) & 1) ? 1.0 : -1.5);
      bool ok = true;
      double r = rtf_atan(x, ok);
      TALLY(8, ok, r, rt_atan(x));
    }
    {  // 9: x^-4.5
      double x = sp ? special(s) : 1.0 + u01(s) * 60.0;
      bool ok = true;
      double r = rtf_pow_m4p5(x, ok);
      TALLY(9, ok, r, rt_pow_m4p5(x));
    }
  }
  for (int f = 0; f < NF; f++)
    for (int k = 0; k < 3; k++)
      if (c[f][k]) atomicAdd(&cnt[f * 3 + k], c[f][k]);
}

// counts[NF*3]; returns 0 on success
extern "C" RT_API int rt_selftest_fastmath(int device, long long per_thread, unsigned long long seed,
                                           unsigned long long *counts, int ncounts) {
  if (!counts || ncounts < NF * 3) return RT_ERR_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return RT_ERR_NO_DEVICE;
  unsigned long long *d = nullptr;
  if (cudaMalloc(&d, NF * 3 * sizeof(unsigned long long)) != cudaSuccess) return RT_ERR_NOMEM;
  cudaMemset(d, 0, NF * 3 * sizeof(unsigned long long));
  rt_selftest_kernel<<<148 * 4, 128>>>(per_thread, seed, d);
  cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(counts, d, NF * 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  cudaFree(d);
  return e == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}

// ---- latency probes: one warp per SM-quadrant, dependent chains -----------------------------------
__global__ void rt_lat_kernel(int op, int iters, double a, double b, double *out, long long *cyc) {
  double x = a + threadIdx.x * 1e-9, y = b;
  long long t0 = clock64();
  if (op == 0) for (int i = 0; i < iters; i++) { x = __fma_rn(x, y, y); x = __fma_rn(x, y, y); x = __fma_rn(x, y, y); x = __fma_rn(x, y, y); }
  if (op == 1) for (int i = 0; i < iters; i++) { x = __dmul_rn(x, y); x = __dmul_rn(x, y); x = __dmul_rn(x, y); x = __dmul_rn(x, y); }
  if (op == 2) for (int i = 0; i < iters; i++) { x = __dadd_rn(x, y); x = __dadd_rn(x, y); x = __dadd_rn(x, y); x = __dadd_rn(x, y); }
  if (op == 3) for (int i = 0; i < iters; i++) {
      asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x)); asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x));
      asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x)); asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x)); }
  if (op == 4) {  // two independent DFMA chains
    double z = x + 1.0;
    for (int i = 0; i < iters; i++) { x = __fma_rn(x, y, y); z = __fma_rn(z, y, y); x = __fma_rn(x, y, y); z = __fma_rn(z, y, y); }
    x += z;
  }
  if (op == 5) {  // four independent DFMA chains
    double z = x + 1.0, w = x + 2.0, v = x + 3.0;
    for (int i = 0; i < iters; i++) { x = __fma_rn(x, y, y); z = __fma_rn(z, y, y); w = __fma_rn(w, y, y); v = __fma_rn(v, y, y); }
    x += z + w + v;
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// cycles per dependent op for op in 0..5 with `warps` warps in one block (1 block): res[6]
extern "C" RT_API int rt_probe_latency(int device, int warps, double *res) {
  if (!res || warps < 1 || warps > 32) return RT_ERR_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return RT_ERR_NO_DEVICE;
  double *out; long long *cyc;
  cudaMalloc(&out, 32 * 32 * sizeof(double));
  cudaMalloc(&cyc, sizeof(long long));
  const int iters = 4096;
  for (int op = 0; op < 6; op++) {
    long long h = 0;
    for (int rep = 0; rep < 2; rep++) {
      rt_lat_kernel<<<1, 32 * warps>>>(op, iters, 0.999, 1e-3, out, cyc);
      cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    }
    res[op] = (double)h / (4.0 * iters);
  }
  cudaError_t e = cudaDeviceSynchronize();
  cudaFree(out); cudaFree(cyc);
  return e == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}

// ---- FP64 pipe throughput probes: what a real instruction stream can reach --------------------------------------------
// variant 0: x_c = fma(x_c, a, b)             (the peak microbenchmark: two operands from the reuse cache)
// variant 1: x_c = fma(x_c+1, x_c+2, x_c+3)   (three distinct register pairs per DFMA, as in real code)
// variant 2: variant 1, alternating DFMA / DMUL / DADD
// variant 3: variant 1 with one integer instruction (LOP3 on a side value) after every DFMA (40 % non-FP64, as the kernel)
// variant 4: variant 1 with one FP32 compare + select after every DFMA (the guard logic of the branch-free divisions)
// `warps_per_sched` x 4 warps per SM (blocks of 128 threads), NCH chains per thread
template <int VAR, int NCH>
__global__ void __launch_bounds__(128) rt_pipe_kernel(double *out, int iters, double a, double b) {
  double x[NCH], kk[NCH];  // kk: small per-chain factors held in registers, so that the values stay bounded
#pragma unroll
  for (int c = 0; c < NCH; c++) {
    x[c] = 1.0 + 1e-3 * (threadIdx.x + c);
    kk[c] = a * 1e-12 * (c + 1);
  }
  unsigned side = threadIdx.x;
  float fs = (float)threadIdx.x;
#pragma unroll 2
  for (int k = 0; k < iters; k++) {
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      if (VAR == 0) x[c] = __fma_rn(x[c], a, b);
      else if (VAR == 2 && (c % 3) == 1) x[c] = __dmul_rn(x[(c + 1) % NCH], kk[(c + 2) % NCH]) + 0.0 * b;
      else if (VAR == 2 && (c % 3) == 2) x[c] = __dadd_rn(x[(c + 3) % NCH], kk[(c + 1) % NCH]);
      else x[c] = __fma_rn(x[(c + 1) % NCH], kk[(c + 2) % NCH], x[(c + 3) % NCH]);
      if (VAR == 3) side = (side ^ (side << 1)) & 0x7fffffffu;
      if (VAR == 4) fs = (fabsf(fs) < 6.5e-37f) ? fs + 1.0f : fs;
    }
  }
  double s = (double)side + (double)fs;
#pragma unroll
  for (int c = 0; c < NCH; c++) s += x[c];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// res[5] = FP64 warp instructions per cycle per scheduler x 2.0 for the five variants (1.0 = nominal pipe rate);
// wall-clock based (CUDA events), SM clock taken as clock_mhz
extern "C" RT_API int rt_probe_pipe(int device, int warps_per_sched, double clock_mhz, double *res) {
  if (!res || warps_per_sched < 1 || warps_per_sched > 12) return RT_ERR_INVALID;
  if (cudaSetDevice(device) != cudaSuccess) return RT_ERR_NO_DEVICE;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  const int blocks = prop.multiProcessorCount * warps_per_sched, iters = 1 << 13;
  constexpr int NCH = 8;
  double *out;
  cudaMalloc(&out, (size_t)blocks * 128 * sizeof(double));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int v = 0; v < 5; v++) {
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
      cudaEventRecord(e0);
      if (v == 0) rt_pipe_kernel<0, NCH><<<blocks, 128>>>(out, iters, 0.9999999, 1e-7);
      if (v == 1) rt_pipe_kernel<1, NCH><<<blocks, 128>>>(out, iters, 0.9999999, 1e-7);
      if (v == 2) rt_pipe_kernel<2, NCH><<<blocks, 128>>>(out, iters, 0.9999999, 1e-7);
      if (v == 3) rt_pipe_kernel<3, NCH><<<blocks, 128>>>(out, iters, 0.9999999, 1e-7);
      if (v == 4) rt_pipe_kernel<4, NCH><<<blocks, 128>>>(out, iters, 0.9999999, 1e-7);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
    }
    const double fp64_inst = (double)warps_per_sched * iters * NCH;  // chain ops only
    const double cycles = best * 1e-3 * clock_mhz * 1e6;
    res[v] = 2.0 * fp64_inst / cycles;
  }
  cudaError_t e = cudaDeviceSynchronize();
  cudaFree(out);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return e == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}
