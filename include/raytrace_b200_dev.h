/* raytrace_b200_dev.h — diagnostics exported ONLY by the developer build of the library (-DRT_DEV:
 * ray-trace_b200/libraytrace_b200_dev.so).  The product library (libraytrace_b200.so) exports the reference-facing
 * surface of raytrace_b200.h and nothing else, reads no environment variable and keeps no census counters.
 * The developer build additionally reads the scheduling switches RT_DEBUG, RT_COOP_POLICY, RT_COOP_SUM, RT_COOP_ATT,
 * RT_WAIT_ATT, RT_SPLIT_MIN, RT_EARLY_MIN, RT_EARLY_DIV from the environment (results never depend on them). */
#ifndef RAYTRACE_B200_DEV_H
#define RAYTRACE_B200_DEV_H

#include "raytrace_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- self-test hooks ------------------------------------------------------------------------------------ */
/* Bit-equality of the kernels' branch-free division / sqrt / elementary functions against the exact ones on
 * pseudo-random operands.  counts[10][3] = {compared, mismatches, flagged-to-exact-path} for
 * div, div(shared rcp), div(constant divisor), sqrt, exp, 10^x, sincos, acos, atan, x^-4.5. */
RT_API int rt_selftest_fastmath(int device, long long per_thread, unsigned long long seed,
                                unsigned long long *counts, int ncounts);
/* how often the kernels left the branch-free right-hand side for the exact one since the last reset:
 * out8 = {calls, r out of range, non-finite input, non-finite result, finite both (conservative guard), 2D calls,
 *         r is NaN, |r| < 2^16} */
RT_API int rt_debug_fallback_counts(rt_ctx *ctx, unsigned long long *out8, int reset);
/* both versions of the 3D right-hand side on n caller-chosen inputs in[n][8] = u[0..6], freq (3D preset's
 * environment): du_fast[n][7] and ok[n] from the branch-free code, du_exact[n][7] from the exact code */
RT_API int rt_debug_rhs3d(rt_ctx *ctx, int n, const double *in, double *du_fast, int *ok, double *du_exact);
/* inputs of the first exact-path calls since the last reset: out[max_rows][8] = u[0..6], freq (3D preset);
 * returns the number of rows copied (at most 8192; rows beyond the call count are stale) */
RT_API int rt_debug_fallback_samples(rt_ctx *ctx, double *out, int max_rows);
/* cycles per dependent operation with `warps` warps resident on one SM:
 * res[6] = DFMA, DMUL, DADD, MUFU.RCP64H chains, then 2 and 4 independent DFMA chains per thread */
RT_API int rt_probe_latency(int device, int warps, double *res);
/* FP64 pipe throughput of synthetic instruction streams with `warps_per_sched` warps per scheduler, as a fraction of the
 * nominal rate (one FP64 warp instruction per 2 cycles per scheduler) at clock_mhz: res[5] = {DFMA with reused operands,
 * DFMA with three distinct register operands, DFMA/DMUL/DADD mix, + one integer instruction per FP64 instruction,
 * + one FP32 compare/select per FP64 instruction}; run it under ncu for sm__pipe_fp64_cycles_active */
RT_API int rt_probe_pipe(int device, int warps_per_sched, double clock_mhz, double *res);

#ifdef __cplusplus
}
#endif
#endif /* RAYTRACE_B200_DEV_H */
