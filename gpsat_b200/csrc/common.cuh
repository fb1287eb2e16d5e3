// Shared device helpers for the GPSat-B200 local-expert GP engine (sm_100a only).
//
// Layout conventions (see DESIGN.md "Data layout in HBM"):
//   * every expert in flight owns one "slot": a row-major FP64 matrix A[rows][ld], NB = 128,
//     ld = T*NB.  After a factor+invert sweep the strict lower block triangle holds the Cholesky
//     factor L of Ks = K + s2 I (augmented with the observation row y, so row n holds
//     a = L^-1 y), the strict upper block triangle holds U = L^-T, and the diagonal tiles hold
//     U_tt.  W_tt = L_tt^-1 (row-major) lives in a side buffer.
//   * the covariance is never materialised: tiles of K are evaluated from the coordinates at
//     the moment a tile is first touched.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define GPSAT_NB 128          // tile edge
#define GPSAT_DMAX 6          // max input dimensions
#define GPSAT_THREADS 256     // threads per CTA for the tile kernels (8 warps, 2 x 4)

enum GpsatKernelKind : int { GPSAT_MATERN12 = 0, GPSAT_MATERN32 = 1, GPSAT_MATERN52 = 2, GPSAT_RBF = 3,
                             GPSAT_PRECOMPUTED = 4 };   // matrix entries read from SlotDesc::pre (SGPR B matrix)

// One expert in flight.  Written by the scheduler / optimiser kernel, read by the tile kernels.
struct SlotDesc {
  int expert;            // index into the batch, -1 = slot idle
  int n;                 // observations
  int m;                 // prediction rows carried in this pass (0 while fitting)
  int T;                 // tile columns to factor: ceil((n + 1) / NB)
  int TR;                // tile rows: ceil((n + 1 + m) / NB)
  int ld;                // leading dimension of A (doubles)
  int kind;              // GpsatKernelKind
  int D;
  const double* X;       // [n][D] observation coordinates (already divided by coords_scale)
  const double* y;       // [n]
  const double* Xs;      // [m][D] prediction coordinates
  double* A;             // slot matrix
  double* Wd;            // [T][NB][NB] inverse diagonal blocks W_tt (row-major)
  double* Ld;            // [T][NB][NB] diagonal blocks L_tt (row-major; tests / debugging)
  double* logdet;        // [T] sum log diag(L_tt)
  double* quad;          // [1] |L^-1 y|^2
  double* gpart;         // [T*(T+1)/2][GPSAT_DMAX + 2] gradient partial sums per lauum tile
  int* notpd;            // [1] set to 1 when a pivot is not positive
  const double* pre;     // kind == GPSAT_PRECOMPUTED: dense symmetric matrix (lower part read), else unused
  const double* pre_s;   // kind == GPSAT_PRECOMPUTED with m > 0: prediction rows [m][ldpre] (Kus^T)
  int ldpre;
  int8_t* Uq;            // int8 slice planes of the upper block triangle (i8mma.cuh), nullptr = DMMA lauum
  int* uex;              // [T*NB] row scale exponents of the planes
  int8_t* Lq;            // int8 slice planes of the strict lower block triangle (step_kernel.cuh), nullptr = DMMA potrf
  int8_t* Uq8;           // same for U = L^-T with a-priori row scales (trtri main loop), nullptr = DMMA trtri
  int8_t* Lq_mem;        // the slot's plane storage; Lq / Uq8 point at it for evaluations the precision rule admits
  int8_t* Uq8_mem;       //   (slot_set_theta in engine.cu) and are nullptr otherwise: FP64 mma.sync for that evaluation
  double ynorm2;         // |y|^2 (a-priori scale of the observation row of the factor)
  int skip_fp64;         // 1: every consumer of this slot's off-diagonal L / U tiles reads the int8 planes (sweep and K^-1
                         //    products on the planes, no debug extraction), so step_kernel does not store the FP64 tiles of
                         //    block rows / columns that lie entirely among the observations
  int i8_low;            // 1: low-rho tier of the precision rule - the int8 k-sums of step_kernel drop their last
                         //    slice-pair group (slot_set_theta in engine.cu)
  double inv_ls[GPSAT_DMAX];
  double var;            // kernel variance
  double noise;          // likelihood variance
};

// ---- kernel functions -----------------------------------------------------------------------
// Reference: gpflow.kernels.{Matern12,Matern32,Matern52,SquaredExponential} as constructed at
// GPSat/models/gpflow_models.py:116-135.  r2 is computed from direct differences of the scaled
// coordinates (no |a|^2+|b|^2-2ab cancellation); Matern kernels clamp r2 at 1e-36 like GPflow.
// kunit = k / variance;  gunit = g / variance with dk/dl_d = g * (scaled diff_d)^2 / l_d.
// FP64 exp / sqrt for the covariance and gradient epilogues.  The CUDA library versions cost about 110 FP64-pipe
// instructions per covariance element together (measured: 15 us per 128 x 128 tile); the arguments here are
// confined (exp: x <= 0, sqrt: a >= 1e-36, both finite), which allows straight-line code of about 30:
//   exp_nonpos: k = rint(x / ln 2) by the add-magic-constant trick, Cody-Waite reduction with a two-part ln 2,
//               degree-13 Taylor polynomial on |r| <= 0.347 (remainder 4e-18), scaling by an exponent-field add.
//               x is clamped at -700 (result 1e-304 instead of something smaller).
//   sqrt_rsqrt_pos: MUFU.RSQ64H seed (rsqrt.approx.ftz.f64, 2^-22) + two coupled Newton steps for g ~ sqrt(a) and
//               h ~ 1 / (2 sqrt(a)) + one final residual correction of g.
// Both are accurate to about 1 ulp, the same level as the differences between libm implementations the oracle
// comparison (rel 1e-8 on LML / predictions) already absorbs.
// -DGPSAT_LIBM_KERNELS builds the library with the CUDA library functions instead (A/B checks).
__device__ __forceinline__ double exp_nonpos(double x) {
#ifdef GPSAT_LIBM_KERNELS
  return exp(x);
#endif
  x = fmax(x, -700.0);
  const double MAGIC = 6755399441055744.0;                    // 1.5 * 2^52
  const double t = fma(x, 1.4426950408889634, MAGIC);
  const int ki = __double2loint(t);
  const double kd = t - MAGIC;
  double r = fma(kd, -6.93147180369123816490e-01, x);
  r = fma(kd, -1.90821492927058770002e-10, r);
  double p = 1.6059043836821613e-10;                          // 1 / 13!
  p = fma(p, r, 2.08767569878681e-09);                        // 1 / 12!
  p = fma(p, r, 2.505210838544172e-08);
  p = fma(p, r, 2.755731922398589e-07);
  p = fma(p, r, 2.7557319223985893e-06);
  p = fma(p, r, 2.48015873015873e-05);
  p = fma(p, r, 1.984126984126984e-04);
  p = fma(p, r, 1.3888888888888889e-03);
  p = fma(p, r, 8.333333333333333e-03);
  p = fma(p, r, 4.1666666666666664e-02);
  p = fma(p, r, 1.6666666666666666e-01);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  return __hiloint2double(__double2hiint(p) + (ki << 20), __double2loint(p));
}

// g = sqrt(a), rinv = 1 / sqrt(a) for finite a >= 1e-36
__device__ __forceinline__ double sqrt_rsqrt_pos(double a, double& rinv) {
#ifdef GPSAT_LIBM_KERNELS
  { const double g0 = sqrt(a); rinv = 1.0 / g0; return g0; }
#endif
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  double h = 0.5 * y;
  double g = a * y;
  double e = fma(-h, g, 0.5);
  g = fma(g, e, g);
  h = fma(h, e, h);
  e = fma(-h, g, 0.5);
  g = fma(g, e, g);
  h = fma(h, e, h);
  g = fma(fma(-g, g, a), h, g);
  rinv = h + h;
  return g;
}

template <int KIND>
__device__ __forceinline__ double kunit_of_r2(double r2) {
  if (KIND == GPSAT_RBF) return exp_nonpos(-0.5 * r2);
  double rinv;
  double r = sqrt_rsqrt_pos(fmax(r2, 1e-36), rinv);
  if (KIND == GPSAT_MATERN12) return exp_nonpos(-r);
  if (KIND == GPSAT_MATERN32) { double s = 1.7320508075688772 * r; return (1.0 + s) * exp_nonpos(-s); }
  double s = 2.23606797749979 * r;
  return (1.0 + s + (5.0 / 3.0) * r * r) * exp_nonpos(-s);
}

template <int KIND>
__device__ __forceinline__ void kunit_gunit_of_r2(double r2, double& ku, double& gu) {
  if (KIND == GPSAT_RBF) { ku = exp_nonpos(-0.5 * r2); gu = ku; return; }
  double rinv;
  double r = sqrt_rsqrt_pos(fmax(r2, 1e-36), rinv);
  if (KIND == GPSAT_MATERN12) {
    ku = exp_nonpos(-r);
    gu = (r2 > 1e-36) ? ku * rinv : 0.0;
    return;
  }
  if (KIND == GPSAT_MATERN32) {
    double s = 1.7320508075688772 * r;
    double e = exp_nonpos(-s);
    ku = (1.0 + s) * e;
    gu = 3.0 * e;
    return;
  }
  double s = 2.23606797749979 * r;
  double e = exp_nonpos(-s);
  ku = (1.0 + s + (5.0 / 3.0) * r * r) * e;
  gu = (5.0 / 3.0) * (1.0 + s) * e;
}

__device__ __forceinline__ double kunit_dyn(int kind, double r2) {
  switch (kind) {
    case GPSAT_MATERN12: return kunit_of_r2<GPSAT_MATERN12>(r2);
    case GPSAT_MATERN32: return kunit_of_r2<GPSAT_MATERN32>(r2);
    case GPSAT_MATERN52: return kunit_of_r2<GPSAT_MATERN52>(r2);
    default: return kunit_of_r2<GPSAT_RBF>(r2);
  }
}

__device__ __forceinline__ void kunit_gunit_dyn(int kind, double r2, double& ku, double& gu) {
  switch (kind) {
    case GPSAT_MATERN12: kunit_gunit_of_r2<GPSAT_MATERN12>(r2, ku, gu); break;
    case GPSAT_MATERN32: kunit_gunit_of_r2<GPSAT_MATERN32>(r2, ku, gu); break;
    case GPSAT_MATERN52: kunit_gunit_of_r2<GPSAT_MATERN52>(r2, ku, gu); break;
    default: kunit_gunit_of_r2<GPSAT_RBF>(r2, ku, gu); break;
  }
}

// ---- FP64 tensor-core MMA (DMMA.8x8x4) ------------------------------------------------------
// D(8x8) += A(8x4, row) * B(4x8, col).  Lane l holds A[l>>2][l&3], B[k=l&3][n=l>>2],
// C[l>>2][2*(l&3) + {0,1}].  All larger f64 mma shapes lower to this one on sm_100a.
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ---- cp.async (LDGSTS) ----------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N)); }
