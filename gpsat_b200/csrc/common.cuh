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
  double ynorm2;         // |y|^2 (a-priori scale of the observation row of the factor)
  double inv_ls[GPSAT_DMAX];
  double var;            // kernel variance
  double noise;          // likelihood variance
};

// ---- kernel functions -----------------------------------------------------------------------
// Reference: gpflow.kernels.{Matern12,Matern32,Matern52,SquaredExponential} as constructed at
// GPSat/models/gpflow_models.py:116-135.  r2 is computed from direct differences of the scaled
// coordinates (no |a|^2+|b|^2-2ab cancellation); Matern kernels clamp r2 at 1e-36 like GPflow.
// kunit = k / variance;  gunit = g / variance with dk/dl_d = g * (scaled diff_d)^2 / l_d.
template <int KIND>
__device__ __forceinline__ double kunit_of_r2(double r2) {
  if (KIND == GPSAT_RBF) return exp(-0.5 * r2);
  double r = sqrt(fmax(r2, 1e-36));
  if (KIND == GPSAT_MATERN12) return exp(-r);
  if (KIND == GPSAT_MATERN32) { double s = 1.7320508075688772 * r; return (1.0 + s) * exp(-s); }
  double s = 2.23606797749979 * r;
  return (1.0 + s + (5.0 / 3.0) * r * r) * exp(-s);
}

template <int KIND>
__device__ __forceinline__ void kunit_gunit_of_r2(double r2, double& ku, double& gu) {
  if (KIND == GPSAT_RBF) { ku = exp(-0.5 * r2); gu = ku; return; }
  double r = sqrt(fmax(r2, 1e-36));
  if (KIND == GPSAT_MATERN12) {
    ku = exp(-r);
    gu = (r2 > 1e-36) ? ku / r : 0.0;
    return;
  }
  if (KIND == GPSAT_MATERN32) {
    double s = 1.7320508075688772 * r;
    double e = exp(-s);
    ku = (1.0 + s) * e;
    gu = 3.0 * e;
    return;
  }
  double s = 2.23606797749979 * r;
  double e = exp(-s);
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
