// Tile kernels of the batched factor / invert / gradient sweep (sm_100a, FP64 DMMA).
//
// Per expert and per evaluation of the objective the reference does, inside TensorFlow,
//   K = k(X, X); L = chol(K + s2 I); lml = ...; autodiff for the gradient
//   (GPSat/models/gpflow_models.py:317-321, 337 -> gpflow.models.GPR.log_marginal_likelihood).
// Here the same quantities come from a left-looking blocked sweep over 128 x 128 tiles:
//   for i in 0 .. T-1:
//     diag_kernel(i)  : L_ii = chol(S_ii), W_ii = L_ii^-1          (in shared memory)
//     step_kernel(i)  : potrf tiles  L_ri = (K_ri - sum_k<i L_rk L_ik^T) W_ii^T,  r > i
//                                    S_rr -= L_ri L_ri^T
//                       trtri tiles  U_ji = -(sum_{j<=k<i} U_jk L_ik^T) W_ii^T,   j < i   (U = L^-T)
//   lauum_grad_kernel : Kinv_ij = sum_k U_ik U_jk^T, reduced on the fly against dK/dtheta
// The observation vector rides along as row n of the matrix, so a = L^-1 y and
// alpha = Ks^-1 y = -U[:, n] fall out of the same tiles; prediction rows ride along as rows
// n+1 .. n+m (potrf tiles only).
#pragma once
#include "common.cuh"

namespace gpsat {

constexpr int NB = GPSAT_NB;
constexpr int KC = 16;                 // k-slab per pipeline stage
constexpr int LDS_S = KC + 4;          // padded smem row stride of a slab (conflict-free DMMA loads)
constexpr int LDC = NB + 4;            // padded smem row stride of a full tile
constexpr int STAGE_D = 2 * NB * LDS_S;     // doubles per stage (A slab + B slab)
constexpr int NSTAGE = 3;
constexpr int SM_TILE_D = NB * LDC;         // 16896 doubles
constexpr int SM_WST_D = 2 * NB * LDS_S;    // two B-only stages for streaming W_ii
constexpr int SM_XY_D = 2 * NB * GPSAT_DMAX + 2 * NB;
constexpr int STEP_SMEM_BYTES = (SM_TILE_D + SM_WST_D + SM_XY_D) * 8;
constexpr int LAUUM_SMEM_BYTES = (NSTAGE * STAGE_D + SM_XY_D + 64 * 8) * 8;
constexpr int LDD = NB + 1;                 // diag kernel smem stride
constexpr int DIAG_SMEM_BYTES = (NB * LDD + NB) * 8;

struct Acc { double c[8][4][2]; };

__device__ __forceinline__ void acc_zero(Acc& a) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { a.c[i][j][0] = 0.0; a.c[i][j][1] = 0.0; }
}

// acc += A[wm*64 .. +64][k] * B[wn*32 .. +32][k]^T over nk4*4 values of k.  As/Bs point at k = 0
// of the slab; both are k-contiguous with row strides lda / ldb (strides = 4 mod 16 doubles).
__device__ __forceinline__ void mma_slab(Acc& acc, const double* As, int lda, const double* Bs, int ldb,
                                         int nk4, int wm, int wn, int lane) {
  const int lr = lane >> 2, lk = lane & 3;
  const double* ap = As + (wm * 64 + lr) * lda + lk;
  const double* bp = Bs + (wn * 32 + lr) * ldb + lk;
#pragma unroll 2
  for (int k4 = 0; k4 < nk4; ++k4) {
    double a[8], b[4];
#pragma unroll
    for (int f = 0; f < 8; ++f) a[f] = ap[f * 8 * lda + k4 * 4];
#pragma unroll
    for (int f = 0; f < 4; ++f) b[f] = bp[f * 8 * ldb + k4 * 4];
#pragma unroll
    for (int fm = 0; fm < 8; ++fm)
#pragma unroll
      for (int fn = 0; fn < 4; ++fn) dmma884(acc.c[fm][fn][0], acc.c[fm][fn][1], a[fm], b[fn]);
  }
}

// cp.async one 128 x KC slab (row-major source, row stride ldg doubles) into smem [128][LDS_S].
__device__ __forceinline__ void load_slab(double* dst, const double* src, int ldg, int tid) {
#pragma unroll
  for (int q = 0; q < (NB * KC / 2) / GPSAT_THREADS; ++q) {
    int id = tid + q * GPSAT_THREADS;
    int row = id >> 3, ch = id & 7;
    cp_async16(dst + row * LDS_S + ch * 2, src + (size_t)row * ldg + ch * 2);
  }
}

// acc += sum_{k in [kbeg, kend)} Ag[m][k] * Bg[n][k]   (kbeg, kend multiples of KC)
__device__ __forceinline__ void gemm_nt_mainloop(Acc& acc, const double* Ag, const double* Bg, int ldg,
                                                 int kbeg, int kend, double* stages, int tid) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  const int nslab = (kend - kbeg) / KC;
  if (nslab <= 0) return;
#pragma unroll
  for (int s = 0; s < NSTAGE - 1; ++s) {
    if (s < nslab) {
      double* st = stages + s * STAGE_D;
      load_slab(st, Ag + kbeg + s * KC, ldg, tid);
      load_slab(st + NB * LDS_S, Bg + kbeg + s * KC, ldg, tid);
    }
    cp_async_commit();
  }
  for (int s = 0; s < nslab; ++s) {
    cp_async_wait<NSTAGE - 2>();
    __syncthreads();
    int nx = s + NSTAGE - 1;
    if (nx < nslab) {
      double* st = stages + (nx % NSTAGE) * STAGE_D;
      load_slab(st, Ag + kbeg + nx * KC, ldg, tid);
      load_slab(st + NB * LDS_S, Bg + kbeg + nx * KC, ldg, tid);
    }
    cp_async_commit();
    const double* st = stages + (s % NSTAGE) * STAGE_D;
    mma_slab(acc, st, LDS_S, st + NB * LDS_S, LDS_S, KC / 4, wm, wn, lane);
  }
  cp_async_wait<0>();
  __syncthreads();
}

// Row / column coordinates of one 128-block, scaled by 1/lengthscale, into smem [128][DMAX];
// yv[128] gets the observation values of the block (0 outside [0, n)).
__device__ __forceinline__ void load_block_coords(const SlotDesc& d, int g0, double* xs, double* yv, int tid) {
  const int D = d.D, n = d.n, m = d.m;
  for (int idx = tid; idx < NB * D; idx += GPSAT_THREADS) {
    int row = idx / D, dd = idx - row * D;
    int g = g0 + row;
    double v = 0.0;
    if (g < n) v = d.X[(size_t)g * D + dd] * d.inv_ls[dd];
    else if (g > n && g < n + 1 + m) v = d.Xs[(size_t)(g - n - 1) * D + dd] * d.inv_ls[dd];
    xs[row * GPSAT_DMAX + dd] = v;
  }
  for (int row = tid; row < NB; row += GPSAT_THREADS) {
    int g = g0 + row;
    yv[row] = (g < n) ? d.y[g] : 0.0;
  }
}

// Entry (gr, gc) of the augmented matrix  [[K + s2 I], [y^T], [Kxs^T], [0 .. I]]  from smem coords.
__device__ __forceinline__ double tile_value(const SlotDesc& d, int gr, int gc, const double* xr,
                                             const double* xc, double yc) {
  const int n = d.n;
  if (gc >= n) return (gr == gc && gr > n) ? 1.0 : 0.0;
  if (gr == n) return yc;
  if (gr >= n + 1 + d.m) return 0.0;
  double r2 = 0.0;
  for (int dd = 0; dd < d.D; ++dd) { double t = xr[dd] - xc[dd]; r2 += t * t; }
  double v = d.var * kunit_dyn(d.kind, r2);
  if (gr == gc) v += d.noise;
  return v;
}

}  // namespace gpsat
