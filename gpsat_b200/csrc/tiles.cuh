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
constexpr int SM_TILE_D = NB * LDC;         // 16896 doubles (>= NSTAGE * STAGE_D: the tile aliases the stages)
constexpr int NWSTAGE = 3;                  // B-only stages for streaming W_ii
constexpr int SM_WST_D = NWSTAGE * NB * LDS_S;
constexpr int SM_XY_D = 2 * NB * GPSAT_DMAX + 2 * NB;
constexpr int STEP_SMEM_BYTES = (SM_TILE_D + SM_WST_D + SM_XY_D) * 8;
constexpr int LAUUM_SMEM_BYTES = (SM_TILE_D + SM_XY_D + 64 * 8) * 8;
static_assert(SM_TILE_D >= NSTAGE * STAGE_D, "tile must cover the pipeline stages it aliases");

struct Acc { double c[8][4][2]; };        // 2 x 4 warp grid: warp tile 64 x 32
struct AccStrip { double c[2][16][2]; };  // strip mapping: warp tile 16 x 128

__device__ __forceinline__ void acc_zero(Acc& a) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { a.c[i][j][0] = 0.0; a.c[i][j][1] = 0.0; }
}
__device__ __forceinline__ void acc_zero(AccStrip& a) {
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 16; ++j) { a.c[i][j][0] = 0.0; a.c[i][j][1] = 0.0; }
}

// Row strip of warp w in the strip mapping.  Warps w and w + 4 share a scheduler (warp id mod 4);
// pairing strips s and 7 - s balances triangular work between the four schedulers.
__device__ __forceinline__ int strip_of_warp(int warp) { return warp < 4 ? warp : 11 - warp; }

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

// Same as mma_slab for a slab of an upper-triangular A block (A[a][k] = 0 for k < a, a and k
// local to the 128-block): fragments whose rows all lie below the slab's k range are skipped.
// kl0 = local k of the slab's first column.
__device__ __forceinline__ void mma_slab_tri(Acc& acc, const double* As, int lda, const double* Bs, int ldb,
                                             int nk4, int wm, int wn, int lane, int kl0) {
  const int lr = lane >> 2, lk = lane & 3;
  const double* ap = As + (wm * 64 + lr) * lda + lk;
  const double* bp = Bs + (wn * 32 + lr) * ldb + lk;
#pragma unroll
  for (int k4 = 0; k4 < nk4; ++k4) {
    const int klast = kl0 + 4 * k4 + 3;
    if (klast < wm * 64) continue;
    double b[4];
#pragma unroll
    for (int f = 0; f < 4; ++f) b[f] = bp[f * 8 * ldb + k4 * 4];
#pragma unroll
    for (int fm = 0; fm < 8; ++fm) {
      if (klast >= wm * 64 + fm * 8) {
        const double a = ap[fm * 8 * lda + k4 * 4];
#pragma unroll
        for (int fn = 0; fn < 4; ++fn) dmma884(acc.c[fm][fn][0], acc.c[fm][fn][1], a, b[fn]);
      }
    }
  }
}

// Mirror case: a slab of a lower-triangular A block (A[a][k] = 0 for k > a): fragments whose rows all
// lie above the slab's k range are skipped.
__device__ __forceinline__ void mma_slab_tri_lo(Acc& acc, const double* As, int lda, const double* Bs, int ldb,
                                                int nk4, int wm, int wn, int lane, int kl0) {
  const int lr = lane >> 2, lk = lane & 3;
  const double* ap = As + (wm * 64 + lr) * lda + lk;
  const double* bp = Bs + (wn * 32 + lr) * ldb + lk;
#pragma unroll
  for (int k4 = 0; k4 < nk4; ++k4) {
    const int kfirst = kl0 + 4 * k4;
    if (kfirst > wm * 64 + 63) continue;
    double b[4];
#pragma unroll
    for (int f = 0; f < 4; ++f) b[f] = bp[f * 8 * ldb + k4 * 4];
#pragma unroll
    for (int fm = 0; fm < 8; ++fm) {
      if (kfirst <= wm * 64 + fm * 8 + 7) {
        const double a = ap[fm * 8 * lda + k4 * 4];
#pragma unroll
        for (int fn = 0; fn < 4; ++fn) dmma884(acc.c[fm][fn][0], acc.c[fm][fn][1], a, b[fn]);
      }
    }
  }
}

// Strip mapping, one block of 4 k values: acc[rf][cf] += A[strip rows][k] * B[8cf .. 8cf+8][k]^T for
// the column fragments cf_lo <= cf < cf_hi (warp-uniform bounds).  ap / bp point at this lane's
// element of the k block: ap = A + (16*strip + lr) * lda + k + lk,  bp = B + lr * ldb + k + lk.
__device__ __forceinline__ void mma_strip_k4(AccStrip& acc, const double* ap, int lda, const double* bp, int ldb,
                                             int cf_lo, int cf_hi) {
  const double a0 = ap[0], a1 = ap[8 * lda];
#pragma unroll
  for (int cf = 0; cf < 16; ++cf) {
    if (cf >= cf_lo && cf < cf_hi) {
      const double b = bp[cf * 8 * ldb];
      dmma884(acc.c[0][cf][0], acc.c[0][cf][1], a0, b);
      dmma884(acc.c[1][cf][0], acc.c[1][cf][1], a1, b);
    }
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
// Pipeline prologue of gemm_nt_mainloop: issue the first NSTAGE-1 slab loads.  Split out so that
// callers can put other global loads (coordinates) behind it instead of in front of it.
__device__ __forceinline__ void gemm_nt_prologue(const double* Ag, const double* Bg, int ldg, int kbeg, int kend,
                                                 double* stages, int tid) {
  const int nslab = (kend - kbeg) / KC;
#pragma unroll
  for (int s = 0; s < NSTAGE - 1; ++s) {
    if (s < nslab) {
      double* st = stages + s * STAGE_D;
      load_slab(st, Ag + kbeg + s * KC, ldg, tid);
      load_slab(st + NB * LDS_S, Bg + kbeg + s * KC, ldg, tid);
    }
    cp_async_commit();
  }
}

// acc += sum_{k in [kbeg, kend)} Ag[m][k] * Bg[n][k]   (kbeg, kend multiples of KC), after
// gemm_nt_prologue with the same arguments.  tri_first: the first 128-block of A
// (k in [kbeg, kbeg + 128)) is upper triangular; tri_last: the last one is lower triangular.
__device__ __forceinline__ void gemm_nt_mainloop(Acc& acc, const double* Ag, const double* Bg, int ldg,
                                                 int kbeg, int kend, double* stages, int tid,
                                                 bool tri_first = false, bool tri_last = false) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  const int nslab = (kend - kbeg) / KC;
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
    if (tri_first && s < NB / KC) mma_slab_tri(acc, st, LDS_S, st + NB * LDS_S, LDS_S, KC / 4, wm, wn, lane, s * KC);
    else if (tri_last && s >= nslab - NB / KC)
      mma_slab_tri_lo(acc, st, LDS_S, st + NB * LDS_S, LDS_S, KC / 4, wm, wn, lane, (s - (nslab - NB / KC)) * KC);
    else mma_slab(acc, st, LDS_S, st + NB * LDS_S, LDS_S, KC / 4, wm, wn, lane);
  }
  cp_async_wait<0>();
  __syncthreads();
}

// Row / column coordinates of one 128-block, scaled by 1/lengthscale, into smem as
// xs[dim][row] (structure of arrays: a warp reading 32 consecutive rows of one dimension is
// conflict-free, a warp reading one row is a broadcast); yv[128] gets the observation values of
// the block (0 outside [0, n)).
// (tid, nthr): the threads that take part, numbered from 0 (default: the whole CTA).
__device__ __forceinline__ void load_block_coords(const SlotDesc& d, int g0, double* xs, double* yv, int tid,
                                                  int nthr = GPSAT_THREADS) {
  const int D = d.D, n = d.n, m = d.m;
  for (int idx = tid; idx < NB * D; idx += nthr) {
    int row = idx / D, dd = idx - row * D;
    int g = g0 + row;
    double v = 0.0;
    if (g < n) v = d.X[(size_t)g * D + dd] * d.inv_ls[dd];
    else if (g > n && g < n + 1 + m) v = d.Xs[(size_t)(g - n - 1) * D + dd] * d.inv_ls[dd];
    xs[dd * NB + row] = v;
  }
  for (int row = tid; row < NB; row += nthr) {
    int g = g0 + row;
    yv[row] = (g < n) ? d.y[g] : 0.0;
  }
}

// ---- compact (rolled) evaluation of covariance tiles ------------------------------------------
// Thread t owns column c = t & 127 and the rows r = (t >> 7), (t >> 7) + 2, ...: the column
// coordinates stay in registers, the row coordinates are warp-wide broadcasts, the tile is
// touched with unit stride.  DD > 0 fixes the input dimension at compile time (DD = 3: x, y, t).
template <int DD>
struct ColCoords {
  double x[DD > 0 ? DD : GPSAT_DMAX];
  __device__ __forceinline__ void load(const double* xc, int c, int D) {
#pragma unroll
    for (int dd = 0; dd < (DD > 0 ? DD : GPSAT_DMAX); ++dd) x[dd] = (DD > 0 || dd < D) ? xc[dd * NB + c] : 0.0;
  }
  __device__ __forceinline__ double r2(const double* xr, int r, int D) const {
    double s = 0.0;
#pragma unroll
    for (int dd = 0; dd < (DD > 0 ? DD : GPSAT_DMAX); ++dd)
      if (DD > 0 || dd < D) { double t = xr[dd * NB + r] - x[dd]; s = fma(t, t, s); }
    return s;
  }
};

// Entry (gr, gc) of the augmented matrix  [[K + s2 I], [y^T], [Kxs^T], [0 .. I]].
template <int KIND, int DD>
__device__ __forceinline__ double aug_value(const SlotDesc& d, int gr, int gc, int r, const double* xr,
                                            const ColCoords<DD>& cc, double yc) {
  const int n = d.n;
  if (gc >= n) return (gr == gc && gr > n) ? 1.0 : 0.0;
  if (gr == n) return yc;
  if (gr >= n + 1 + d.m) return 0.0;
  double v = d.var * kunit_of_r2<KIND>(cc.r2(xr, r, d.D));
  if (gr == gc) v += d.noise;
  return v;
}

// Cs[r][c] += aug(g0r + r, g0c + c) for the whole tile (lower_only: c <= r only).
// Tiles whose rows are all observation rows (or all prediction rows) and whose columns are all
// observation columns take a branch-free path: var * k(r2) (+ noise on the global diagonal).
template <int KIND, int DD>
__device__ __forceinline__ void add_aug_tile_t(const SlotDesc& d, int g0r, int g0c, const double* xr,
                                               const double* xc, const double* yc, double* Cs, int tid,
                                               bool lower_only) {
  const int c = tid & (NB - 1);
  const int n = d.n;
  ColCoords<DD> cc;
  cc.load(xc, c, d.D);
  const bool rows_plain = (g0r + NB <= n) || (g0r > n && g0r + NB <= n + 1 + d.m);
  if (rows_plain && g0c + NB <= n) {
    const double var = d.var;
    if (g0r != g0c) {
      // batches of 8 rows: the 8 kernel values are independent dependency chains (the read-modify-write of Cs
      // would otherwise serialise the iterations: one ~36-instruction FP64 chain at a time per warp)
      constexpr int UB = 8, RS = GPSAT_THREADS / NB;
#pragma unroll 1
      for (int r0 = tid >> 7; r0 < NB; r0 += RS * UB) {
        double v[UB];
#pragma unroll
        for (int u = 0; u < UB; ++u) v[u] = var * kunit_of_r2<KIND>(cc.r2(xr, r0 + RS * u, d.D));
#pragma unroll
        for (int u = 0; u < UB; ++u) Cs[(r0 + RS * u) * LDC + c] += v[u];
      }
    } else {
      // diagonal tile of observation rows: + noise on the diagonal; with lower_only the batches that lie above
      // the diagonal for the whole warp are skipped
      const double noise = d.noise;
      constexpr int UB = 8, RS = GPSAT_THREADS / NB;
      const int wc0 = c & ~31;
#pragma unroll 1
      for (int r0 = tid >> 7; r0 < NB; r0 += RS * UB) {
        if (lower_only && r0 + RS * (UB - 1) < wc0) continue;
        double v[UB];
#pragma unroll
        for (int u = 0; u < UB; ++u) v[u] = var * kunit_of_r2<KIND>(cc.r2(xr, r0 + RS * u, d.D));
#pragma unroll
        for (int u = 0; u < UB; ++u) {
          const int r = r0 + RS * u;
          if (!(lower_only && c > r)) Cs[r * LDC + c] += (r == c) ? v[u] + noise : v[u];
        }
      }
    }
    return;
  }
  // tiles that contain the observation row, prediction rows next to padding, or identity columns: the kernel value
  // is formed for every element (coordinates outside the data are zeros) and then selected, so that the batches stay
  // branch-free
  const double ycv = yc[c];
  const double var = d.var, noise = d.noise;
  const int gc = g0c + c, nm = n + 1 + d.m;
  constexpr int UB = 8, RS = GPSAT_THREADS / NB;
  const int wc0 = c & ~31;
#pragma unroll 1
  for (int r0 = tid >> 7; r0 < NB; r0 += RS * UB) {
    if (lower_only && r0 + RS * (UB - 1) < wc0) continue;
    double v[UB];
#pragma unroll
    for (int u = 0; u < UB; ++u) v[u] = var * kunit_of_r2<KIND>(cc.r2(xr, r0 + RS * u, d.D));
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const int r = r0 + RS * u, gr = g0r + r;
      double w = (gr == gc) ? v[u] + noise : v[u];
      if (gr >= nm) w = 0.0;
      if (gr == n) w = ycv;
      if (gc >= n) w = (gr == gc && gr > n) ? 1.0 : 0.0;
      if (!(lower_only && c > r)) Cs[r * LDC + c] += w;
    }
  }
}

// Precomputed matrix instead of a covariance function (the B matrix of sparse GPR): same augmented
// layout, entries below row n come from d.pre.
__device__ __forceinline__ void add_pre_tile(const SlotDesc& d, int g0r, int g0c, const double* yc, double* Cs,
                                             int tid, bool lower_only) {
  const int c = tid & (NB - 1), gc = g0c + c, n = d.n;
  const double ycv = yc[c];
#pragma unroll 4
  for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
    if (lower_only && c > r) continue;
    const int gr = g0r + r;
    double v;
    if (gc >= n) v = (gr == gc && gr > n) ? 1.0 : 0.0;
    else if (gr == n) v = ycv;
    else if (gr < n) v = d.pre[(size_t)gr * d.ldpre + gc];
    else v = (gr < n + 1 + d.m) ? d.pre_s[(size_t)(gr - n - 1) * d.ldpre + gc] : 0.0;
    Cs[r * LDC + c] += v;
  }
}

__device__ __noinline__ void add_aug_tile(const SlotDesc& d, int g0r, int g0c, const double* xr, const double* xc,
                                          const double* yc, double* Cs, int tid, bool lower_only) {
  if (d.kind == GPSAT_PRECOMPUTED) { add_pre_tile(d, g0r, g0c, yc, Cs, tid, lower_only); return; }
  const bool d3 = d.D == 3;
  switch (d.kind) {
    case GPSAT_MATERN12:
      if (d3) add_aug_tile_t<GPSAT_MATERN12, 3>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      else add_aug_tile_t<GPSAT_MATERN12, 0>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      break;
    case GPSAT_MATERN32:
      if (d3) add_aug_tile_t<GPSAT_MATERN32, 3>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      else add_aug_tile_t<GPSAT_MATERN32, 0>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      break;
    case GPSAT_MATERN52:
      if (d3) add_aug_tile_t<GPSAT_MATERN52, 3>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      else add_aug_tile_t<GPSAT_MATERN52, 0>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      break;
    default:
      if (d3) add_aug_tile_t<GPSAT_RBF, 3>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      else add_aug_tile_t<GPSAT_RBF, 0>(d, g0r, g0c, xr, xc, yc, Cs, tid, lower_only);
      break;
  }
}

// Accumulators to a smem tile [128][LDC] (2 x 4 warp grid layout).
__device__ __forceinline__ void acc_to_smem(const Acc& acc, double* Cs, int tid, double sign) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  const int lr = lane >> 2, lk = lane & 3;
#pragma unroll
  for (int fm = 0; fm < 8; ++fm)
#pragma unroll
    for (int fn = 0; fn < 4; ++fn) {
      int r = wm * 64 + fm * 8 + lr, c = wn * 32 + fn * 8 + 2 * lk;
      double2 v = make_double2(sign * acc.c[fm][fn][0], sign * acc.c[fm][fn][1]);
      *reinterpret_cast<double2*>(Cs + r * LDC + c) = v;
    }
}

}  // namespace gpsat
