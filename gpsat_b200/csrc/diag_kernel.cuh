// diag_kernel: in-shared-memory Cholesky + triangular inverse of one 128 x 128 diagonal tile
// per expert ("batched in-SM Cholesky of K + s2 I" for experts with n < 128, and the panel
// step of the blocked factorisation for larger ones).
//
// Blocked right-looking factorisation with 32-wide panels, all in shared memory:
//   for p in 0..3:  warp 0 factors the 32 x 32 diagonal block in registers (one row per lane);
//                   full panels: warps 0..2 solve the rows below against L_pp (one row per thread) while warp 7
//                   inverts the block off the critical path; partial panels (the tile with the y row): warp 0
//                   inverts it (one column per lane) and all warps multiply the rows below by W_pp^T (DMMA);
//                   trailing block -= panel panel^T (DMMA, lower fragments only)
//   then U = L^-T block by block (the 32-granular twin of the trtri tiles of step_kernel).
// The tile is "augmented": local row ncol (global row n) is the observation row y^T and rides
// along, later rows are prediction rows / padding.  Columns >= ncol are identity columns.
#pragma once
#include "tiles.cuh"
#include "i8slice.cuh"

namespace gpsat {

__device__ unsigned long long g_step_phase[32];   // see step_kernel.cuh
#ifdef GPSAT_STEP_PHASES
#define DIAG_PHASE(k) do { if (tid == 0) { long long now_ = clock64(); atomicAdd(&g_step_phase[k], (unsigned long long)(now_ - tdiag)); tdiag = now_; } } while (0)
#else
#define DIAG_PHASE(k) do { } while (0)
#endif

constexpr int LDD = NB + 4;                 // tile stride: DMMA fragment loads are conflict-free
constexpr int PB = 32;                      // panel width
constexpr int LDP = PB + 4;                 // stride of the 32 x 32 side blocks
constexpr int DIAG_CS_D = NB * LDD;
constexpr int DIAG_WP_D = 4 * PB * LDP;     // W_pp, row-major, p = 0..3
constexpr int DIAG_TT_D = 3 * PB * LDP;     // temporaries of the inverse sweep
constexpr int DIAG_AUX_D = DIAG_WP_D + DIAG_TT_D + NB + 2 * PB + 8;   // scratch of diag_factor_tile
constexpr int DIAG_XY_D = NB * GPSAT_DMAX + NB;
constexpr int DIAG_SMEM_BYTES = (DIAG_CS_D + DIAG_AUX_D + DIAG_XY_D) * 8;
static_assert(LDD == LDC, "add_aug_tile writes with stride LDC");
static_assert(DIAG_AUX_D <= SM_WST_D + SM_XY_D, "step_kernel lends its W stages + coordinate arrays as scratch");

// One 8 x 8 fragment: c += sa * A[8 rows][4*nk4 k] * B[8 rows][same k]^T (both k-contiguous).
__device__ __forceinline__ void frag_nt(double& c0, double& c1, const double* A, int lda, const double* B,
                                        int ldb, int nk4, int lane, double sa) {
  const double* ap = A + (lane >> 2) * lda + (lane & 3);
  const double* bp = B + (lane >> 2) * ldb + (lane & 3);
#pragma unroll 4
  for (int k4 = 0; k4 < nk4; ++k4) dmma884(c0, c1, sa * ap[4 * k4], bp[4 * k4]);
}

// ---- full 32-wide panels (every column an observation column): the three pieces of the fast path ----------------
// (1) Cholesky of the 32 x 32 diagonal block by one warp, lane = row.  The dependent chain per pivot is
//     shuffle -> rsqrt -> one multiply -> one FMA: every lane tracks its own diagonal entry in `dcur` (a_ll -= v^2 needs
//     no other lane's value, so the next pivot does not wait for the column's round trip through shared memory), and
//     L_jj = sqrt(pivot) and 1 / L_jj are refined once per lane after the loop, all 32 pivots in parallel, instead
//     of five dependent FP64 operations inside every step.  Returns log(L_ll) of the lane's pivot.
__device__ __forceinline__ double chol32_fast(double* Cs, int c0, double* colb, double* wdiag, int lane, bool& bad) {
  double a[PB];
#pragma unroll
  for (int k = 0; k < PB; ++k) a[k] = Cs[(c0 + lane) * LDD + c0 + k];
  double dcur = 0.0, mypiv = 1.0;
#pragma unroll
  for (int k = 0; k < PB; ++k) if (k == lane) dcur = a[k];
#pragma unroll
  for (int j = 0; j < PB; ++j) {
    double piv = __shfl_sync(0xffffffffu, dcur, j);
    const bool nok = !(piv > 0.0) || !isfinite(piv);
    bad = bad || nok;
    piv = nok ? 1.0 : piv;
    if (lane == j) mypiv = piv;
    const double inv0 = rsqrt(piv);
    const double v = (lane > j) ? a[j] * inv0 : 0.0;
    dcur = fma(-v, v, dcur);
    a[j] = v;
    double* cb = colb + (j & 1) * PB;
    cb[lane] = v;
    __syncwarp();
    // a[k] -= L[lane][j] * L[k][j]; entries with k > lane are never read, so no predicate
#pragma unroll
    for (int k = j + 1; k < PB; ++k) a[k] = fma(-v, cb[k], a[k]);
  }
  const double inv0 = rsqrt(mypiv);
  double dg = mypiv * inv0;
  dg = fma(fma(-dg, dg, mypiv), 0.5 * inv0, dg);
  const double inv = fma(fma(-dg, inv0, 1.0), inv0, inv0);
  wdiag[c0 + lane] = inv;
#pragma unroll
  for (int k = 0; k < PB; ++k) if (k == lane) a[k] = dg;
#pragma unroll
  for (int k = 0; k < PB; ++k) if (k <= lane) Cs[(c0 + lane) * LDD + c0 + k] = a[k];
  return log(dg);
}

// (2) inverse of the factored block by one warp, lane = column c of W_pp: x[r] = W[r][c]; W_pp goes to W (row-major,
//     stride LDP) and its transpose U_pp into the strict upper part of the block in Cs.
__device__ __forceinline__ void inv32(double* Cs, int c0, const double* wdiag, double* W, int lane) {
  double x[PB], acc[PB];
#pragma unroll
  for (int r = 0; r < PB; ++r) { x[r] = 0.0; acc[r] = 0.0; }
  const double myinv = wdiag[c0 + lane];
#pragma unroll
  for (int r = 0; r < PB; ++r) if (r == lane) x[r] = myinv;
#pragma unroll
  for (int k = 0; k < PB - 1; ++k) {
    const double xk = x[k];
#pragma unroll
    for (int r = k + 1; r < PB; ++r) acc[r] = fma(Cs[(c0 + r) * LDD + c0 + k], xk, acc[r]);
    if (k + 1 > lane) x[k + 1] = -acc[k + 1] * wdiag[c0 + k + 1];
  }
#pragma unroll
  for (int r = 0; r < PB; ++r) {
    W[r * LDP + lane] = x[r];                                  // W_pp[r][c], zero above the diagonal
    if (r > lane) Cs[(c0 + lane) * LDD + c0 + r] = x[r];      // U_pp = W_pp^T, strict upper
  }
}

// (3) one row below the block: x L_pp^T = a by forward substitution (two dependent operations per column), in place.
//     Replaces the product with the explicit inverse, which therefore leaves the critical path.
__device__ __forceinline__ void rowsolve32(double* Cs, int c0, const double* wdiag, int R) {
  double x[PB];
  double* row = Cs + R * LDD + c0;
#pragma unroll
  for (int k = 0; k < PB; ++k) x[k] = row[k];
#pragma unroll
  for (int j = 0; j < PB; ++j) {
    x[j] *= wdiag[c0 + j];
#pragma unroll
    for (int k = j + 1; k < PB; ++k) x[k] = fma(-x[j], Cs[(c0 + k) * LDD + c0 + j], x[k]);
  }
#pragma unroll
  for (int k = 0; k < PB; ++k) row[k] = x[k];
}

__device__ __forceinline__ void diag_bar_w06() { asm volatile("bar.sync 3, 224;\n" ::: "memory"); }

// Factor the diagonal tile i of slot d held in shared memory: on entry the lower triangle of
// Cs[128][LDD] is the Schur complement S_ii (the strict upper part may hold anything finite); on
// exit L_ii / W_ii are in the slot's side buffers, U_ii in the slot matrix, logdet[i] (and quad,
// notpd) are set.  aux: DIAG_AUX_D doubles of scratch.  All 256 threads of the CTA must call it.
__device__ __noinline__ void diag_factor_tile(const SlotDesc& d, int i, double* Cs, double* aux, int tid) {
  double* Wp = aux;                         // [4][32][LDP]  W_pp (lower triangular, with diagonal)
  double* Tt = Wp + DIAG_WP_D;              // [3][32][LDP]
  double* wdiag = Tt + DIAG_TT_D;           // [NB] 1 / L_cc
  double* colb = wdiag + NB;                // [2][32] column broadcast buffer of the panel warp
  double* misc = colb + 2 * PB;             // [0] log-determinant accumulator
  const int warp = tid >> 5, lane = tid & 31, lr = lane >> 2, lk = lane & 3;
  const int row0 = i * NB;
  const int n = d.n, ld = d.ld;
  int ncol = n - row0;               // real columns in this tile; local index ncol is the y row
  if (ncol > NB) ncol = NB;
  if (ncol < 0) ncol = 0;
  if (tid == 0) misc[0] = 0.0;
  __syncthreads();
#ifdef GPSAT_STEP_PHASES
  long long tdiag = clock64();
#endif

  bool bad = false;
  for (int p = 0; p < NB / PB; ++p) {
    const int c0 = p * PB;
    int jmax = ncol - c0;
    if (jmax > PB) jmax = PB;
    if (jmax < 0) jmax = 0;
    if (jmax == PB) {
      // ---- full panel: warp 0 factors the block; then warp 7 inverts it off the critical path (nothing before the
      // U sweep below reads W_pp) while warps 0..6 solve the rows below against L_pp and update the trailing block ----
      if (warp == 0) {
        double lg = chol32_fast(Cs, c0, colb, wdiag, lane, bad);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) lg += __shfl_xor_sync(0xffffffffu, lg, o);
        if (lane == 0) misc[0] += lg;
      }
      __syncthreads();                 // L_pp and 1 / L_jj visible; warp 7 is done with the previous block's inverse
      const int nbelow = NB - c0 - PB;
      if (warp == 7) {
        inv32(Cs, c0, wdiag, Wp + p * PB * LDP, lane);
      } else if (nbelow > 0) {
        if (tid < nbelow) rowsolve32(Cs, c0, wdiag, c0 + PB + tid);
        diag_bar_w06();
        const int nf = nbelow / 8;
        const double* P0 = Cs + (c0 + PB) * LDD + c0;
        for (int f = warp; f < nf * (nf + 1) / 2; f += 7) {
          int fr = (int)((sqrtf(8.0f * f + 1.0f) - 1.0f) * 0.5f);
          while (fr * (fr + 1) / 2 > f) --fr;
          while ((fr + 1) * (fr + 2) / 2 <= f) ++fr;
          const int fc = f - fr * (fr + 1) / 2;
          double* Cg = Cs + (c0 + PB + 8 * fr + lr) * LDD + c0 + PB + 8 * fc + 2 * lk;
          double2 cv = *reinterpret_cast<double2*>(Cg);
          frag_nt(cv.x, cv.y, P0 + 8 * fr * LDD, LDD, P0 + 8 * fc * LDD, LDD, PB / 4, lane, -1.0);
          *reinterpret_cast<double2*>(Cg) = cv;
        }
        diag_bar_w06();                // trailing block complete before warp 0 factors the next diagonal block
      }
      continue;
    }
    // ---- (a) warp 0: Cholesky of the 32 x 32 diagonal block, lane = row; then its inverse ----
    if (warp == 0) {
      double a[PB];
#pragma unroll
      for (int k = 0; k < PB; ++k) a[k] = Cs[(c0 + lane) * LDD + c0 + k];
#pragma unroll
      for (int j = 0; j < PB; ++j) {
        if (j < jmax) {
          double piv = __shfl_sync(0xffffffffu, a[j], j);
          if (!(piv > 0.0) || !isfinite(piv)) { bad = true; piv = 1.0; }
          // 1/sqrt first (one MUFU + Newton), then L_jj = piv * inv with one correction step:
          // shorter dependent chain than sqrt followed by a division
          double inv = rsqrt(piv);
          double dg = piv * inv;
          dg = fma(fma(-dg, dg, piv), 0.5 * inv, dg);
          inv = fma(fma(-dg, inv, 1.0), inv, inv);
          const double v = (lane > j) ? a[j] * inv : (lane == j ? dg : 0.0);
          a[j] = v;
          double* cb = colb + (j & 1) * PB;
          cb[lane] = v;
          if (lane == 0) wdiag[c0 + j] = inv;
          __syncwarp();
          // a[k] -= L[lane][j] * L[k][j]; entries with k > lane are never read, so no predicate
#pragma unroll
          for (int k = j + 1; k < PB; ++k) a[k] = fma(-v, cb[k], a[k]);
        } else {
          if (j == jmax && lane == j && c0 + j == ncol) *d.quad = -a[j];   // Schur complement at the y row
          a[j] = (lane == j) ? 1.0 : 0.0;
          if (lane == 0) wdiag[c0 + j] = 1.0;
        }
      }
      // log-determinant: lane l holds L_ll
      {
        double dsel = 1.0;          // a[lane] through static indices only (a[] must stay in registers)
#pragma unroll
        for (int k = 0; k < PB; ++k) if (k == lane) dsel = a[k];
        double lg = (lane < jmax) ? log(dsel) : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) lg += __shfl_xor_sync(0xffffffffu, lg, o);
        if (lane == 0) misc[0] += lg;
      }
#pragma unroll
      for (int k = 0; k < PB; ++k) if (k <= lane) Cs[(c0 + lane) * LDD + c0 + k] = a[k];
      __syncwarp();
      // inverse, lane = column c of W_pp: x[r] = W[r][c]
      {
        double x[PB], acc[PB];
#pragma unroll
        for (int r = 0; r < PB; ++r) { x[r] = 0.0; acc[r] = 0.0; }
        const double myinv = wdiag[c0 + lane];
#pragma unroll
        for (int r = 0; r < PB; ++r) if (r == lane) x[r] = myinv;
#pragma unroll
        for (int k = 0; k < PB - 1; ++k) {
          const double xk = x[k];
#pragma unroll
          for (int r = k + 1; r < PB; ++r) acc[r] = fma(Cs[(c0 + r) * LDD + c0 + k], xk, acc[r]);
          if (k + 1 > lane) x[k + 1] = -acc[k + 1] * wdiag[c0 + k + 1];
        }
        double* W = Wp + p * PB * LDP;
#pragma unroll
        for (int r = 0; r < PB; ++r) {
          W[r * LDP + lane] = x[r];                                  // W_pp[r][c], zero above the diagonal
          if (r > lane) Cs[(c0 + lane) * LDD + c0 + r] = x[r];      // U_pp = W_pp^T, strict upper
        }
      }
    }
    __syncthreads();
    const int nbelow = NB - c0 - PB;          // rows below the diagonal block
    if (nbelow > 0) {
      // ---- (b) panel: rows below *= W_pp^T; a warp owns whole 8-row blocks (in-place safe) ----
      const double* W = Wp + p * PB * LDP;
      for (int rb = warp; rb < nbelow / 8; rb += GPSAT_THREADS / 32) {
        double* Arow = Cs + (c0 + PB + rb * 8) * LDD + c0;
        double c[4][2];
#pragma unroll
        for (int f = 0; f < 4; ++f) { c[f][0] = 0.0; c[f][1] = 0.0; }
#pragma unroll
        for (int k4 = 0; k4 < PB / 4; ++k4) {
          const double av = Arow[lr * LDD + 4 * k4 + lk];
#pragma unroll
          for (int f = 0; f < 4; ++f)
            if (8 * f + 7 >= 4 * k4)             // W_pp lower triangular: k <= c
              dmma884(c[f][0], c[f][1], av, W[(8 * f + lr) * LDP + 4 * k4 + lk]);
        }
        __syncwarp();
#pragma unroll
        for (int f = 0; f < 4; ++f) {
          const int cc = c0 + 8 * f + 2 * lk;      // identity columns get zeros
          double2 v = make_double2(cc < ncol ? c[f][0] : 0.0, cc + 1 < ncol ? c[f][1] : 0.0);
          *reinterpret_cast<double2*>(Arow + lr * LDD + 8 * f + 2 * lk) = v;
        }
      }
      __syncthreads();
      // ---- (c) trailing block -= panel panel^T, lower 8 x 8 fragments ----
      const int nf = nbelow / 8;
      const double* P0 = Cs + (c0 + PB) * LDD + c0;
      for (int f = warp; f < nf * (nf + 1) / 2; f += GPSAT_THREADS / 32) {
        int fr = (int)((sqrtf(8.0f * f + 1.0f) - 1.0f) * 0.5f);
        while (fr * (fr + 1) / 2 > f) --fr;
        while ((fr + 1) * (fr + 2) / 2 <= f) ++fr;
        const int fc = f - fr * (fr + 1) / 2;
        double* Cg = Cs + (c0 + PB + 8 * fr + lr) * LDD + c0 + PB + 8 * fc + 2 * lk;
        double2 cv = *reinterpret_cast<double2*>(Cg);
        frag_nt(cv.x, cv.y, P0 + 8 * fr * LDD, LDD, P0 + 8 * fc * LDD, LDD, PB / 4, lane, -1.0);
        *reinterpret_cast<double2*>(Cg) = cv;
      }
      __syncthreads();
    }
  }
  __syncthreads();                       // the last block's inverse (warp 7 on the fast path) is complete
  DIAG_PHASE(11);
  if (bad && tid == 0) *d.notpd = 1;     // `bad` is only ever set in warp 0 (all lanes agree)
  if (tid == 0) d.logdet[i] = misc[0];

  // ---- U = L^-T off-diagonal 32-blocks:  U_ji = -(sum_{j<=k<i} U_jk L_ik^T) W_ii^T ----
  for (int dlt = 1; dlt < NB / PB; ++dlt) {
    const int nblk = NB / PB - dlt;
    // T_b = sum_k U_jk L_ik^T
    for (int f = warp; f < nblk * 16; f += GPSAT_THREADS / 32) {
      const int b = f >> 4, fa = (f >> 2) & 3, fc = f & 3;
      const int j0 = b * PB, i0 = (b + dlt) * PB;
      double c0v = 0.0, c1v = 0.0;
      // k = j: U_jj = W_jj^T, read transposed from Wp
      {
        const double* Wj = Wp + b * PB * LDP;
        const double* Lb = Cs + (i0 + 8 * fc + lr) * LDD + j0 + lk;
#pragma unroll
        for (int k4 = 0; k4 < PB / 4; ++k4)
          dmma884(c0v, c1v, Wj[(4 * k4 + lk) * LDP + 8 * fa + lr], Lb[4 * k4]);
      }
      if (dlt > 1)
        frag_nt(c0v, c1v, Cs + (j0 + 8 * fa) * LDD + j0 + PB, LDD, Cs + (i0 + 8 * fc) * LDD + j0 + PB, LDD,
                (dlt - 1) * PB / 4, lane, 1.0);
      *reinterpret_cast<double2*>(Tt + b * PB * LDP + (8 * fa + lr) * LDP + 8 * fc + 2 * lk) = make_double2(c0v, c1v);
    }
    __syncthreads();
    for (int f = warp; f < nblk * 16; f += GPSAT_THREADS / 32) {
      const int b = f >> 4, fa = (f >> 2) & 3, fc = f & 3;
      const int j0 = b * PB, i0 = (b + dlt) * PB;
      double c0v = 0.0, c1v = 0.0;
      frag_nt(c0v, c1v, Tt + b * PB * LDP + 8 * fa * LDP, LDP, Wp + (b + dlt) * PB * LDP + 8 * fc * LDP, LDP,
              PB / 4, lane, -1.0);
      *reinterpret_cast<double2*>(Cs + (j0 + 8 * fa + lr) * LDD + i0 + 8 * fc + 2 * lk) = make_double2(c0v, c1v);
    }
    __syncthreads();
  }

  DIAG_PHASE(12);
  // ---- write L_ii, W_ii (side buffers) and U_ii = W_ii^T (into the slot matrix) ----
  double* Ld = d.Ld + (size_t)i * NB * NB;
  double* Wd = d.Wd + (size_t)i * NB * NB;
  double* Ut = d.A + (size_t)row0 * ld + row0;
  {
    // W_ii is what the tiles of step i multiply by.  L_ii is read back only for the last diagonal block (prediction
    // reduction) and by the debug extraction, the FP64 copy of U_ii only by the FP64 K^-1 kernel, the debug extraction and -
    // the block that holds column n - the alpha column of the gradient: slots whose consumers all read the planes
    // (SlotDesc::skip_fp64) store neither.
    const bool st_l = !(d.skip_fp64 && i < d.T - 1);
    const bool st_u = !(d.skip_fp64 && d.Uq8 != nullptr && (i + 1) * NB <= n);
    const int c = tid & (NB - 1);
    for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
      double lv, wv, uv;
      if (c < r) { lv = Cs[r * LDD + c]; wv = Cs[c * LDD + r]; uv = 0.0; }
      else if (c == r) { lv = Cs[r * LDD + r]; wv = wdiag[r]; uv = wv; }
      else { lv = 0.0; wv = 0.0; uv = Cs[r * LDD + c]; }
      if (st_l) Ld[r * NB + c] = lv;
      Wd[r * NB + c] = wv;
      if (st_u) Ut[(size_t)r * ld + c] = uv;
    }
  }
  // later trtri tiles and the gradient's K^-1 products read U_ii on the int8 path
  if (d.Uq8 != nullptr)
    slice_tile_udiag(d, ncol, Cs, wdiag, d.Uq8 + (size_t)utile_index(d.T, i, i) * I8L_TILE_B, tid);
  DIAG_PHASE(13);
}

// Stand-alone launch: tile (0,0) of every slot (evaluates the covariance itself); the later
// diagonal tiles are factored inside step_kernel by the CTA that completes their Schur complement.
__global__ void __launch_bounds__(GPSAT_THREADS, 1) diag_kernel(const SlotDesc* __restrict__ slots, int i) {
  const SlotDesc& d = slots[blockIdx.x];
  if (d.expert < 0 || i >= d.T) return;
  extern __shared__ double smem[];
  double* Cs = smem;
  double* aux = Cs + DIAG_CS_D;
  double* xr = aux + DIAG_AUX_D;            // [DMAX][NB] coordinates (tile (0,0) only)
  double* yr = xr + NB * GPSAT_DMAX;        // [NB]
  const int tid = threadIdx.x;
  if (i == 0) {
    load_block_coords(d, 0, xr, yr, tid);
    for (int idx = tid; idx < NB * LDD; idx += GPSAT_THREADS) Cs[idx] = 0.0;
    __syncthreads();
    add_aug_tile(d, 0, 0, xr, xr, yr, Cs, tid, true);
  } else {
    // whole tile through cp.async (all requests in flight at once)
    const int row0 = i * NB, ld = d.ld;
    const double* src = d.A + (size_t)row0 * ld + row0;
    for (int id = tid; id < NB * NB / 2; id += GPSAT_THREADS) {
      const int r = id >> 6, ch = id & 63;
      cp_async16(Cs + r * LDD + ch * 2, src + (size_t)r * ld + ch * 2);
    }
    cp_async_commit();
    cp_async_wait<0>();
  }
  __syncthreads();
  diag_factor_tile(d, i, Cs, aux, tid);
}

}  // namespace gpsat