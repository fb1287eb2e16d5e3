// lauum_grad_kernel: Ks^-1 = U U^T tile by tile, contracted at once with dKs/dtheta so that the
// inverse is never written to memory.  Output: per-tile partial sums
//   gpart[tile][d]     = sum w Q_ab gunit_ab sdiff_d^2      (d < D)       -> d lml/d l_d = var/(2 l_d) * sum
//   gpart[tile][D]     = sum w Q_ab kunit_ab                               -> d lml/d var = 1/2 * sum
//   gpart[tile][D + 1] = sum_a Q_aa                                        -> d lml/d s2  = 1/2 * sum
// with Q = alpha alpha^T - Ks^-1 and w = 2 off the diagonal (lower triangle only), 1 on it.
// (analytic form of the gradient TensorFlow autodiff produces for training_loss,
//  GPSat/models/gpflow_models.py:317-321)
#pragma once
#include "tiles.cuh"

namespace gpsat {

// Rolled gradient reduction over one tile of Q = alpha alpha^T - Ks^-1 held in shared memory.
// Same thread -> element mapping as add_aug_tile_t (column in registers, rows broadcast).
template <int KIND, int DD>
__device__ __forceinline__ void grad_tile_t(const SlotDesc& d, int g0r, int g0c, const double* xr, const double* xc,
                                            const double* ar, const double* ac, const double* Cs, int tid,
                                            double* part, double aac) {
  const int n = d.n, D = d.D;
  const int c = tid & (NB - 1);
  const int gj = g0c + c;
  ColCoords<DD> cc;
  cc.load(xc, c, D);
  const double acv = ac[c];
  constexpr int ND = DD > 0 ? DD : GPSAT_DMAX;
  double pd[ND], pk = 0.0, pn = 0.0;
#pragma unroll
  for (int dd = 0; dd < ND; ++dd) pd[dd] = 0.0;
  if (g0r + NB <= n && g0r > g0c) {
    // interior off-diagonal tile: every element counts twice, no bounds, no diagonal.  Batches of 4 rows keep 4
    // independent exp / sqrt chains in flight per thread.
    constexpr int UB = 4, RS = GPSAT_THREADS / NB;
    // the Q values of the next batch are fetched before the current batch is evaluated: with the tile in shared memory
    // that costs four registers, with the tile in the L2-resident scratch of lauum_grad_i8q_kernel it hides the latency
    double qn[UB];
#pragma unroll
    for (int u = 0; u < UB; ++u) qn[u] = Cs[((tid >> 7) + RS * u) * LDC + c];
#pragma unroll 1
    for (int r0 = tid >> 7; r0 < NB; r0 += RS * UB) {
      double wq[UB], sd[UB][ND], ku[UB], gu[UB], qv[UB];
#pragma unroll
      for (int u = 0; u < UB; ++u) qv[u] = qn[u];
      if (r0 + RS * UB < NB) {
#pragma unroll
        for (int u = 0; u < UB; ++u) qn[u] = Cs[(r0 + RS * UB + RS * u) * LDC + c];
      }
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const int r = r0 + RS * u;
        const double aa = ar[r] * acv;
        wq[u] = 2.0 * (aa - (qv[u] - aac * aa));
        double r2 = 0.0;
#pragma unroll
        for (int dd = 0; dd < ND; ++dd) {
          double t = (DD > 0 || dd < D) ? xr[dd * NB + r] - cc.x[dd] : 0.0;
          sd[u][dd] = t * t;
          r2 += sd[u][dd];
        }
        kunit_gunit_of_r2<KIND>(r2, ku[u], gu[u]);
      }
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const double wg = wq[u] * gu[u];
#pragma unroll
        for (int dd = 0; dd < ND; ++dd) pd[dd] = fma(wg, sd[u][dd], pd[dd]);
        pk = fma(wq[u], ku[u], pk);
      }
    }
  } else if (gj < n) {
    // diagonal tiles and the tiles of the last block row: same batches, dead elements (rows >= n, columns above
    // the diagonal) enter with weight zero; batches that are dead for the whole warp are skipped
    constexpr int UB = 4, RS = GPSAT_THREADS / NB;
    const int wc0 = g0c + (c & ~31);                 // first column of this warp
#pragma unroll 1
    for (int r0 = tid >> 7; r0 < NB; r0 += RS * UB) {
      if (g0r + r0 >= n) break;
      if (g0r + r0 + RS * (UB - 1) < wc0) continue;
      double wq[UB], qd[UB], sd[UB][ND], ku[UB], gu[UB], qv[UB];
#pragma unroll
      for (int u = 0; u < UB; ++u) qv[u] = Cs[(r0 + RS * u) * LDC + c];       // four loads in flight before the first use
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const int r = r0 + RS * u, gi = g0r + r;
        const bool live = gi < n && gj <= gi;
        const double aa = ar[r] * acv;
        // aac = 1: the tile holds Ks^-1 + alpha alpha^T (column n of U is -alpha); aac = 0: Ks^-1 alone
        const double q = aa - (qv[u] - aac * aa);
        wq[u] = live ? ((gi == gj) ? q : 2.0 * q) : 0.0;
        qd[u] = (live && gi == gj) ? q : 0.0;
        double r2 = 0.0;
#pragma unroll
        for (int dd = 0; dd < ND; ++dd) {
          double t = (DD > 0 || dd < D) ? xr[dd * NB + r] - cc.x[dd] : 0.0;
          sd[u][dd] = t * t;
          r2 += sd[u][dd];
        }
        kunit_gunit_of_r2<KIND>(r2, ku[u], gu[u]);
      }
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const double wg = wq[u] * gu[u];
#pragma unroll
        for (int dd = 0; dd < ND; ++dd) pd[dd] = fma(wg, sd[u][dd], pd[dd]);
        pk = fma(wq[u], ku[u], pk);
        pn += qd[u];
      }
    }
  }
#pragma unroll
  for (int dd = 0; dd < GPSAT_DMAX; ++dd) part[dd] = dd < ND ? pd[dd < ND ? dd : 0] : 0.0;
  part[GPSAT_DMAX] = pk;
  part[GPSAT_DMAX + 1] = pn;
}

__device__ __noinline__ void grad_tile(const SlotDesc& d, int g0r, int g0c, const double* xr, const double* xc,
                                       const double* ar, const double* ac, const double* Cs, int tid, double* part,
                                       double aac = 1.0) {
  const bool d3 = d.D == 3;
  switch (d.kind) {
    case GPSAT_MATERN12:
      if (d3) grad_tile_t<GPSAT_MATERN12, 3>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      else grad_tile_t<GPSAT_MATERN12, 0>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      break;
    case GPSAT_MATERN32:
      if (d3) grad_tile_t<GPSAT_MATERN32, 3>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      else grad_tile_t<GPSAT_MATERN32, 0>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      break;
    case GPSAT_MATERN52:
      if (d3) grad_tile_t<GPSAT_MATERN52, 3>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      else grad_tile_t<GPSAT_MATERN52, 0>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      break;
    default:
      if (d3) grad_tile_t<GPSAT_RBF, 3>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      else grad_tile_t<GPSAT_RBF, 0>(d, g0r, g0c, xr, xc, ar, ac, Cs, tid, part, aac);
      break;
  }
}

// skip_planes != 0: slots whose evaluation runs on the int8 planes (d.Uq8 set) belong to lauum_grad_i8_kernel.
__global__ void __launch_bounds__(GPSAT_THREADS, 1) lauum_grad_kernel(const SlotDesc* __restrict__ slots, int skip_planes) {
  const SlotDesc& d = slots[blockIdx.y];
  if (d.expert < 0 || (skip_planes && d.Uq8 != nullptr)) return;
  const int T = d.T;
  const int t = blockIdx.x;
  if (t >= T * (T + 1) / 2) return;
  int i = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
  while (i * (i + 1) / 2 > t) --i;
  while ((i + 1) * (i + 2) / 2 <= t) ++i;
  const int j = t - i * (i + 1) / 2;

  extern __shared__ double smem[];
  double* Cs = smem;                     // pipeline stages, then the Q tile
  double* xr = smem + SM_TILE_D;
  double* xc = xr + NB * GPSAT_DMAX;
  double* ar = xc + NB * GPSAT_DMAX;     // alpha of the row block
  double* ac = ar + NB;                  // alpha of the column block
  double* red = ac + NB;                 // [8 warps][8]
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int n = d.n, ld = d.ld;

  const int kend = ((n + 1 + KC - 1) / KC) * KC;     // includes column n (the -alpha column)
  gemm_nt_prologue(d.A + (size_t)i * NB * ld, d.A + (size_t)j * NB * ld, ld, i * NB, kend, smem, tid);
  load_block_coords(d, i * NB, xr, ar, tid);      // ar / ac overwritten with alpha below
  load_block_coords(d, j * NB, xc, ac, tid);
  __syncthreads();
  for (int r = tid; r < NB; r += GPSAT_THREADS) {
    int gi = i * NB + r, gj = j * NB + r;
    ar[r] = (gi < n) ? -d.A[(size_t)gi * ld + n] : 0.0;
    ac[r] = (gj < n) ? -d.A[(size_t)gj * ld + n] : 0.0;
  }
  {
    Acc acc;
    acc_zero(acc);
    gemm_nt_mainloop(acc, d.A + (size_t)i * NB * ld, d.A + (size_t)j * NB * ld, ld, i * NB, kend, smem, tid, true);
    __syncthreads();
    acc_to_smem(acc, Cs, tid, 1.0);
  }
  __syncthreads();

  double part[GPSAT_DMAX + 2];
  grad_tile(d, i * NB, j * NB, xr, xc, ar, ac, Cs, tid, part);
  // deterministic block reduction: warp shuffle tree, then warp 0 adds the 8 warp sums in order
#pragma unroll
  for (int p = 0; p < GPSAT_DMAX + 2; ++p) {
    double v = part[p];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp * 8 + p] = v;
  }
  __syncthreads();
  if (tid < GPSAT_DMAX + 2) {
    double v = 0.0;
    for (int w8 = 0; w8 < 8; ++w8) v += red[w8 * 8 + tid];
    d.gpart[(size_t)t * (GPSAT_DMAX + 2) + tid] = v;
  }
}

// predict_reduce_kernel: one warp per prediction row s:  f* = <A_s, a>, f*_var = var - |A_s|^2,
// y_var = f*_var + s2, with A_s = row n+1+s of the factored matrix (= (L^-1 k_s)^T) and
// a = row n (= (L^-1 y)^T).   Reference: GPR.predict_f / predict_y called at
// GPSat/models/gpflow_models.py:229-230.
__global__ void predict_reduce_kernel(const SlotDesc* __restrict__ slots, const int64_t* __restrict__ off_m,
                                      double* __restrict__ fmean, double* __restrict__ fvar,
                                      double* __restrict__ yvar) {
  const SlotDesc& d = slots[blockIdx.y];
  if (d.expert < 0) return;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= d.m) return;
  const int n = d.n, ld = d.ld;
  const int g = n + 1 + warp;
  const double* As = d.A + (size_t)g * ld;
  const double* a = d.A + (size_t)n * ld;
  // columns of the last factored block: rows of block row T-1 (the y row, and the first few
  // prediction rows) keep their L entries in the diagonal side buffer, not in the slot matrix
  const int c0 = (d.T - 1) * NB;
  const double* Ldt = d.Ld + (size_t)(d.T - 1) * NB * NB;
  const double* a_last = Ldt + (size_t)(n - c0) * NB;
  const double* As_last = (g < d.T * NB) ? Ldt + (size_t)(g - c0) * NB : nullptr;
  double s1 = 0.0, s2 = 0.0;
  for (int k = lane; k < c0; k += 32) { double v = As[k]; s1 += v * a[k]; s2 += v * v; }
  for (int k = c0 + lane; k < n; k += 32) {
    double v = As_last ? As_last[k - c0] : As[k];
    s1 += v * a_last[k - c0];
    s2 += v * v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (lane == 0) {
    int64_t o = off_m[d.expert] + warp;
    double fv = d.var - s2;
    fmean[o] = s1;
    fvar[o] = fv;
    yvar[o] = fv + d.noise;
  }
}

// predict_cov_kernel: full posterior covariance of the prediction rows of one slot,
//   cov[s][t] = k(xs_s, xs_t) - <A_s, A_t>,   A_s = row n+1+s of the factored matrix (columns < n),
// written as a dense m x m block at fcov + off_c[expert] (GPR.predict_f(full_cov=True), called at
// GPSat/models/gpflow_models.py:230).  64 x 64 output tile per CTA, 4 x 4 per thread, plain DFMA:
// this is an occasional diagnostic output, not part of the timed hot path.
__device__ __forceinline__ double factor_entry(const SlotDesc& d, int g, int k) {
  const int c0 = (d.T - 1) * NB;
  if (g < d.T * NB && k >= c0) return d.Ld[(size_t)(d.T - 1) * NB * NB + (size_t)(g - c0) * NB + (k - c0)];
  return d.A[(size_t)g * d.ld + k];
}

__global__ void __launch_bounds__(256) predict_cov_kernel(const SlotDesc* __restrict__ slots,
                                                          const int64_t* __restrict__ off_c, double* __restrict__ fcov) {
  const SlotDesc& d = slots[blockIdx.z];
  if (d.expert < 0) return;
  const int m = d.m, n = d.n, D = d.D;
  const int s0 = blockIdx.y * 64, t0 = blockIdx.x * 64;
  if (s0 >= m || t0 >= m || t0 > s0) return;          // lower tiles; mirrored on store
  __shared__ double As[64][17], Bs[64][17];
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  for (int k0 = 0; k0 < n; k0 += 16) {
    for (int idx = tid; idx < 64 * 16; idx += 256) {
      const int r = idx >> 4, k = idx & 15;
      As[r][k] = (s0 + r < m && k0 + k < n) ? factor_entry(d, n + 1 + s0 + r, k0 + k) : 0.0;
      Bs[r][k] = (t0 + r < m && k0 + k < n) ? factor_entry(d, n + 1 + t0 + r, k0 + k) : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      double av[4], bv[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) av[a] = As[ty * 4 + a][k];
#pragma unroll
      for (int b = 0; b < 4; ++b) bv[b] = Bs[tx * 4 + b][k];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
    }
    __syncthreads();
  }
  double* out = fcov + off_c[d.expert];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int s = s0 + ty * 4 + a, t = t0 + tx * 4 + b;
      if (s >= m || t >= m || t > s) continue;
      double r2 = 0.0;
      for (int dd = 0; dd < D; ++dd) {
        double df = (d.Xs[(size_t)s * D + dd] - d.Xs[(size_t)t * D + dd]) * d.inv_ls[dd];
        r2 += df * df;
      }
      const double v = d.var * kunit_dyn(d.kind, r2) - acc[a][b];
      out[(size_t)s * m + t] = v;
      out[(size_t)t * m + s] = v;
    }
}

}  // namespace gpsat