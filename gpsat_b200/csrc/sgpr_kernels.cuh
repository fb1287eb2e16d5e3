// Sparse GPR (Titsias) on the same tile machinery: the reference's GPflowSGPRModel
// (GPSat/models/gpflow_models.py:666-901 -> gpflow.models.SGPR.elbo / predict_f, Z fixed :896-897).
//
// Forward (ELBO):
//   Kuu = k(Z,Z) + jitter I = L L^T         sweep of the "K slot" (M inducing points as observations)
//   W   = L^-1, U = W^T                      transpose of the trtri tiles
//   Ct  = Kuf^T  [n][Mp]                     kuf_build_kernel (column M carries y, see below)
//   A'  = W Ct^T / sigma                     GEMM; row M of A' comes out as y^T / sigma; stored as
//                                            Aq [Mp][npad] and At [npad][Mp]
//   B'  = A' A'^T (+ I)   [Mp][Mp]           split-K SYRK + ordered reduce:
//                                            B'[u][v] = B, B'[M][u] = (A y)_u / sigma, B'[M][M] = y'y / s2
//   chol(B') with row M riding along         sweep of the "B slot" on the precomputed matrix:
//                                            log|LB|, c = LB^-1 A y / sigma, quad = c'c
// Prediction: T1 = Kus^T W^T, T2 = T1 WB^T, mean = T2 c, var = k** + |T2_s|^2 - |T1_s|^2
//
// Gradient w.r.t. theta (analytic form of what TensorFlow autodiff gives for -elbo; derivation and
// the NumPy statement in oracle/sgpr.py:elbo_and_grad).  With WB = LB^-1, UB = WB^T, gamma = UB c,
// r = y - sigma A^T gamma:
//   G_C^T = [ (At - At WB^T UB^T) / sigma + r gamma^T / s2 ] U^T        [n][M]
//   G_K   = 1/2 U [ 2 I - B - B^-1 - gamma gamma^T ] U^T,   B^-1 = UB UB^T
//   d elbo / d l_d = var / l_d * ( sum G_C gunit sd_d + sum G_K gunit sd_d )
//   d elbo / d var = sum G_C kunit + sum G_K kunit - n / (2 s2)
//   d elbo / d s2  = -n / (2 s2) + [ r'r + n var - s2 (tr B - 2 M + tr B^-1) ] / (2 s2^2)
// G_C and G_K are never written: the GEMM epilogues contract each 128 x 128 tile with the kernel
// derivatives and keep [D + 1] partial sums per tile.
#pragma once
#include "tiles.cuh"

namespace gpsat {

constexpr double SGPR_JITTER = 1e-6;     // gpflow.config.default_jitter()
constexpr int SGPR_RES_ROWS = 64;        // rows of the residual kernel per CTA
constexpr int SGPR_KCHUNK = 2048;        // k range of one split of the SYRK: a property of the expert alone,
                                         // so results do not depend on what else is in the batch

struct SgprDesc {
  int expert;            // -1 = idle
  int n, m, M, D, kind;
  int npad, mpad, Mp;    // n, m, M + 1 rounded up to multiples of 128
  const double* X; const double* y; const double* Z; const double* Xs;
  double inv_ls[GPSAT_DMAX];
  double var, noise;
  double* Ct;            // [npad][Mp]   Kuf^T | y        (dead after A'; reused as Ht)
  double* Aq;            // [Mp][npad]   A'               (dead after the SYRK; reused as Vt)
  double* At;            // [npad][Mp]   A'^T             (gradient only)
  double* W;             // [Mp][Mp]     L^-1
  double* U;             // [Mp][Mp]     L^-T             (gradient only)
  double* WB;            // [Mp][Mp]     LB^-1
  double* UB;            // [Mp][Mp]     LB^-T            (gradient only)
  double* Bm;            // [Mp][Mp]     B' (lower tiles)
  double* Binv;          // [Mp][Mp]     B^-1 (lower tiles)
  double* Q;             // [Mp][Mp]     2I - B - B^-1 - gamma gamma^T (full, zero outside [0,M)^2)
  double* TT;            // [Mp][Mp]     U Q
  double* part;          // [nsplit][Mp][Mp]
  double* gam;           // [Mp]
  double* bid;           // [Mp]         diag(B^-1)
  double* res;           // [npad]       r
  double* rrp;           // [npad / 64]  partial sums of r'r
  double* gpc;           // [npad/128 * Mp/128][8] partial gradient sums of the Kuf term
  double* gpk;           // [(Mp/128)^2][8]        partial gradient sums of the Kuu term
  double* sums;          // [12] block-reduced scalars of one evaluation (sgpr_sums_kernel)
  double* zpc;           // [Mp/128 * npad/128][2][128][DMAX] per-row partial sums of d elbo / d Z, Kuf term (or nullptr)
  double* zpk;           // [(Mp/128)^2][2][128][DMAX]        the Kuu term (trainable inducing points only)
  double* Cts;           // [mpad][Mp]
  double* T1;            // [mpad][Mp]
  double* T2;            // [mpad][Mp]
};

// Offsets (in doubles) of one slot's workspace; the same for every slot of a call.
struct SgprLay {
  size_t ct, aq, at, w, u, wb, ub, bm, binv, q, tt, part, gam, bid, res, rrp, gpc, gpk, sums, cts, t1, t2, ak, ab,
      side_k, side_b, total, tile_d, zpc, zpk;
  int with_z;
  int Tmax, lgd;         // tiles per side of the M x M slots; aligned length of logdet[]
};

struct SgprIn {
  const int64_t* off_z;  // device copy
  const double* Z;
  const double* zero;    // [Mp] zeros: the "observations" of the K slot
};

__device__ __forceinline__ void sgpr_fill_slot(SlotDesc& s, int expert, int M, int kind, int D, double* A,
                                               double* side, const SgprLay& L) {
  s.expert = expert;
  s.n = M; s.m = 0;
  s.T = s.TR = (M + 1 + NB - 1) / NB;
  s.ld = s.T * NB;
  s.kind = kind; s.D = D;
  s.A = A;
  s.Wd = side; s.Ld = side + L.tile_d; s.logdet = side + 2 * L.tile_d;
  s.quad = s.logdet + L.lgd;
  s.gpart = nullptr;
  s.notpd = reinterpret_cast<int*>(s.quad + 2);
  s.pre = nullptr; s.pre_s = nullptr; s.ldpre = 0;
  s.Uq = nullptr; s.uex = nullptr; s.Lq = nullptr; s.Uq8 = nullptr; s.ynorm2 = 0.0; s.i8_low = 0; s.skip_fp64 = 0;
  s.Xs = nullptr;
}

// Bind expert e to a slot: the SGPR descriptor, the K slot (Kuu + jitter I, y = 0) and the B slot.
__device__ __forceinline__ void sgpr_set_expert(SgprDesc& d, SlotDesc& k, SlotDesc& b, double* base, const SgprLay& L,
                                                const BatchDev& bd, const SgprIn& in, int e, int with_pred) {
  d.expert = e;
  d.n = (int)(bd.off_n[e + 1] - bd.off_n[e]);
  d.m = (with_pred && bd.off_m) ? (int)(bd.off_m[e + 1] - bd.off_m[e]) : 0;
  d.M = (int)(in.off_z[e + 1] - in.off_z[e]);
  d.D = bd.D; d.kind = bd.kind;
  d.npad = (d.n + NB - 1) / NB * NB;
  d.mpad = (max(d.m, 1) + NB - 1) / NB * NB;
  d.Mp = (d.M + 1 + NB - 1) / NB * NB;
  d.X = bd.X + bd.off_n[e] * bd.D; d.y = bd.y + bd.off_n[e];
  d.Z = in.Z + in.off_z[e] * bd.D;
  d.Xs = d.m > 0 ? bd.Xs + bd.off_m[e] * bd.D : nullptr;
  d.Ct = base + L.ct; d.Aq = base + L.aq; d.At = base + L.at; d.W = base + L.w; d.U = base + L.u;
  d.WB = base + L.wb; d.UB = base + L.ub; d.Bm = base + L.bm; d.Binv = base + L.binv; d.Q = base + L.q;
  d.TT = base + L.tt; d.part = base + L.part; d.gam = base + L.gam; d.bid = base + L.bid; d.res = base + L.res;
  d.rrp = base + L.rrp; d.gpc = base + L.gpc; d.gpk = base + L.gpk; d.sums = base + L.sums;
  d.zpc = L.with_z ? base + L.zpc : nullptr; d.zpk = L.with_z ? base + L.zpk : nullptr;
  d.Cts = base + L.cts; d.T1 = base + L.t1;
  d.T2 = base + L.t2;
  sgpr_fill_slot(k, e, d.M, d.kind, d.D, base + L.ak, base + L.side_k, L);
  k.X = d.Z; k.y = in.zero;
  sgpr_fill_slot(b, e, d.M, GPSAT_PRECOMPUTED, 0, base + L.ab, base + L.side_b, L);
  b.X = nullptr; b.y = d.Bm + (size_t)d.M * d.Mp;
  b.var = 0.0; b.noise = 0.0;
  b.pre = d.Bm; b.ldpre = d.Mp;
}

__device__ __forceinline__ void sgpr_set_theta(SgprDesc& d, SlotDesc& k, SlotDesc& b, const double* theta) {
  for (int q = 0; q < d.D; ++q) { d.inv_ls[q] = 1.0 / theta[q]; k.inv_ls[q] = d.inv_ls[q]; }
  d.var = theta[d.D]; d.noise = theta[d.D + 1];
  k.var = d.var; k.noise = SGPR_JITTER;
  *k.quad = 0.0; *k.notpd = 0;
  *b.quad = 0.0; *b.notpd = 0;
}

// ---- Ct[i][k] = k(x_i, z_k); column M = y_i (train) or 0; zero padding elsewhere ----------------
template <int KIND, int DD>
__device__ __forceinline__ void kuf_tile(const SgprDesc& d, const double* P, int np, const double* yv, double* out,
                                         int i0, int k0, double* xi, double* zk, int tid) {
  const int D = d.D, M = d.M;
  for (int idx = tid; idx < NB * D; idx += GPSAT_THREADS) {
    const int row = idx / D, dd = idx - row * D;
    xi[dd * NB + row] = (i0 + row < np) ? P[(size_t)(i0 + row) * D + dd] * d.inv_ls[dd] : 0.0;
    zk[dd * NB + row] = (k0 + row < M) ? d.Z[(size_t)(k0 + row) * D + dd] * d.inv_ls[dd] : 0.0;
  }
  __syncthreads();
  const int c = tid & (NB - 1), k = k0 + c;
  ColCoords<DD> cc;
  cc.load(zk, c, D);
  const double var = d.var;
  // batches of 8 rows: 8 independent exp / sqrt chains per thread, values outside the data (coordinates read as
  // zeros) are formed and then replaced, so the batch is branch-free (same scheme as add_aug_tile_t in tiles.cuh)
  constexpr int UB = 8, RS = GPSAT_THREADS / NB;
#pragma unroll 1
  for (int r0 = tid >> 7; r0 < NB; r0 += RS * UB) {
    double v[UB];
#pragma unroll
    for (int u = 0; u < UB; ++u) v[u] = var * kunit_of_r2<KIND>(cc.r2(xi, r0 + RS * u, D));
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const int i = i0 + r0 + RS * u;
      double w = v[u];
      if (k > M || i >= np) w = 0.0;
      else if (k == M) w = yv ? yv[i] : 0.0;
      out[(size_t)i * d.Mp + k] = w;
    }
  }
}

// which = 0: training inputs X -> Ct (with y in column M); which = 1: prediction inputs Xs -> Cts
__global__ void __launch_bounds__(GPSAT_THREADS) kuf_build_kernel(const SgprDesc* __restrict__ descs, int which) {
  const SgprDesc& d = descs[blockIdx.z];
  if (d.expert < 0) return;
  const int np = which ? d.m : d.n, npad = which ? d.mpad : d.npad;
  const int i0 = blockIdx.x * NB, k0 = blockIdx.y * NB;
  if (i0 >= npad || k0 >= d.Mp) return;
  __shared__ double xi[NB * GPSAT_DMAX], zk[NB * GPSAT_DMAX];
  const double* P = which ? d.Xs : d.X;
  const double* yv = which ? nullptr : d.y;
  double* out = which ? d.Cts : d.Ct;
#define KUF_CASE(K) \
  if (d.D == 3) kuf_tile<K, 3>(d, P, np, yv, out, i0, k0, xi, zk, threadIdx.x); \
  else kuf_tile<K, 0>(d, P, np, yv, out, i0, k0, xi, zk, threadIdx.x); \
  break;
  switch (d.kind) {
    case GPSAT_MATERN12: KUF_CASE(GPSAT_MATERN12)
    case GPSAT_MATERN32: KUF_CASE(GPSAT_MATERN32)
    case GPSAT_MATERN52: KUF_CASE(GPSAT_MATERN52)
    default: KUF_CASE(GPSAT_RBF)
  }
#undef KUF_CASE
}

// ---- W = L^-1 (dense lower, row-major [Mp][Mp]) and optionally U = W^T from a factored slot ------
// Off-diagonal 128-blocks come from the strict upper block triangle of the slot matrix (which holds
// U), the diagonal blocks from the W_tt side buffer.  which = 0: descs[].W / U from the K slots,
// 1: descs[].WB / UB from the B slots.  One CTA owns the 32 x 32 tile W[u0.., k0..] and its mirror
// U[k0.., u0..].  U is the transpose of the M x M factor inverse only: its columns >= M (the rows of
// W that belong to the y row and the padding) are zero, so that products over U / UB see exactly
// L^-T / LB^-T.
__global__ void __launch_bounds__(256) transpose_w_kernel(const SgprDesc* __restrict__ descs,
                                                          const SlotDesc* __restrict__ slots, int which, int with_u) {
  const SgprDesc& d = descs[blockIdx.z];
  if (d.expert < 0) return;
  const SlotDesc& s = slots[blockIdx.z];
  const int Mp = d.Mp;
  const int u0 = blockIdx.y * 32, k0 = blockIdx.x * 32;
  if (u0 >= Mp || k0 >= Mp) return;
  double* W = which ? d.WB : d.W;
  double* U = with_u ? (which ? d.UB : d.U) : nullptr;
  __shared__ double t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
  const int ub = u0 / NB, kb = k0 / NB;
  const bool ucol = u0 + tx < d.M;
  if (kb > ub) {
    for (int r = ty; r < 32; r += 8) {
      W[(size_t)(u0 + r) * Mp + k0 + tx] = 0.0;
      if (U) U[(size_t)(k0 + r) * Mp + u0 + tx] = 0.0;
    }
    return;
  }
  if (kb == ub) {
    const double* Wd = s.Wd + (size_t)ub * NB * NB;
    for (int r = ty; r < 32; r += 8) {
      const double v = Wd[(size_t)(u0 - ub * NB + r) * NB + (k0 - kb * NB + tx)];
      W[(size_t)(u0 + r) * Mp + k0 + tx] = v;
      t[r][tx] = v;
    }
    if (U) {
      __syncthreads();
      for (int r = ty; r < 32; r += 8) U[(size_t)(k0 + r) * Mp + u0 + tx] = ucol ? t[tx][r] : 0.0;
    }
    return;
  }
  for (int r = ty; r < 32; r += 8) {
    const double v = s.A[(size_t)(k0 + r) * s.ld + u0 + tx];    // U[k][u]
    t[r][tx] = v;
    if (U) U[(size_t)(k0 + r) * Mp + u0 + tx] = ucol ? v : 0.0;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) W[(size_t)(u0 + r) * Mp + k0 + tx] = t[tx][r];
}

// ---- contraction of one 128 x 128 tile of G (in shared memory) with the kernel derivatives -------
//   part[d] = sum G_rc gunit(r2_rc) sd_d,   part[DMAX] = sum G_rc kunit(r2_rc)
// rows r < nr carry the scaled coordinates xr, columns c < nc the scaled coordinates xc.
template <int KIND>
__device__ __forceinline__ void cross_grad_tile_t(int D, const double* xr, const double* xc, int nr, int nc,
                                                  const double* Cs, int tid, double* part) {
  const int c = tid & (NB - 1);
  ColCoords<0> cc;
  cc.load(xc, c, D);
  double pd[GPSAT_DMAX], pk = 0.0;
#pragma unroll
  for (int dd = 0; dd < GPSAT_DMAX; ++dd) pd[dd] = 0.0;
  if (c < nc) {
#pragma unroll 2
    for (int r = tid >> 7; r < nr; r += GPSAT_THREADS / NB) {
      const double gv = Cs[r * LDC + c];
      double sd[GPSAT_DMAX], r2 = 0.0;
#pragma unroll
      for (int dd = 0; dd < GPSAT_DMAX; ++dd) {
        const double t = dd < D ? xr[dd * NB + r] - cc.x[dd] : 0.0;
        sd[dd] = t * t;
        r2 += sd[dd];
      }
      double ku, gu;
      kunit_gunit_of_r2<KIND>(r2, ku, gu);
      const double wg = gv * gu;
#pragma unroll
      for (int dd = 0; dd < GPSAT_DMAX; ++dd) pd[dd] = fma(wg, sd[dd], pd[dd]);
      pk = fma(gv, ku, pk);
    }
  }
#pragma unroll
  for (int dd = 0; dd < GPSAT_DMAX; ++dd) part[dd] = pd[dd];
  part[GPSAT_DMAX] = pk;
  part[GPSAT_DMAX + 1] = 0.0;
}

__device__ __noinline__ void cross_grad_tile(int kind, int D, const double* xr, const double* xc, int nr, int nc,
                                             const double* Cs, int tid, double* part) {
  switch (kind) {
    case GPSAT_MATERN12: cross_grad_tile_t<GPSAT_MATERN12>(D, xr, xc, nr, nc, Cs, tid, part); break;
    case GPSAT_MATERN32: cross_grad_tile_t<GPSAT_MATERN32>(D, xr, xc, nr, nc, Cs, tid, part); break;
    case GPSAT_MATERN52: cross_grad_tile_t<GPSAT_MATERN52>(D, xr, xc, nr, nc, Cs, tid, part); break;
    default: cross_grad_tile_t<GPSAT_RBF>(D, xr, xc, nr, nc, Cs, tid, part); break;
  }
}

// ---- the same tile contracted with d k / d z of its ROW points (trainable inducing points) -----------------------
//   zp[half][r][d] = sum_{c in half} G_rc gunit(r2_rc) (xc_d[c] - xr_d[r]),   d k(z_r, p_c) / d z_rd = var gunit (xc_d - xr_d) / l_d
// thread = (row r, column parity): 64 columns each, the two parities are summed by sgpr_gradz_kernel in a fixed order.
template <int KIND>
__device__ __forceinline__ void cross_gradz_tile_t(int D, const double* xr, const double* xc, int nr, int nc,
                                                   const double* Cs, int tid, double* zp) {
  const int r = tid & (NB - 1), half = tid >> 7;
  double x0[GPSAT_DMAX], pz[GPSAT_DMAX];
#pragma unroll
  for (int dd = 0; dd < GPSAT_DMAX; ++dd) { x0[dd] = dd < D ? xr[dd * NB + r] : 0.0; pz[dd] = 0.0; }
  if (r < nr) {
#pragma unroll 2
    for (int c = half; c < nc; c += GPSAT_THREADS / NB) {
      const double gv = Cs[r * LDC + c];
      double df[GPSAT_DMAX], r2 = 0.0;
#pragma unroll
      for (int dd = 0; dd < GPSAT_DMAX; ++dd) {
        df[dd] = dd < D ? xc[dd * NB + c] - x0[dd] : 0.0;
        r2 = fma(df[dd], df[dd], r2);
      }
      double ku, gu;
      kunit_gunit_of_r2<KIND>(r2, ku, gu);
      const double wg = gv * gu;
#pragma unroll
      for (int dd = 0; dd < GPSAT_DMAX; ++dd) pz[dd] = fma(wg, df[dd], pz[dd]);
    }
  }
#pragma unroll
  for (int dd = 0; dd < GPSAT_DMAX; ++dd) zp[((size_t)half * NB + r) * GPSAT_DMAX + dd] = pz[dd];
}

__device__ __noinline__ void cross_gradz_tile(int kind, int D, const double* xr, const double* xc, int nr, int nc,
                                              const double* Cs, int tid, double* zp) {
  switch (kind) {
    case GPSAT_MATERN12: cross_gradz_tile_t<GPSAT_MATERN12>(D, xr, xc, nr, nc, Cs, tid, zp); break;
    case GPSAT_MATERN32: cross_gradz_tile_t<GPSAT_MATERN32>(D, xr, xc, nr, nc, Cs, tid, zp); break;
    case GPSAT_MATERN52: cross_gradz_tile_t<GPSAT_MATERN52>(D, xr, xc, nr, nc, Cs, tid, zp); break;
    default: cross_gradz_tile_t<GPSAT_RBF>(D, xr, xc, nr, nc, Cs, tid, zp); break;
  }
}

// scaled coordinates of rows [g0, g0 + 128) of a point set with np valid rows -> xs[dim][row]
__device__ __forceinline__ void load_scaled_coords(const double* P, int np, int D, const double* inv_ls, int g0,
                                                   double* xs, int tid) {
  for (int idx = tid; idx < NB * D; idx += GPSAT_THREADS) {
    const int row = idx / D, dd = idx - row * D;
    xs[dd * NB + row] = (g0 + row < np) ? P[(size_t)(g0 + row) * D + dd] * inv_ls[dd] : 0.0;
  }
}

// ---- batched NT GEMM on 128 x 128 tiles:  out = Aop Bop^T  (both operands k-contiguous) ----------
enum SgprGemmMode : int {
  SGPR_GEMM_A = 0,     // Aq[Mp][npad] (and At) = W Ct^T / sigma     (W lower triangular: k <= row tile)
  SGPR_GEMM_SYRK = 1,  // part[split] = Aq Aq^T over one k range, lower tiles
  SGPR_GEMM_T1 = 2,    // T1[mpad][Mp] = Cts W^T                      (k <= column tile)
  SGPR_GEMM_T2 = 3,    // T2[mpad][Mp] = T1 WB^T                      (k <= column tile)
  SGPR_GEMM_V = 4,     // V = WB At^T, stored as Vt[npad][Mp]         (k <= row tile, last block triangular)
  SGPR_GEMM_H = 5,     // UB Vt^T -> Ht[npad][Mp] = (At - Vt UB^T) / sigma + r gamma^T / s2   (k >= row tile)
  SGPR_GEMM_GC = 6,    // G_C tile = U Ht^T, contracted with dKuf/dtheta -> gpc                (k >= row tile)
  SGPR_GEMM_BINV = 7,  // Binv = UB UB^T, lower tiles                 (k >= row tile)
  SGPR_GEMM_TT = 8,    // TT = U Q                                    (k >= row tile)
  SGPR_GEMM_GK = 9     // G_K tile = 1/2 U TT^T, contracted with dKuu/dtheta -> gpk   (k >= row tile)
};
constexpr int SGPR_GEMM_SMEM_BYTES = (SM_TILE_D + 2 * NB * GPSAT_DMAX + 64) * 8;

__device__ __forceinline__ void tri_tile(int tile, int& rt, int& ct) {
  rt = (int)((sqrtf(8.0f * tile + 1.0f) - 1.0f) * 0.5f);
  while (rt * (rt + 1) / 2 > tile) --rt;
  while ((rt + 1) * (rt + 2) / 2 <= tile) ++rt;
  ct = tile - rt * (rt + 1) / 2;
}

__global__ void __launch_bounds__(GPSAT_THREADS, 1)
sgpr_gemm_kernel(const SgprDesc* __restrict__ descs, int mode, int nsplit, int with_t) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int tile = blockIdx.x / nsplit, split = blockIdx.x - tile * nsplit;
  const int Mp = d.Mp, MT = Mp / NB;
  const double *Aop, *Bop;
  double* out = nullptr;
  int ldg, ldo = Mp, rt, ct, kbeg = 0, kend = Mp;
  bool tri_first = false, tri_last = false;
  switch (mode) {
    case SGPR_GEMM_A: {
      const int nct = d.npad / NB;
      if (tile >= MT * nct) return;
      rt = tile / nct; ct = tile - rt * nct;
      Aop = d.W; Bop = d.Ct; ldg = Mp; out = d.Aq; ldo = d.npad;
      kend = (rt + 1) * NB; tri_last = true;
      break;
    }
    case SGPR_GEMM_SYRK: {
      if (tile >= MT * (MT + 1) / 2) return;
      tri_tile(tile, rt, ct);
      Aop = d.Aq; Bop = d.Aq; ldg = d.npad; out = d.part + (size_t)split * Mp * Mp;
      if (split * SGPR_KCHUNK >= d.npad) return;
      kbeg = split * SGPR_KCHUNK;
      kend = min(d.npad, kbeg + SGPR_KCHUNK);
      break;
    }
    case SGPR_GEMM_V: case SGPR_GEMM_H: case SGPR_GEMM_GC: {
      // M-side operand (triangular) as A, data-side operand as B: the tile is the transpose of the
      // [n][M] quantity it stands for, and the skipped half of the triangular 128-block is skipped
      // by row fragments (balanced over the four schedulers)
      const int nct = d.npad / NB;
      if (tile >= MT * nct) return;
      rt = tile / nct; ct = tile - rt * nct;
      ldg = Mp;
      if (mode == SGPR_GEMM_V) { Aop = d.WB; Bop = d.At; kend = (rt + 1) * NB; tri_last = true; }
      else if (mode == SGPR_GEMM_H) { Aop = d.UB; Bop = d.Aq; kbeg = rt * NB; tri_first = true; }
      else { Aop = d.U; Bop = d.Ct; kbeg = rt * NB; tri_first = true; }
      break;
    }
    case SGPR_GEMM_T1: case SGPR_GEMM_T2: {
      const int nrt = d.mpad / NB;
      if (tile >= nrt * MT) return;
      rt = tile / MT; ct = tile - rt * MT;
      ldg = Mp;
      if (mode == SGPR_GEMM_T1) { Aop = d.Cts; Bop = d.W; out = d.T1; kend = (ct + 1) * NB; }
      else { Aop = d.T1; Bop = d.WB; out = d.T2; kend = (ct + 1) * NB; }
      break;
    }
    case SGPR_GEMM_BINV: {
      if (tile >= MT * (MT + 1) / 2) return;
      tri_tile(tile, rt, ct);
      Aop = d.UB; Bop = d.UB; ldg = Mp; out = d.Binv; kbeg = rt * NB; tri_first = true;
      break;
    }
    default: {
      if (tile >= MT * MT) return;
      rt = tile / MT; ct = tile - rt * MT;
      ldg = Mp; kbeg = rt * NB; tri_first = true;
      Aop = d.U;
      if (mode == SGPR_GEMM_TT) { Bop = d.Q; out = d.TT; }
      else Bop = d.TT;
      break;
    }
  }
  extern __shared__ double smem[];
  double* Cs = smem;                       // aliases the pipeline stages
  double* xr = smem + SM_TILE_D;
  double* xc = xr + NB * GPSAT_DMAX;
  double* red = xc + NB * GPSAT_DMAX;
  const int tid = threadIdx.x;
  Acc acc;
  acc_zero(acc);
  const double* Ag = Aop + (size_t)rt * NB * ldg;
  const double* Bg = Bop + (size_t)ct * NB * ldg;
  gemm_nt_prologue(Ag, Bg, ldg, kbeg, kend, smem, tid);
  if (mode == SGPR_GEMM_GC) {
    load_scaled_coords(d.Z, d.M, d.D, d.inv_ls, rt * NB, xr, tid);
    load_scaled_coords(d.X, d.n, d.D, d.inv_ls, ct * NB, xc, tid);
  } else if (mode == SGPR_GEMM_GK) {
    load_scaled_coords(d.Z, d.M, d.D, d.inv_ls, rt * NB, xr, tid);
    load_scaled_coords(d.Z, d.M, d.D, d.inv_ls, ct * NB, xc, tid);
  }
  gemm_nt_mainloop(acc, Ag, Bg, ldg, kbeg, kend, smem, tid, tri_first, tri_last);
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3, lr = lane >> 2, lk = lane & 3;

  if (mode == SGPR_GEMM_GC || mode == SGPR_GEMM_GK) {
    // ---- contraction epilogue: the tile of G never leaves the SM ----
    acc_to_smem(acc, Cs, tid, mode == SGPR_GEMM_GK ? 0.5 : 1.0);
    __syncthreads();
    const int nr = min(NB, max(0, d.M - rt * NB));
    const int nc = min(NB, max(0, (mode == SGPR_GEMM_GC ? d.n : d.M) - ct * NB));
    double part[GPSAT_DMAX + 2];
    cross_grad_tile(d.kind, d.D, xr, xc, nr, nc, Cs, tid, part);
    if (d.zpc != nullptr) {
      const int nct = (mode == SGPR_GEMM_GC) ? d.npad / NB : MT;
      double* zp = (mode == SGPR_GEMM_GC ? d.zpc : d.zpk) + (size_t)(rt * nct + ct) * (2 * NB * GPSAT_DMAX);
      cross_gradz_tile(d.kind, d.D, xr, xc, nr, nc, Cs, tid, zp);
    }
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
      double* gp = (mode == SGPR_GEMM_GC) ? d.gpc : d.gpk;
      const int ncols = (mode == SGPR_GEMM_GC) ? d.npad / NB : MT;
      gp[(size_t)(rt * ncols + ct) * (GPSAT_DMAX + 2) + tid] = v;
    }
    return;
  }

  const int row0 = rt * NB + wm * 64 + lr, col0 = ct * NB + wn * 32 + 2 * lk;
  if (mode == SGPR_GEMM_V) {
    // tile (u, i) of V = WB A, stored as Vt[i][u] (lanes with equal lk write 8 consecutive doubles)
#pragma unroll
    for (int fm = 0; fm < 8; ++fm)
#pragma unroll
      for (int fn = 0; fn < 4; ++fn) {
        double* o = d.Aq + (size_t)(col0 + fn * 8) * Mp + row0 + fm * 8;
        o[0] = acc.c[fm][fn][0];
        o[Mp] = acc.c[fm][fn][1];
      }
    return;
  }
  if (mode == SGPR_GEMM_H) {
    // tile (u, i) of UB Vt^T; Ht[i][u] = (At[i][u] - acc) / sigma + r_i gamma_u / s2, zero for u >= M
    const double isig = rsqrt(d.noise), is2 = 1.0 / d.noise;
#pragma unroll
    for (int fm = 0; fm < 8; ++fm) {
      const int u = row0 + fm * 8;
      const double gu = u < d.M ? d.gam[u] * is2 : 0.0;
#pragma unroll
      for (int fn = 0; fn < 4; ++fn) {
        const int i = col0 + fn * 8;
        const size_t o = (size_t)i * Mp + u;
        const double v0 = u < d.M ? (d.At[o] - acc.c[fm][fn][0]) * isig + d.res[i] * gu : 0.0;
        const double v1 = u < d.M ? (d.At[o + Mp] - acc.c[fm][fn][1]) * isig + d.res[i + 1] * gu : 0.0;
        d.Ct[o] = v0;
        d.Ct[o + Mp] = v1;
      }
    }
    return;
  }
  const double alpha = (mode == SGPR_GEMM_A) ? rsqrt(d.noise) : 1.0;
#pragma unroll
  for (int fm = 0; fm < 8; ++fm)
#pragma unroll
    for (int fn = 0; fn < 4; ++fn)
      *reinterpret_cast<double2*>(out + (size_t)(row0 + fm * 8) * ldo + col0 + fn * 8) =
          make_double2(alpha * acc.c[fm][fn][0], alpha * acc.c[fm][fn][1]);
  if (mode == SGPR_GEMM_A && with_t) {
    // transposed copy At[i][u]: lanes with equal lk write 8 consecutive doubles
#pragma unroll
    for (int fm = 0; fm < 8; ++fm)
#pragma unroll
      for (int fn = 0; fn < 4; ++fn) {
        double* o = d.At + (size_t)(col0 + fn * 8) * Mp + row0 + fm * 8;
        o[0] = alpha * acc.c[fm][fn][0];
        o[Mp] = alpha * acc.c[fm][fn][1];
      }
  }
}

// Bm = sum of the split-K partials in a fixed order (+ I on the first M diagonal entries when add_eye),
// lower tiles.
__global__ void __launch_bounds__(256) sgpr_reduce_kernel(const SgprDesc* __restrict__ descs, int nsplit, int add_eye) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int Mp = d.Mp;
  const size_t total = (size_t)Mp * Mp;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx / Mp), c = (int)(idx - (size_t)r * Mp);
    if (c / NB > r / NB) continue;
    double v = 0.0;
    const int own = (d.npad + SGPR_KCHUNK - 1) / SGPR_KCHUNK;
    for (int s = 0; s < own; ++s) v += d.part[(size_t)s * total + idx];
    if (add_eye && r == c && r < d.M) v += 1.0;
    d.Bm[idx] = v;
  }
}

// entry k < M of c = LB^-1 A y / sigma: row M of the factored B slot
__device__ __forceinline__ double sgpr_c_entry(const SlotDesc& sb, int M, int k) {
  const int c0 = (sb.T - 1) * NB;
  return (k >= c0) ? sb.Ld[(size_t)(sb.T - 1) * NB * NB + (size_t)(M - c0) * NB + (k - c0)]
                   : sb.A[(size_t)M * sb.ld + k];
}

// One warp per row u < Mp: gamma_u = <UB_u, c>, bid_u = |UB_u|^2 = (B^-1)_uu  (0 for u >= M).
__global__ void __launch_bounds__(256) sgpr_gamma_kernel(const SgprDesc* __restrict__ descs,
                                                         const SlotDesc* __restrict__ slotsB) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const SlotDesc& sb = slotsB[blockIdx.y];
  const int u = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (u >= d.Mp) return;
  double s1 = 0.0, s2 = 0.0;
  if (u < d.M) {
    const double* row = d.UB + (size_t)u * d.Mp;
    for (int k = u - (u & 31) + lane; k < d.M; k += 32) {
      if (k < u) continue;
      const double v = row[k];
      s1 = fma(v, sgpr_c_entry(sb, d.M, k), s1);
      s2 = fma(v, v, s2);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (lane == 0) { d.gam[u] = s1; d.bid[u] = s2; }
}

// r_i = y_i - sigma <At_i, gamma> (0 for i >= n); one warp per row, 64 rows per CTA, and the CTA's
// share of r'r in a fixed order.
__global__ void __launch_bounds__(256) sgpr_resid_kernel(const SgprDesc* __restrict__ descs) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int i0 = blockIdx.x * SGPR_RES_ROWS;
  if (i0 >= d.npad) return;
  __shared__ double rs[SGPR_RES_ROWS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const double sigma = sqrt(d.noise);
  for (int q = 0; q < SGPR_RES_ROWS / 8; ++q) {
    const int i = i0 + warp * (SGPR_RES_ROWS / 8) + q;
    double s = 0.0;
    if (i < d.n) {
      const double* row = d.At + (size_t)i * d.Mp;
      for (int k = lane; k < d.M; k += 32) s = fma(row[k], d.gam[k], s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      const double r = (i < d.n) ? d.y[i] - sigma * s : 0.0;
      d.res[i] = r;
      rs[i - i0] = r * r;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
    for (int q = 0; q < SGPR_RES_ROWS; ++q) v += rs[q];
    d.rrp[blockIdx.x] = v;
  }
}

// Q = 2 I - B - B^-1 - gamma gamma^T on [0, M)^2 (both triangles), zero elsewhere.
__global__ void __launch_bounds__(256) sgpr_q_kernel(const SgprDesc* __restrict__ descs) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int Mp = d.Mp, M = d.M;
  const size_t total = (size_t)Mp * Mp;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx / Mp), c = (int)(idx - (size_t)r * Mp);
    double v = 0.0;
    if (r < M && c < M) {
      const size_t lo = (r >= c) ? idx : (size_t)c * Mp + r;
      v = (r == c ? 2.0 : 0.0) - d.Bm[lo] - d.Binv[lo] - d.gam[r] * d.gam[c];
    }
    d.Q[idx] = v;
  }
}

// Block-reduced scalars of one evaluation, one CTA per slot, fixed summation order:
//   sums[p], p = 0..8 : sum over the tiles of gpc + gpk (gradient partial sums; with_grad)
//   sums[9] = r'r, sums[10] = tr B^-1 (with_grad), sums[11] = tr B
constexpr int SGPR_NSUMS = 12;
__global__ void __launch_bounds__(256) sgpr_sums_kernel(const SgprDesc* __restrict__ descs, int with_grad) {
  const SgprDesc& d = descs[blockIdx.x];
  if (d.expert < 0) return;
  const int tid = threadIdx.x;
  double acc[SGPR_NSUMS];
#pragma unroll
  for (int p = 0; p < SGPR_NSUMS; ++p) acc[p] = 0.0;
  for (int u = tid; u < d.M; u += 256) acc[11] += d.Bm[(size_t)u * d.Mp + u];
  if (with_grad) {
    const int MT = d.Mp / NB, ntc = (d.npad / NB) * MT, ntk = MT * MT;
    for (int t = tid; t < ntc; t += 256)
#pragma unroll
      for (int p = 0; p <= GPSAT_DMAX; ++p) acc[p] += d.gpc[(size_t)t * (GPSAT_DMAX + 2) + p];
    for (int t = tid; t < ntk; t += 256)
#pragma unroll
      for (int p = 0; p <= GPSAT_DMAX; ++p) acc[p] += d.gpk[(size_t)t * (GPSAT_DMAX + 2) + p];
    for (int q = tid; q < d.npad / SGPR_RES_ROWS; q += 256) acc[9] += d.rrp[q];
    for (int u = tid; u < d.M; u += 256) acc[10] += d.bid[u];
  }
  __shared__ double red[256];
  for (int p = 0; p < SGPR_NSUMS; ++p) {
    red[tid] = acc[p];
    __syncthreads();
    for (int h = 128; h > 0; h >>= 1) {
      if (tid < h) red[tid] += red[tid + h];
      __syncthreads();
    }
    if (tid == 0) d.sums[p] = red[0];
    __syncthreads();
  }
}

// elbo (gpflow.models.SGPR.elbo) and, with_grad, d elbo / d theta from the by-products of one
// evaluation (after sgpr_sums_kernel).  Returns -inf when a factorisation failed.
__device__ __forceinline__ double sgpr_finish_eval(const SgprDesc& d, const SlotDesc& sk, const SlotDesc& sb,
                                                   double* gtheta, bool with_grad) {
  double logdet = 0.0;
  for (int t = 0; t < sb.T; ++t) logdet += sb.logdet[t];
  const double trB = d.sums[11];
  const double yy = d.Bm[(size_t)d.M * d.Mp + d.M];          // y'y / s2
  const double n = (double)d.n, s2 = d.noise;
  double v = -0.5 * n * 1.8378770664093453 - logdet - 0.5 * n * log(s2) - 0.5 * yy + 0.5 * (*sb.quad) -
             0.5 * n * d.var / s2 + 0.5 * (trB - (double)d.M);
  const bool bad = *sk.notpd || *sb.notpd;
  if (with_grad) {
    const double rr = d.sums[9], trBinv = d.sums[10];
    for (int k = 0; k < d.D; ++k) gtheta[k] = d.var * d.sums[k] * d.inv_ls[k];
    gtheta[d.D] = d.sums[GPSAT_DMAX] - 0.5 * n / s2;
    gtheta[d.D + 1] = -0.5 * n / s2 + 0.5 * (rr + n * d.var - s2 * (trB - 2.0 * d.M + trBinv)) / (s2 * s2);
  }
  return bad ? -INFINITY : v;
}

// d elbo / d Z of one slot from the per-tile row sums (fixed order: Kuf tiles left to right, then the Kuu tiles):
//   gradz[z_k][d] = var / l_d * ( sum_tiles zpc + 2 sum_tiles zpk )     (G_K is symmetric: row and column of Kuu move with z_k)
// written at gradz + (off_z[expert] + k) * D + d.  One thread per (k, d).
__global__ void __launch_bounds__(256) sgpr_gradz_kernel(const SgprDesc* __restrict__ descs,
                                                         const int64_t* __restrict__ off_z, double* __restrict__ gradz) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0 || d.zpc == nullptr) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = idx / GPSAT_DMAX, dd = idx - k * GPSAT_DMAX;
  if (k >= d.M || dd >= d.D) return;
  const int rt = k / NB, r = k - rt * NB, NT = d.npad / NB, MT = d.Mp / NB;
  constexpr size_t TS = 2 * NB * GPSAT_DMAX;
  double sc = 0.0, sk = 0.0;
  for (int ct = 0; ct < NT; ++ct) {
    const double* zp = d.zpc + (size_t)(rt * NT + ct) * TS + (size_t)r * GPSAT_DMAX + dd;
    sc += zp[0] + zp[NB * GPSAT_DMAX];
  }
  for (int ct = 0; ct < MT; ++ct) {
    const double* zp = d.zpk + (size_t)(rt * MT + ct) * TS + (size_t)r * GPSAT_DMAX + dd;
    sk += zp[0] + zp[NB * GPSAT_DMAX];
  }
  gradz[(size_t)(off_z[d.expert] + k) * d.D + dd] = d.var * d.inv_ls[dd] * (sc + 2.0 * sk);
}

// One warp per prediction point s: mean = <T2_s, c>, var = var_f + |T2_s|^2 - |T1_s|^2
// (SGPR.predict_f, full_cov = False; y_var adds the likelihood variance).
__global__ void sgpr_predict_reduce_kernel(const SgprDesc* __restrict__ descs, const SlotDesc* __restrict__ slotsB,
                                           const int64_t* __restrict__ off_m, double* __restrict__ fmean,
                                           double* __restrict__ fvar, double* __restrict__ yvar) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const SlotDesc& sb = slotsB[blockIdx.y];
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= d.m) return;
  const double* t1 = d.T1 + (size_t)warp * d.Mp;
  const double* t2 = d.T2 + (size_t)warp * d.Mp;
  const int M = d.M;
  double s1 = 0.0, s2 = 0.0, s3 = 0.0;
  for (int k = lane; k < M; k += 32) {
    const double cv = sgpr_c_entry(sb, M, k);
    const double a = t2[k], b = t1[k];
    s1 = fma(a, cv, s1); s2 = fma(a, a, s2); s3 = fma(b, b, s3);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    s3 += __shfl_xor_sync(0xffffffffu, s3, o);
  }
  if (lane == 0) {
    const int64_t o = off_m[d.expert] + warp;
    const double fv = d.var + s2 - s3;
    fmean[o] = s1; fvar[o] = fv; yvar[o] = fv + d.noise;
  }
}

// sgpr_predict_cov_kernel: full posterior covariance of the prediction points of one slot (SGPR.predict_f with
// full_cov = True, reached through GPflowSGPRModel.predict(full_cov=True), gpflow_models.py:229-230, 245-263):
//   cov[s][t] = k(xs_s, xs_t) - <T1_s, T1_t> + <T2_s, T2_t>,  T1 = Kus^T L^-T, T2 = T1 LB^-T  (rows of d.T1 / d.T2, columns < M)
// written as a dense m x m block at fcov + off_c[expert].  64 x 64 output tile per CTA, 4 x 4 per thread, plain DFMA:
// an occasional diagnostic output like predict_cov_kernel, not part of the timed hot path.
__global__ void __launch_bounds__(256) sgpr_predict_cov_kernel(const SgprDesc* __restrict__ descs,
                                                               const int64_t* __restrict__ off_c, double* __restrict__ fcov) {
  const SgprDesc& d = descs[blockIdx.z];
  if (d.expert < 0) return;
  const int m = d.m, M = d.M, D = d.D, Mp = d.Mp;
  const int s0 = blockIdx.y * 64, t0 = blockIdx.x * 64;
  if (s0 >= m || t0 >= m || t0 > s0) return;          // lower tiles; mirrored on store
  __shared__ double A1[64][17], B1[64][17], A2[64][17], B2[64][17];
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  for (int k0 = 0; k0 < M; k0 += 16) {
    for (int idx = tid; idx < 64 * 16; idx += 256) {
      const int r = idx >> 4, k = idx & 15;
      const bool ka = k0 + k < M;
      const bool ra = s0 + r < m && ka, rb = t0 + r < m && ka;
      A1[r][k] = ra ? d.T1[(size_t)(s0 + r) * Mp + k0 + k] : 0.0;
      A2[r][k] = ra ? d.T2[(size_t)(s0 + r) * Mp + k0 + k] : 0.0;
      B1[r][k] = rb ? d.T1[(size_t)(t0 + r) * Mp + k0 + k] : 0.0;
      B2[r][k] = rb ? d.T2[(size_t)(t0 + r) * Mp + k0 + k] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      double a1[4], a2[4], b1[4], b2[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) { a1[a] = A1[ty * 4 + a][k]; a2[a] = A2[ty * 4 + a][k]; }
#pragma unroll
      for (int b = 0; b < 4; ++b) { b1[b] = B1[tx * 4 + b][k]; b2[b] = B2[tx * 4 + b][k]; }
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = fma(a2[a], b2[b], fma(-a1[a], b1[b], acc[a][b]));
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
        const double df = (d.Xs[(size_t)s * D + dd] - d.Xs[(size_t)t * D + dd]) * d.inv_ls[dd];
        r2 += df * df;
      }
      const double v = d.var * kunit_dyn(d.kind, r2) + acc[a][b];
      out[(size_t)s * m + t] = v;
      out[(size_t)t * m + s] = v;
    }
}

}  // namespace gpsat
