// Sparse GPR (Titsias) on the same tile machinery: the reference's GPflowSGPRModel
// (GPSat/models/gpflow_models.py:666-901 -> gpflow.models.SGPR.elbo / predict_f).
//
//   Kuu = k(Z,Z) + jitter I = L L^T         sweep of the "K slot" (M inducing points as observations)
//   W   = L^-1                               transpose of U = L^-T produced by the trtri tiles
//   Ct  = Kuf^T  [n][Mp]                     kuf_build_kernel (column M carries y, see below)
//   A'  = W Ct^T / sigma  [Mp][npad]         gemm_nt_kernel; row M of A' comes out as y^T / sigma
//   B'  = A' A'^T (+ I)   [Mp][Mp]           gemm_nt_kernel (split-K) + splitk_reduce_kernel:
//                                            B'[u][v] = B, B'[M][u] = (A y)_u / sigma, B'[M][M] = y'y / s2
//   chol(B') with row M riding along         sweep of the "B slot" on the precomputed matrix:
//                                            log|LB|, c = LB^-1 A y / sigma, quad = c'c
//   predict: T1 = Kus^T W^T, T2 = T1 WB^T, mean = T2 c, var = k** + |T2_s|^2 - |T1_s|^2
#pragma once
#include "tiles.cuh"

namespace gpsat {

constexpr double SGPR_JITTER = 1e-6;     // gpflow.config.default_jitter()

struct SgprDesc {
  int expert;            // -1 = idle
  int n, m, M, D, kind;
  int npad, mpad, Mp;    // n, m, M + 1 rounded up to multiples of 128
  const double* X; const double* y; const double* Z; const double* Xs;
  double inv_ls[GPSAT_DMAX];
  double var, noise;
  double* Ct;            // [npad][Mp]
  double* Aq;            // [Mp][npad]
  double* W;             // [Mp][Mp]  L^-1
  double* WB;            // [Mp][Mp]  LB^-1
  double* Bm;            // [Mp][Mp]
  double* part;          // [nsplit][Mp][Mp]
  double* Cts;           // [mpad][Mp]
  double* T1;            // [mpad][Mp]
  double* T2;            // [mpad][Mp]
};

// ---- Ct[i][k] = k(x_i, z_k); column M = y_i (train) or 0; zero padding elsewhere ----------------
template <int KIND>
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
  ColCoords<0> cc;
  cc.load(zk, c, D);
#pragma unroll 4
  for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
    const int i = i0 + r;
    double v = 0.0;
    if (i < np) {
      if (k < M) v = d.var * kunit_of_r2<KIND>(cc.r2(xi, r, D));
      else if (k == M && yv) v = yv[i];
    }
    out[(size_t)i * d.Mp + k] = v;
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
  switch (d.kind) {
    case GPSAT_MATERN12: kuf_tile<GPSAT_MATERN12>(d, P, np, yv, out, i0, k0, xi, zk, threadIdx.x); break;
    case GPSAT_MATERN32: kuf_tile<GPSAT_MATERN32>(d, P, np, yv, out, i0, k0, xi, zk, threadIdx.x); break;
    case GPSAT_MATERN52: kuf_tile<GPSAT_MATERN52>(d, P, np, yv, out, i0, k0, xi, zk, threadIdx.x); break;
    default: kuf_tile<GPSAT_RBF>(d, P, np, yv, out, i0, k0, xi, zk, threadIdx.x); break;
  }
}

// ---- W = U^T: dense lower-triangular inverse factor of a slot, row-major [Mp][Mp] ---------------
// Off-diagonal 128-blocks come from the strict upper block triangle of the slot matrix (U), the
// diagonal blocks from the W_tt side buffer.  which = 0: descs[].W from slots, 1: descs[].WB.
__global__ void __launch_bounds__(256) transpose_w_kernel(const SgprDesc* __restrict__ descs,
                                                          const SlotDesc* __restrict__ slots, int which) {
  const SgprDesc& d = descs[blockIdx.z];
  if (d.expert < 0) return;
  const SlotDesc& s = slots[blockIdx.z];
  const int Mp = d.Mp;
  const int u0 = blockIdx.y * 32, k0 = blockIdx.x * 32;       // W tile rows u, cols k
  if (u0 >= Mp || k0 >= Mp) return;
  double* W = which ? d.WB : d.W;
  __shared__ double t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
  const int ub = u0 / NB, kb = k0 / NB;
  if (kb > ub) {
    for (int r = ty; r < 32; r += 8) W[(size_t)(u0 + r) * Mp + k0 + tx] = 0.0;
    return;
  }
  if (kb == ub) {
    const double* Wd = s.Wd + (size_t)ub * NB * NB;
    for (int r = ty; r < 32; r += 8) W[(size_t)(u0 + r) * Mp + k0 + tx] = Wd[(size_t)(u0 - ub * NB + r) * NB + (k0 - kb * NB + tx)];
    return;
  }
  for (int r = ty; r < 32; r += 8) t[r][tx] = s.A[(size_t)(k0 + r) * s.ld + u0 + tx];    // U[k][u]
  __syncthreads();
  for (int r = ty; r < 32; r += 8) W[(size_t)(u0 + r) * Mp + k0 + tx] = t[tx][r];
}

// ---- batched NT GEMM on 128 x 128 tiles:  out = alpha * Aop Bop^T  (both operands k-contiguous) ----
enum SgprGemmMode : int {
  SGPR_GEMM_A = 0,     // Aq[Mp][npad]  = W  Ct^T / sigma        (W lower triangular: k <= row tile)
  SGPR_GEMM_SYRK = 1,  // part[split]   = Aq Aq^T over one k range, lower tiles
  SGPR_GEMM_T1 = 2,    // T1[mpad][Mp]  = Cts W^T                 (k <= column tile)
  SGPR_GEMM_T2 = 3     // T2[mpad][Mp]  = T1  WB^T                (k <= column tile)
};
constexpr int SGPR_GEMM_SMEM_BYTES = NSTAGE * STAGE_D * 8;

__global__ void __launch_bounds__(GPSAT_THREADS, 1)
sgpr_gemm_kernel(const SgprDesc* __restrict__ descs, int mode, int nsplit) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int tile = blockIdx.x / nsplit, split = blockIdx.x - tile * nsplit;
  const int Mp = d.Mp, MT = Mp / NB;
  const double *Aop, *Bop;
  double* out;
  int ldg, ldo, rt, ct, kbeg = 0, kend;
  double alpha = 1.0;
  if (mode == SGPR_GEMM_A) {
    const int nct = d.npad / NB;
    if (tile >= MT * nct) return;
    rt = tile / nct; ct = tile - rt * nct;
    Aop = d.W; Bop = d.Ct; ldg = Mp; out = d.Aq; ldo = d.npad;
    kend = (rt + 1) * NB;
    alpha = rsqrt(d.noise);
  } else if (mode == SGPR_GEMM_SYRK) {
    if (tile >= MT * (MT + 1) / 2) return;
    rt = (int)((sqrtf(8.0f * tile + 1.0f) - 1.0f) * 0.5f);
    while (rt * (rt + 1) / 2 > tile) --rt;
    while ((rt + 1) * (rt + 2) / 2 <= tile) ++rt;
    ct = tile - rt * (rt + 1) / 2;
    Aop = d.Aq; Bop = d.Aq; ldg = d.npad; out = d.part + (size_t)split * Mp * Mp; ldo = Mp;
    const int nslab = d.npad / KC, per = (nslab + nsplit - 1) / nsplit;
    kbeg = split * per * KC;
    kend = min(d.npad, (split + 1) * per * KC);
    if (kend < kbeg) kend = kbeg;
  } else {
    const int nrt = d.mpad / NB;
    if (tile >= nrt * MT) return;
    rt = tile / MT; ct = tile - rt * MT;
    Aop = (mode == SGPR_GEMM_T1) ? d.Cts : d.T1;
    Bop = (mode == SGPR_GEMM_T1) ? d.W : d.WB;
    ldg = Mp; out = (mode == SGPR_GEMM_T1) ? d.T1 : d.T2; ldo = Mp;
    kend = (ct + 1) * NB;
  }
  extern __shared__ double smem[];
  const int tid = threadIdx.x;
  Acc acc;
  acc_zero(acc);
  const double* Ag = Aop + (size_t)rt * NB * ldg;
  const double* Bg = Bop + (size_t)ct * NB * ldg;
  gemm_nt_prologue(Ag, Bg, ldg, kbeg, kend, smem, tid);
  gemm_nt_mainloop(acc, Ag, Bg, ldg, kbeg, kend, smem, tid);
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3, lr = lane >> 2, lk = lane & 3;
  double* Og = out + ((size_t)rt * NB + wm * 64 + lr) * ldo + (size_t)ct * NB + wn * 32 + 2 * lk;
#pragma unroll
  for (int fm = 0; fm < 8; ++fm)
#pragma unroll
    for (int fn = 0; fn < 4; ++fn)
      *reinterpret_cast<double2*>(Og + (size_t)fm * 8 * ldo + fn * 8) =
          make_double2(alpha * acc.c[fm][fn][0], alpha * acc.c[fm][fn][1]);
}

// Bm = sum of the split-K partials in a fixed order (+ I on the first M diagonal entries), lower tiles.
__global__ void __launch_bounds__(256) sgpr_reduce_kernel(const SgprDesc* __restrict__ descs, int nsplit) {
  const SgprDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int Mp = d.Mp;
  const size_t total = (size_t)Mp * Mp;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx / Mp), c = (int)(idx - (size_t)r * Mp);
    if (c / NB > r / NB) continue;
    double v = 0.0;
    for (int s = 0; s < nsplit; ++s) v += d.part[(size_t)s * total + idx];
    if (r == c && r < d.M) v += 1.0;
    d.Bm[idx] = v;
  }
}

__global__ void sgpr_reset_kernel(const SlotDesc* __restrict__ slots, int S2) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S2 || slots[s].expert < 0) return;
  *slots[s].quad = 0.0;
  *slots[s].notpd = 0;
}

// elbo[e] from the by-products of the B sweep (gpflow.models.SGPR.elbo).
__global__ void sgpr_finish_kernel(const SgprDesc* __restrict__ descs, const SlotDesc* __restrict__ slotsK,
                                   const SlotDesc* __restrict__ slotsB, double* __restrict__ elbo,
                                   int* __restrict__ status, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const SgprDesc& d = descs[s];
  if (d.expert < 0) return;
  const SlotDesc& sb = slotsB[s];
  double logdet = 0.0;
  for (int t = 0; t < sb.T; ++t) logdet += sb.logdet[t];
  double trB = 0.0;
  for (int u = 0; u < d.M; ++u) trB += d.Bm[(size_t)u * d.Mp + u];
  const double yy = d.Bm[(size_t)d.M * d.Mp + d.M];          // y'y / s2
  const double n = (double)d.n;
  double v = -0.5 * n * 1.8378770664093453 - logdet - 0.5 * n * log(d.noise) - 0.5 * yy + 0.5 * (*sb.quad) -
             0.5 * n * d.var / d.noise + 0.5 * (trB - (double)d.M);
  const bool bad = *slotsK[s].notpd || *sb.notpd;
  elbo[d.expert] = bad ? -INFINITY : v;
  if (status) status[d.expert] = bad ? 1 : 0;
}

// One warp per prediction point s: mean = <T2_s, c>, var = var_f + |T2_s|^2 - |T1_s|^2, c = row M of
// the factored B slot (SGPR.predict_f, full_cov = False; y_var adds the likelihood variance).
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
  const int M = d.M, c0 = (sb.T - 1) * NB;
  const double* crow = sb.A + (size_t)M * sb.ld;
  const double* clast = sb.Ld + (size_t)(sb.T - 1) * NB * NB + (size_t)(M - c0) * NB;
  double s1 = 0.0, s2 = 0.0, s3 = 0.0;
  for (int k = lane; k < M; k += 32) {
    const double cv = (k >= c0) ? clast[k - c0] : crow[k];
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

}  // namespace gpsat