// Variational Fourier features (the reference's GPflowVFFModel -> GPR_kron, GPSat/vff.py:564-675,
// GPSat/models/vff_model.py:14-264) on the tile machinery of the exact path.
//
// Static per expert (theta independent, vff.py:600-610):
//   feat_d [M_d][n]   cos / sin features of dimension d (make_Kuf_np, :518-523)
//   Kuf    [Mp][npad] rows I = (i_1, .., i_D) = prod_d feat_d[i_d] (make_kvs_np, :553-559); row M = y
//   Phi'   [Mp][Mp]   Kuf Kuf^T by the split-K SYRK of the sparse path: Phi'[M][J] = (Kuf y)_J,
//                     Phi'[M][M] = y'y
// Per evaluation (GPR_kron.elbo, :612-643):
//   per dimension K_d = diag(d) + W W^T (make_Kuu, :381-454), its inverse and their derivatives
//   w.r.t. (lengthscale_d, variance_d) from the low-rank structure (vff_dim_kernel);
//   P' = Phi / s2 + kron_d K_d, row M = Kuf y / s2  -> factored by the sweep as a PRECOMPUTED slot:
//   log|P|, c = L^-1 Kuf y / s2, c'c, and U = L_aug^-T for the gradient;
//   T = P^-1 + beta beta^T tile by tile (U U^T including the -beta column, as lauum_grad_kernel),
//   contracted with kron(.., dK_d, ..); Phi contracted with kron(.., K_d^-1 or dK_d^-1, ..).
// Prediction (GPR_kron.predict_f, :645-675): prediction rows Kus^T ride along in the sweep
// (tmp = L^-1 Kus), var = prod var_d + |tmp_s|^2 - prod_d phi_d' K_d^-1 phi_d.
// theta layout: [lengthscale_1..D, variance_1..D, likelihood variance], D <= 3.
#pragma once
#include "tiles.cuh"
#include "lauum_kernel.cuh"
#include "sgpr_kernels.cuh"

namespace gpsat {

constexpr int VFF_DMAX = 3;
constexpr int VFF_MD_MAX = 256;          // features per dimension (2 * frequencies - 1): up to 128 frequencies
constexpr int VFF_DIM_SMEM_BYTES = 30 * VFF_MD_MAX * 8;      // per-feature tables of vff_dim_kernel (dynamic: 60 KB)
constexpr int VFF_NPART = 16;            // per-tile partial sums, see vff_grad_kernel
constexpr int VFF_NMAT = 6;              // K, dK/dl, dK/dv, K^-1, dK^-1/dl, dK^-1/dv

struct VffDesc {
  int expert;            // -1 = idle
  int fresh;             // 1 until the static quantities of this expert have been built
  int n, m, M, D, kind, npad, mpad, Mp;
  int Md[VFF_DMAX], nf[VFF_DMAX];
  double a[VFF_DMAX], b[VFF_DMAX];
  double ell[VFF_DMAX], var[VFF_DMAX], noise;
  const double* X; const double* y; const double* Xs;
  double* feat;          // [D][VFF_MD_MAX][npad]
  double* featS;         // [D][VFF_MD_MAX][mpad]
  double* Kuf;           // [Mp][npad]
  double* part;          // [nsplit][Mp][Mp]
  double* Phi;           // [Mp][Mp] lower tiles
  double* Pm;            // [Mp][Mp] lower tiles
  double* KusT;          // [mpad][Mp]
  double* dimw;          // [D][VFF_NMAT][VFF_MD_MAX^2]
  double* dims;          // [D][4]: log|K_d|, tr(K_d^-1 dK_d/dl), tr(K_d^-1 dK_d/dv), unused
  double* gp;            // [tiles][VFF_NPART]
};

__device__ __forceinline__ const double* vff_mat(const VffDesc& d, int dim, int which) {
  return d.dimw + ((size_t)dim * VFF_NMAT + which) * VFF_MD_MAX * VFF_MD_MAX;
}

// I -> (i_1, .., i_D), first dimension slowest (make_kvs_np)
__device__ __forceinline__ void vff_unravel(const VffDesc& d, int I, int* idx) {
  for (int dd = d.D - 1; dd >= 0; --dd) { idx[dd] = I % d.Md[dd]; I /= d.Md[dd]; }
}

// ---- feature tables: which = 0 data (make_Kuf_np), which = 1 prediction points (make_Kuf with the
// kernel's edge values outside [a, b], vff.py:472-497; Matern52 has none, :498-511) -----------------
__global__ void __launch_bounds__(256) vff_feat_kernel(const VffDesc* __restrict__ descs, int which) {
  const VffDesc& d = descs[blockIdx.z];
  if (d.expert < 0 || (which == 0 && !d.fresh)) return;
  const int dd = blockIdx.y;
  if (dd >= d.D) return;
  const int np = which ? d.m : d.n, npad = which ? d.mpad : d.npad;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npad) return;
  double* out = (which ? d.featS : d.feat) + (size_t)dd * VFF_MD_MAX * npad + i;
  const int Md = d.Md[dd], nf = d.nf[dd];
  if (i >= np) {
    for (int r = 0; r < Md; ++r) out[(size_t)r * npad] = 0.0;
    return;
  }
  const double a = d.a[dd], b = d.b[dd];
  const double x = (which ? d.Xs : d.X)[(size_t)i * d.D + dd];
  const bool lt = which && x < a, gt = which && x > b;
  double ecos = 0.0, esin = 0.0;      // edge values: cos rows = ecos, sin rows = esin * omega
  if (lt || gt) {
    const double edge = lt ? a : b;
    if (d.kind == GPSAT_MATERN12) { ecos = exp(-fabs(x - edge) / d.ell[dd]); esin = 0.0; }
    else {
      const double arg = 1.7320508075688772 * fabs(x - edge) / d.ell[dd];
      const double e = exp(-arg);
      ecos = (1.0 + arg) * e;
      esin = (x - edge) * e;
    }
  }
  for (int r = 0; r < Md; ++r) {
    const int fr = r < nf ? r : r - nf + 1;
    const double om = 2.0 * 3.141592653589793 * (double)fr / (b - a);
    double v;
    if (lt || gt) v = r < nf ? ecos : esin * om;
    else v = r < nf ? cos(om * (x - a)) : sin(om * (x - a));
    out[(size_t)r * npad] = v;
  }
}

// ---- which = 0: Kuf[I][i] (row M = y); which = 1: KusT[s][I] ---------------------------------------
__global__ void __launch_bounds__(GPSAT_THREADS) vff_kuf_build_kernel(const VffDesc* __restrict__ descs, int which) {
  const VffDesc& d = descs[blockIdx.z];
  if (d.expert < 0 || (which == 0 && !d.fresh)) return;
  const int tid = threadIdx.x, c = tid & (NB - 1);
  if (which == 0) {
    const int i0 = blockIdx.x * NB, I0 = blockIdx.y * NB;
    if (i0 >= d.npad || I0 >= d.Mp) return;
    const int i = i0 + c;
    for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
      const int I = I0 + r;
      double v = 0.0;
      if (I < d.M) {
        int idx[VFF_DMAX];
        vff_unravel(d, I, idx);
        v = 1.0;
        for (int dd = 0; dd < d.D; ++dd) v *= d.feat[((size_t)dd * VFF_MD_MAX + idx[dd]) * d.npad + i];
      } else if (I == d.M && i < d.n) v = d.y[i];
      d.Kuf[(size_t)I * d.npad + i] = v;
    }
  } else {
    const int s0 = blockIdx.x * NB, I0 = blockIdx.y * NB;
    if (s0 >= d.mpad || I0 >= d.Mp) return;
    const int I = I0 + c;
    int idx[VFF_DMAX] = {0, 0, 0};
    if (I < d.M) vff_unravel(d, I, idx);
    for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
      const int s = s0 + r;
      double v = 0.0;
      if (I < d.M && s < d.m) {
        v = 1.0;
        for (int dd = 0; dd < d.D; ++dd) v *= d.featS[((size_t)dd * VFF_MD_MAX + idx[dd]) * d.mpad + s];
      }
      d.KusT[(size_t)s * d.Mp + I] = v;
    }
  }
}

// ---- per-dimension Kuu matrices from the diag + low-rank structure (make_Kuu, vff.py:381-454) -----
// K = diag(dv) + W W^T (W has <= 3 columns, block structured: cos rows / sin rows).
//   K^-1 = D^-1 - R C R^T,  R = D^-1 W,  C = (I + W^T R)^-1,  log|K| = sum log dv + log det(I + W^T R)
//   dK_p = diag(dd_p) + dW_p W^T + W dW_p^T
//   K^-1 dK_p K^-1 = diag(dd_p / dv^2) - (E_p C R^T + R C E_p^T) + R (C G_p C) R^T + Yd_p Y^T + Y Yd_p^T
//   with Y = K^-1 W = R C, Yd_p = K^-1 dW_p = D^-1 dW_p - R C (R^T dW_p), E_p = diag(dd_p / dv) R,
//   G_p = R^T diag(dd_p) R.
// One CTA per (dimension, slot) writes the six dense M_d x M_d matrices and three scalars.
__device__ __forceinline__ void inv3_sym(const double* S, double* C, double& det) {
  const double a = S[0], b = S[1], c = S[2], d = S[4], e = S[5], f = S[8];
  const double c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
  det = a * c00 + b * c01 + c * c02;
  const double id = 1.0 / det;
  C[0] = c00 * id; C[1] = c01 * id; C[2] = c02 * id;
  C[3] = C[1]; C[4] = (a * f - c * c) * id; C[5] = (b * c - a * e) * id;
  C[6] = C[2]; C[7] = C[5]; C[8] = (a * d - b * b) * id;
}

__global__ void __launch_bounds__(128) vff_dim_kernel(const VffDesc* __restrict__ descs) {
  const VffDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const int dim = blockIdx.x;
  if (dim >= d.D) return;
  constexpr int MX = VFF_MD_MAX;
  // per-feature tables (30 rows of MX doubles) in dynamic shared memory
  extern __shared__ double vff_dim_sm[];
  double* dv = vff_dim_sm;                                                          // [MX]
  double (*ddp)[MX] = reinterpret_cast<double (*)[MX]>(vff_dim_sm + MX);            // [2][MX]
  double (*W)[MX] = reinterpret_cast<double (*)[MX]>(vff_dim_sm + 3 * MX);          // [3][MX]
  double (*dW)[3][MX] = reinterpret_cast<double (*)[3][MX]>(vff_dim_sm + 6 * MX);   // [2][3][MX]
  double (*R)[MX] = reinterpret_cast<double (*)[MX]>(vff_dim_sm + 12 * MX);         // [3][MX]
  double (*Y)[MX] = reinterpret_cast<double (*)[MX]>(vff_dim_sm + 15 * MX);         // [3][MX]
  double (*Yd)[3][MX] = reinterpret_cast<double (*)[3][MX]>(vff_dim_sm + 18 * MX);  // [2][3][MX]
  double (*Ep)[3][MX] = reinterpret_cast<double (*)[3][MX]>(vff_dim_sm + 24 * MX);  // [2][3][MX]
  __shared__ double S3[9], C3[9], H[2][9], G[2][9], CGC[2][9], CH[2][9], red[128];
  __shared__ double logdetS;
  const int tid = threadIdx.x, Md = d.Md[dim], nf = d.nf[dim];
  const double a = d.a[dim], b = d.b[dim], ell = d.ell[dim], var = d.var[dim];
  int pw, qw; double c0, coef0, coefn;
  if (d.kind == GPSAT_MATERN12) { pw = 1; qw = 1; c0 = 1.0; coef0 = 0.5; coefn = 0.25; }
  else if (d.kind == GPSAT_MATERN32) { pw = 2; qw = 3; c0 = 1.7320508075688772; coef0 = 0.25; coefn = 0.125; }
  else { pw = 3; qw = 5; c0 = 2.23606797749979; coef0 = 3.0 / 16.0; coefn = 3.0 / 32.0; }
  const double lamb = c0 / ell, dlamb = -lamb / ell, isv = rsqrt(var);
  for (int i = tid; i < Md; i += blockDim.x) {
    const bool is_cos = i < nf;
    const int fr = is_cos ? i : i - nf + 1;
    const double om = 2.0 * 3.141592653589793 * (double)fr / (b - a);
    const double s = lamb * lamb + om * om;
    double sp = s, lq = lamb;
    for (int k = 1; k < pw; ++k) sp *= s;
    for (int k = 1; k < qw; ++k) lq *= lamb;
    const double dd0 = (b - a) * ((is_cos && fr == 0) ? coef0 : coefn) * sp / lq / var;
    dv[i] = dd0;
    ddp[0][i] = dd0 * (2.0 * pw * lamb / s - qw / lamb) * dlamb;
    ddp[1][i] = -dd0 / var;
    double w[3] = {0.0, 0.0, 0.0}, wl[3] = {0.0, 0.0, 0.0};
    if (d.kind == GPSAT_MATERN12) { if (is_cos) w[0] = isv; }
    else if (d.kind == GPSAT_MATERN32) {
      if (is_cos) w[0] = isv;
      else { w[1] = om / lamb * isv; wl[1] = -w[1] / lamb * dlamb; }
    } else {
      if (is_cos) {
        w[0] = (3.0 * (om / lamb) * (om / lamb) - 1.0) * rsqrt(8.0 * var);
        w[1] = isv;
        wl[0] = -6.0 * om * om / (lamb * lamb * lamb) * rsqrt(8.0 * var) * dlamb;
      } else { w[2] = 1.7320508075688772 * om / lamb * isv; wl[2] = -w[2] / lamb * dlamb; }
    }
    for (int k = 0; k < 3; ++k) {
      W[k][i] = w[k]; dW[0][k][i] = wl[k]; dW[1][k][i] = -0.5 * w[k] / var;
      R[k][i] = w[k] / dd0;
      Ep[0][k][i] = ddp[0][i] / dd0 * R[k][i];
      Ep[1][k][i] = ddp[1][i] / dd0 * R[k][i];
    }
  }
  __syncthreads();
  // 3 x 3 sums over the features: S = I + W^T R, H_p = R^T dW_p, G_p = R^T diag(dd_p) R
  if (tid < 45) {
    const int which = tid / 9, ab = tid - which * 9, ai = ab / 3, bi = ab - ai * 3;
    double s = 0.0;
    for (int i = 0; i < Md; ++i) {
      if (which == 0) s += W[ai][i] * R[bi][i];
      else if (which <= 2) s += R[ai][i] * dW[which - 1][bi][i];
      else s += R[ai][i] * ddp[which - 3][i] * R[bi][i];
    }
    if (which == 0) S3[ab] = s + (ai == bi ? 1.0 : 0.0);
    else if (which <= 2) H[which - 1][ab] = s;
    else G[which - 3][ab] = s;
  }
  __syncthreads();
  if (tid == 0) {
    double det;
    inv3_sym(S3, C3, det);
    logdetS = log(det);
    for (int p = 0; p < 2; ++p)
      for (int ai = 0; ai < 3; ++ai)
        for (int bi = 0; bi < 3; ++bi) {
          double ch = 0.0, cgc = 0.0;
          for (int k = 0; k < 3; ++k) {
            ch += C3[ai * 3 + k] * H[p][k * 3 + bi];
            for (int l = 0; l < 3; ++l) cgc += C3[ai * 3 + k] * G[p][k * 3 + l] * C3[l * 3 + bi];
          }
          CH[p][ai * 3 + bi] = ch;
          CGC[p][ai * 3 + bi] = cgc;
        }
  }
  __syncthreads();
  for (int i = tid; i < Md; i += blockDim.x)
    for (int k = 0; k < 3; ++k) {
      double y = 0.0, y0 = 0.0, y1 = 0.0;
      for (int l = 0; l < 3; ++l) {
        y += R[l][i] * C3[l * 3 + k];
        y0 += R[l][i] * CH[0][l * 3 + k];
        y1 += R[l][i] * CH[1][l * 3 + k];
      }
      Y[k][i] = y;
      Yd[0][k][i] = dW[0][k][i] / dv[i] - y0;
      Yd[1][k][i] = dW[1][k][i] / dv[i] - y1;
    }
  __syncthreads();
  double* out = d.dimw + (size_t)dim * VFF_NMAT * MX * MX;
  double tr0 = 0.0, tr1 = 0.0, ld = 0.0;
  for (int e = tid; e < Md * Md; e += blockDim.x) {
    const int i = e / Md, j = e - i * Md;
    const double dij = (i == j) ? 1.0 : 0.0;
    double k = dij * dv[i], ki = dij / dv[i];
    double dk[2] = {dij * ddp[0][i], dij * ddp[1][i]};
    double q[2] = {dij * ddp[0][i] / (dv[i] * dv[i]), dij * ddp[1][i] / (dv[i] * dv[i])};
    for (int x = 0; x < 3; ++x) {
      k += W[x][i] * W[x][j];
      for (int p = 0; p < 2; ++p) {
        dk[p] += dW[p][x][i] * W[x][j] + W[x][i] * dW[p][x][j];
        q[p] += Yd[p][x][i] * Y[x][j] + Y[x][i] * Yd[p][x][j];
      }
      for (int z = 0; z < 3; ++z) {
        const double cxz = C3[x * 3 + z];
        ki -= R[x][i] * cxz * R[z][j];
        for (int p = 0; p < 2; ++p)
          q[p] += R[x][i] * CGC[p][x * 3 + z] * R[z][j] - cxz * (Ep[p][x][i] * R[z][j] + R[x][i] * Ep[p][z][j]);
      }
    }
    const size_t o = (size_t)i * MX + j;
    out[o] = k; out[(size_t)MX * MX + o] = dk[0]; out[(size_t)2 * MX * MX + o] = dk[1];
    out[(size_t)3 * MX * MX + o] = ki; out[(size_t)4 * MX * MX + o] = -q[0]; out[(size_t)5 * MX * MX + o] = -q[1];
    tr0 += ki * dk[0];
    tr1 += ki * dk[1];
    if (i == j) ld += log(dv[i]);
  }
  // fixed-order block sums of the three scalars
  double vals[3] = {ld, tr0, tr1};
  for (int v = 0; v < 3; ++v) {
    red[tid] = vals[v];
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
      for (int t = 0; t < 128; ++t) s += red[t];
      d.dims[dim * 4 + v] = s + (v == 0 ? logdetS : 0.0);
    }
    __syncthreads();
  }
}

// ---- P' (lower tiles): Phi / s2 + kron_d K_d on [0, M)^2, row M = Kuf y / s2 -----------------------
__global__ void __launch_bounds__(GPSAT_THREADS) vff_build_p_kernel(const VffDesc* __restrict__ descs) {
  const VffDesc& d = descs[blockIdx.z];
  if (d.expert < 0) return;
  const int J0 = blockIdx.x * NB, I0 = blockIdx.y * NB;
  if (I0 >= d.Mp || J0 > I0) return;
  const int tid = threadIdx.x, c = tid & (NB - 1), J = J0 + c, M = d.M;
  int jdx[VFF_DMAX] = {0, 0, 0};
  if (J < M) vff_unravel(d, J, jdx);
  const double is2 = 1.0 / d.noise;
  for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
    const int I = I0 + r;
    double v = 0.0;
    if (J < M && I <= M) {
      v = d.Phi[(size_t)I * d.Mp + J] * is2;
      if (I < M) {
        int idx[VFF_DMAX];
        vff_unravel(d, I, idx);
        double k = 1.0;
        for (int dd = 0; dd < d.D; ++dd) k *= vff_mat(d, dd, 0)[idx[dd] * VFF_MD_MAX + jdx[dd]];
        v += k;
      }
    }
    d.Pm[(size_t)I * d.Mp + J] = v;
  }
}

// ---- gradient / trace contractions over the lower tiles of the M x M matrices ---------------------
// part[2 d + p]     = sum w T_IJ   (prod_{e != d} K_e)    dK_{d,p}     [i_d, j_d]      (with_grad)
// part[6 + 2 d + p] = sum w Phi_IJ (prod_{e != d} K_e^-1) dK^-1_{d,p}  [i_d, j_d]      (with_grad)
// part[12] = sum w T_IJ Phi_IJ (with_grad),   part[13] = sum w Phi_IJ prod_e K_e^-1 = tr(Kuu^-1 Phi)
// T = P^-1 + beta beta^T = (U U^T)[0:M, 0:M] with U = L_aug^-T (column M of U is -beta), w = 2 below
// the diagonal, 1 on it.
__global__ void __launch_bounds__(GPSAT_THREADS, 1)
vff_grad_kernel(const VffDesc* __restrict__ descs, const SlotDesc* __restrict__ slots, int with_grad) {
  const VffDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const SlotDesc& s = slots[blockIdx.y];
  const int MT = d.Mp / NB;
  const int t = blockIdx.x;
  if (t >= MT * (MT + 1) / 2) return;
  int ti, tj;
  tri_tile(t, ti, tj);
  extern __shared__ double smem[];
  double* Cs = smem;
  double* red = smem + SM_TILE_D;          // [8 warps][VFF_NPART]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int M = d.M, ld = s.ld, D = d.D;
  if (with_grad) {
    const int kend = ((M + 1 + KC - 1) / KC) * KC;
    const double* Ag = s.A + (size_t)ti * NB * ld;
    const double* Bg = s.A + (size_t)tj * NB * ld;
    Acc acc;
    acc_zero(acc);
    gemm_nt_prologue(Ag, Bg, ld, ti * NB, kend, smem, tid);
    gemm_nt_mainloop(acc, Ag, Bg, ld, ti * NB, kend, smem, tid, true);
    __syncthreads();
    acc_to_smem(acc, Cs, tid, 1.0);
    __syncthreads();
  }
  const int c = tid & (NB - 1), J = tj * NB + c;
  double part[VFF_NPART];
#pragma unroll
  for (int p = 0; p < VFF_NPART; ++p) part[p] = 0.0;
  if (J < M) {
    int jdx[VFF_DMAX] = {0, 0, 0};
    vff_unravel(d, J, jdx);
    for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB) {
      const int I = ti * NB + r;
      if (I >= M || J > I) continue;
      int idx[VFF_DMAX] = {0, 0, 0};
      vff_unravel(d, I, idx);
      const double w = (I == J) ? 1.0 : 2.0;
      const double phi = w * d.Phi[(size_t)I * d.Mp + J];
      double k[VFF_DMAX] = {1.0, 1.0, 1.0}, ki[VFF_DMAX] = {1.0, 1.0, 1.0};
      int off[VFF_DMAX] = {0, 0, 0};
      for (int dd = 0; dd < D; ++dd) {
        off[dd] = idx[dd] * VFF_MD_MAX + jdx[dd];
        k[dd] = vff_mat(d, dd, 0)[off[dd]];
        ki[dd] = vff_mat(d, dd, 3)[off[dd]];
      }
      part[13] += phi * ki[0] * ki[1] * ki[2];
      if (with_grad) {
        const double tv = w * Cs[r * LDC + c];
        part[12] += tv * phi / w;
        for (int dd = 0; dd < D; ++dd) {
          const double ko = (dd == 0 ? 1.0 : k[0]) * (dd == 1 ? 1.0 : k[1]) * (dd == 2 ? 1.0 : k[2]);
          const double kio = (dd == 0 ? 1.0 : ki[0]) * (dd == 1 ? 1.0 : ki[1]) * (dd == 2 ? 1.0 : ki[2]);
          part[2 * dd] += tv * ko * vff_mat(d, dd, 1)[off[dd]];
          part[2 * dd + 1] += tv * ko * vff_mat(d, dd, 2)[off[dd]];
          part[6 + 2 * dd] += phi * kio * vff_mat(d, dd, 4)[off[dd]];
          part[6 + 2 * dd + 1] += phi * kio * vff_mat(d, dd, 5)[off[dd]];
        }
      }
    }
  }
#pragma unroll
  for (int p = 0; p < VFF_NPART; ++p) {
    double v = part[p];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp * VFF_NPART + p] = v;
  }
  __syncthreads();
  if (tid < VFF_NPART) {
    double v = 0.0;
    for (int w8 = 0; w8 < 8; ++w8) v += red[w8 * VFF_NPART + tid];
    d.gp[(size_t)t * VFF_NPART + tid] = v;
  }
}
constexpr int VFF_GRAD_SMEM_BYTES = (SM_TILE_D + 8 * VFF_NPART) * 8;

// elbo (GPR_kron.elbo) and d elbo / d theta (layout [l_1..D, v_1..D, s2]) from the by-products.
__device__ __forceinline__ double vff_finish_eval(const VffDesc& d, const SlotDesc& s, double* gtheta, bool with_grad) {
  double logdetL = 0.0;
  for (int t = 0; t < s.T; ++t) logdetL += s.logdet[t];
  const int MT = d.Mp / NB, nt = MT * (MT + 1) / 2, D = d.D;
  double sum[VFF_NPART];
  for (int p = 0; p < VFF_NPART; ++p) sum[p] = 0.0;
  for (int t = 0; t < nt; ++t)
    for (int p = 0; p < VFF_NPART; ++p) sum[p] += d.gp[(size_t)t * VFF_NPART + p];
  const double n = (double)d.n, s2 = d.noise, yy = d.Phi[(size_t)d.M * d.Mp + d.M], cc = *s.quad;
  double pv = 1.0, kuu_logdet = 0.0;
  for (int dd = 0; dd < D; ++dd) { pv *= d.var[dd]; kuu_logdet += ((double)d.M / d.Md[dd]) * d.dims[dd * 4]; }
  const double trKiPhi = sum[13];
  double v = -0.5 * n * log(2.0 * 3.141592653589793 * s2) - logdetL + 0.5 * kuu_logdet - 0.5 * yy / s2 + 0.5 * cc -
             0.5 * n * pv / s2 + 0.5 * trKiPhi / s2;
  if (with_grad) {
    for (int dd = 0; dd < D; ++dd)
      for (int p = 0; p < 2; ++p) {
        double g = -0.5 * sum[2 * dd + p] + 0.5 * ((double)d.M / d.Md[dd]) * d.dims[dd * 4 + 1 + p] +
                   0.5 * sum[6 + 2 * dd + p] / s2;
        if (p == 1) g -= 0.5 * n * pv / d.var[dd] / s2;
        gtheta[dd + p * D] = g;
      }
    gtheta[2 * D] = -0.5 * n / s2 + 0.5 * sum[12] / (s2 * s2) - cc / s2 + 0.5 * (yy + n * pv - trKiPhi) / (s2 * s2);
  }
  return *s.notpd ? -INFINITY : v;
}

// One warp per prediction point: mean = <tmp_s, c>, var = prod var_d + |tmp_s|^2 - prod_d phi_d' K_d^-1 phi_d
// (GPR_kron.predict_f, vff.py:662-674), tmp_s = row M + 1 + s of the factored slot.
__global__ void __launch_bounds__(256) vff_predict_reduce_kernel(const VffDesc* __restrict__ descs,
                                                                 const SlotDesc* __restrict__ slots,
                                                                 const int64_t* __restrict__ off_m,
                                                                 double* __restrict__ fmean, double* __restrict__ fvar,
                                                                 double* __restrict__ yvar) {
  const VffDesc& d = descs[blockIdx.y];
  if (d.expert < 0) return;
  const SlotDesc& s = slots[blockIdx.y];
  const int sidx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (sidx >= d.m) return;
  const int M = d.M;
  double s1 = 0.0, s2 = 0.0;
  for (int k = lane; k < M; k += 32) {
    const double cv = sgpr_c_entry(s, M, k), a = factor_entry(s, M + 1 + sidx, k);
    s1 = fma(a, cv, s1); s2 = fma(a, a, s2);
  }
  double q = 1.0, pv = 1.0;
  for (int dd = 0; dd < d.D; ++dd) {
    const int Md = d.Md[dd];
    const double* f = d.featS + (size_t)dd * VFF_MD_MAX * d.mpad + sidx;
    const double* Ki = vff_mat(d, dd, 3);
    double qd = 0.0;
    for (int e = lane; e < Md * Md; e += 32) {
      const int i = e / Md, j = e - i * Md;
      qd = fma(f[(size_t)i * d.mpad] * Ki[i * VFF_MD_MAX + j], f[(size_t)j * d.mpad], qd);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) qd += __shfl_xor_sync(0xffffffffu, qd, o);
    q *= qd;
    pv *= d.var[dd];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (lane == 0) {
    const int64_t o = off_m[d.expert] + sidx;
    const double fv = pv + s2 - q;
    fmean[o] = s1; fvar[o] = fv; yvar[o] = fv + d.noise;
  }
}

}  // namespace gpsat
