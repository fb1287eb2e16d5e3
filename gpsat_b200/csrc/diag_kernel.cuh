// diag_kernel: in-shared-memory Cholesky + triangular inverse of one 128 x 128 diagonal tile
// per expert ("batched in-SM Cholesky of K + s2 I" for experts with n < 128, and the panel
// step of the blocked factorisation for larger ones).
#pragma once
#include "tiles.cuh"

namespace gpsat {

// global-memory variant of tile_value (used for tile (0,0) only)
__device__ __forceinline__ double tile_value_g(const SlotDesc& d, int gr, int gc) {
  const int n = d.n;
  if (gc >= n) return (gr == gc && gr > n) ? 1.0 : 0.0;
  if (gr == n) return d.y[gc];
  if (gr >= n + 1 + d.m) return 0.0;
  const double* pr = (gr < n) ? d.X + (size_t)gr * d.D : d.Xs + (size_t)(gr - n - 1) * d.D;
  const double* pc = d.X + (size_t)gc * d.D;
  double r2 = 0.0;
  for (int dd = 0; dd < d.D; ++dd) {
    double t = pr[dd] * d.inv_ls[dd] - pc[dd] * d.inv_ls[dd];
    r2 += t * t;
  }
  double v = d.var * kunit_dyn(d.kind, r2);
  if (gr == gc) v += d.noise;
  return v;
}

__global__ void __launch_bounds__(GPSAT_THREADS, 1) diag_kernel(const SlotDesc* __restrict__ slots, int i) {
  const SlotDesc& d = slots[blockIdx.x];
  if (d.expert < 0 || i >= d.T) return;
  extern __shared__ double smem[];
  double* Cs = smem;                 // [NB][LDD]: lower = S_ii -> L_ii, strict upper = W^T
  double* wdiag = smem + NB * LDD;   // [NB] 1 / L_cc
  const int tid = threadIdx.x;
  const int row0 = i * NB;
  const int n = d.n, ld = d.ld;
  int ncol = n - row0;               // real columns in this tile; local index ncol is the y row
  if (ncol > NB) ncol = NB;
  if (ncol < 0) ncol = 0;

  // ---- load the running Schur complement (tile (0,0): evaluate the kernel directly) ----
  if (i == 0) {
    for (int idx = tid; idx < NB * NB; idx += GPSAT_THREADS) {
      int r = idx >> 7, c = idx & (NB - 1);
      if (c <= r) Cs[r * LDD + c] = tile_value_g(d, r, c);
    }
  } else {
    const double* src = d.A + (size_t)row0 * ld + row0;
    for (int idx = tid; idx < NB * NB; idx += GPSAT_THREADS) {
      int r = idx >> 7, c = idx & (NB - 1);
      if (c <= r) Cs[r * LDD + c] = src[(size_t)r * ld + c];
    }
  }
  __syncthreads();

  // ---- right-looking Cholesky of the first ncol columns (all 128 rows ride along) ----
  const int tx = tid & 15, ty = tid >> 4;
  bool bad = false;
  for (int c = 0; c < ncol; ++c) {
    double piv = Cs[c * LDD + c];
    if (!(piv > 0.0) || !isfinite(piv)) { bad = true; piv = 1.0; }
    double dd = sqrt(piv);
    __syncthreads();                     // everyone has read the pivot
    for (int r = c + 1 + tid; r < NB; r += GPSAT_THREADS) Cs[r * LDD + c] /= dd;
    if (tid == 0) Cs[c * LDD + c] = dd;
    __syncthreads();
    for (int r = c + 1 + ty; r < NB; r += 16) {
      double lr = Cs[r * LDD + c];
      for (int k = c + 1 + tx; k <= r; k += 16) Cs[r * LDD + k] -= lr * Cs[k * LDD + c];
    }
    // the next iteration's pivot read is ordered by the __syncthreads at its top-of-loop below
    __syncthreads();
  }
  if (bad && tid == 0) *d.notpd = 1;

  // ---- quad term, identity on the padding, log-determinant ----
  if (tid == 0 && ncol < NB) *d.quad = -Cs[ncol * LDD + ncol];
  __syncthreads();
  for (int p = ncol + tid; p < NB; p += GPSAT_THREADS) Cs[p * LDD + p] = 1.0;
  __syncthreads();
  if (tid < 32) {
    double s = 0.0;
    for (int c = tid; c < ncol; c += 32) s += log(Cs[c * LDD + c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (tid == 0) d.logdet[i] = s;
  }
  for (int c = tid; c < NB; c += GPSAT_THREADS) wdiag[c] = 1.0 / Cs[c * LDD + c];
  __syncthreads();

  // ---- W = L^-1, column c by forward substitution; two threads per column split the dot ----
  // W[k][c] (k > c) is kept transposed at Cs[c][k] (strict upper triangle).
  {
    const int c = tid >> 1, half = tid & 1;
    for (int r = 1; r < NB; ++r) {
      double s = 0.0;
      if (c < r) {
        // k runs over [c, r-1]; k == c uses W[c][c] = wdiag[c]
        for (int k = r - 1 - half; k > c; k -= 2) s += Cs[r * LDD + k] * Cs[c * LDD + k];
        if (((r - 1 - c) & 1) == half) s += Cs[r * LDD + c] * wdiag[c];
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      if (c < r && half == 0) Cs[c * LDD + r] = -s * wdiag[r];
      __syncthreads();
    }
  }

  // ---- write L_ii, W_ii (side buffers) and U_ii = W_ii^T (into the slot matrix) ----
  double* Ld = d.Ld + (size_t)i * NB * NB;
  double* Wd = d.Wd + (size_t)i * NB * NB;
  double* Ut = d.A + (size_t)row0 * ld + row0;
  for (int idx = tid; idx < NB * NB; idx += GPSAT_THREADS) {
    int r = idx >> 7, c = idx & (NB - 1);
    double lv, wv, uv;
    if (c < r) { lv = Cs[r * LDD + c]; wv = Cs[c * LDD + r]; uv = 0.0; }
    else if (c == r) { lv = Cs[r * LDD + r]; wv = wdiag[r]; uv = wv; }
    else { lv = 0.0; wv = 0.0; uv = Cs[r * LDD + c]; }
    Ld[idx] = lv;
    Wd[idx] = wv;
    Ut[(size_t)r * ld + c] = uv;
  }
}

}  // namespace gpsat
