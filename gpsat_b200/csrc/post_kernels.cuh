// Post-processing reductions of the local-expert workflow ("next" rows of SURVEY.md 8f):
//   gaussian_smooth_kernel  - GPSat/postprocessing.py:22-52 (gaussian_2d_weight, the numba kernel
//                             behind smooth_hyperparameters, :96-395)
//   glue_kernel             - GPSat/postprocessing.py:462-593 (glue_local_predictions_1d / _2d:
//                             Gaussian-weighted blend of overlapping expert predictions)
// Both are all-pairs / segmented weighted sums with a fixed summation order.
#pragma once
#include "common.cuh"

namespace gpsat {

// out[f][i] = sum_j w_ij vals[f][j] / sum_j w_ij over the j with non-NaN vals[f][j];
// w_ij = exp(-d2 / 2), d2 = ((x_j - x0_i) / lx)^2 + ((y_j - y0_i) / ly)^2; NaN when the weights sum to 0.
__global__ void __launch_bounds__(128) gaussian_smooth_kernel(int64_t n0, const double* __restrict__ xq,
                                                              const double* __restrict__ yq, int64_t n,
                                                              const double* __restrict__ x, const double* __restrict__ y,
                                                              double lx, double ly, const double* __restrict__ vals,
                                                              double* __restrict__ out) {
  const int64_t i = blockIdx.x;
  const double* v = vals + (size_t)blockIdx.y * n;
  const double x0 = xq[i], y0 = yq[i];
  double ws = 0.0, wv = 0.0;
  for (int64_t j = threadIdx.x; j < n; j += blockDim.x) {
    const double val = v[j];
    if (isnan(val)) continue;
    const double dx = (x[j] - x0) / lx, dy = (y[j] - y0) / ly;
    const double w = exp(-(dx * dx + dy * dy) / 2.0);
    wv = fma(w, val, wv);
    ws += w;
  }
  __shared__ double s0[128], s1[128];
  s0[threadIdx.x] = ws; s1[threadIdx.x] = wv;
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int t = 0; t < 128; ++t) { a += s0[t]; b += s1[t]; }
    out[(size_t)blockIdx.y * n0 + i] = (a == 0.0) ? nan("") : b / a;
  }
}

// Rows are sorted by prediction location; group g owns rows seg[g] .. seg[g + 1).
// w_i = prod_d normpdf(pred[d][i]; mean xprt[d][i], std sigma[i]);  out[v][g] = sum w vals[v] / sum w.
__global__ void __launch_bounds__(128) glue_kernel(int64_t G, const int64_t* __restrict__ seg, int ndim, int64_t n,
                                                   const double* __restrict__ pred, const double* __restrict__ xprt,
                                                   const double* __restrict__ sigma, int64_t V,
                                                   const double* __restrict__ vals, double* __restrict__ out) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= G) return;
  const double INV_SQRT_2PI = 0.3989422804014327;
  for (int64_t v = 0; v < V; ++v) {
    double ws = 0.0, wv = 0.0;
    for (int64_t i = seg[g]; i < seg[g + 1]; ++i) {
      double w = 1.0;
      const double sd = sigma[i];
      for (int d = 0; d < ndim; ++d) {
        const double z = (pred[(size_t)d * n + i] - xprt[(size_t)d * n + i]) / sd;
        w *= exp(-z * z / 2.0) * INV_SQRT_2PI / sd;
      }
      ws += w;
      wv = fma(w, vals[(size_t)v * n + i], wv);
    }
    out[(size_t)v * G + g] = wv / ws;
  }
}

}  // namespace gpsat
