// C ABI of the post-processing reductions (post_kernels.cuh).  Included at the end of engine_vff.inl.
extern "C" {

int gpsat_gaussian_smooth(gpsat_engine* h, int mem, int64_t n0, const double* x0, const double* y0, int64_t n,
                          int64_t F, const double* x, const double* y, double lx, double ly, const double* vals,
                          double* out) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(n >= 0 && n0 >= 0 && F >= 0 && F < 65536, "bad sizes");
  REQUIRE(lx > 0.0 && ly > 0.0, "smoothing length scales must be positive");
  Engine& g = h->e;
  begin_call(g);
  if (n0 == 0 || F == 0) return end_call(g);
  REQUIRE(x0 && y0 && out && (n == 0 || (x && y && vals)), "NULL argument");
  const bool dev = mem == GPSAT_MEM_DEVICE;
  const double *dx0 = x0, *dy0 = y0, *dx = x, *dy = y, *dv = vals;
  double* dout = out;
  if (!dev) {
    if (g.bout.ensure((size_t)((2 + F) * (n + n0)) * 8 + 64)) return 1;
    double* base = g.bout.as<double>();
    dx0 = base; dy0 = base + n0; dx = base + 2 * n0; dy = base + 2 * n0 + n; dv = base + 2 * n0 + 2 * n;
    dout = base + 2 * n0 + (2 + F) * n;
    CUDA_TRY(cudaMemcpyAsync((void*)dx0, x0, (size_t)n0 * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync((void*)dy0, y0, (size_t)n0 * 8, cudaMemcpyHostToDevice, g.stream));
    if (n > 0) {
      CUDA_TRY(cudaMemcpyAsync((void*)dx, x, (size_t)n * 8, cudaMemcpyHostToDevice, g.stream));
      CUDA_TRY(cudaMemcpyAsync((void*)dy, y, (size_t)n * 8, cudaMemcpyHostToDevice, g.stream));
      CUDA_TRY(cudaMemcpyAsync((void*)dv, vals, (size_t)F * n * 8, cudaMemcpyHostToDevice, g.stream));
    }
    g.stats.h2d_bytes += (double)(2 * n0 + (2 + F) * n) * 8;
  }
  gaussian_smooth_kernel<<<dim3((unsigned)n0, (unsigned)F), 128, 0, g.stream>>>(n0, dx0, dy0, n, dx, dy, lx, ly, dv, dout);
  g.stats.kernel_launches++;
  if (!dev) {
    CUDA_TRY(cudaMemcpyAsync(out, dout, (size_t)F * n0 * 8, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)F * n0 * 8;
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

int gpsat_glue_predictions(gpsat_engine* h, int mem, int64_t n, int64_t G, const int64_t* seg, int ndim,
                           const double* pred, const double* xprt, const double* sigma, int64_t V,
                           const double* vals, double* out) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(n >= 0 && G >= 0 && V >= 0 && ndim >= 1 && ndim <= 3, "bad sizes");
  Engine& g = h->e;
  begin_call(g);
  if (G == 0 || V == 0) return end_call(g);
  REQUIRE(seg && pred && xprt && sigma && vals && out, "NULL argument");
  REQUIRE(seg[0] == 0 && seg[G] == n, "seg must cover [0, n)");
  for (int64_t k = 0; k < G; ++k) REQUIRE(seg[k + 1] > seg[k], "every group needs at least one row");
  const bool dev = mem == GPSAT_MEM_DEVICE;
  if (g.bmisc.ensure((size_t)(G + 1) * 8)) return 1;
  CUDA_TRY(cudaMemcpyAsync(g.bmisc.p, seg, (size_t)(G + 1) * 8, cudaMemcpyHostToDevice, g.stream));
  const double *dp = pred, *dx = xprt, *ds = sigma, *dv = vals;
  double* dout = out;
  if (!dev) {
    const size_t tot = (size_t)(2 * ndim + 1 + V) * n + (size_t)V * G;
    if (g.bout.ensure(tot * 8)) return 1;
    double* base = g.bout.as<double>();
    CUDA_TRY(cudaMemcpyAsync(base, pred, (size_t)ndim * n * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync(base + (size_t)ndim * n, xprt, (size_t)ndim * n * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync(base + (size_t)2 * ndim * n, sigma, (size_t)n * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync(base + (size_t)(2 * ndim + 1) * n, vals, (size_t)V * n * 8, cudaMemcpyHostToDevice, g.stream));
    g.stats.h2d_bytes += (double)(2 * ndim + 1 + V) * n * 8;
    dp = base; dx = base + (size_t)ndim * n; ds = base + (size_t)2 * ndim * n; dv = base + (size_t)(2 * ndim + 1) * n;
    dout = base + (size_t)(2 * ndim + 1 + V) * n;
  }
  glue_kernel<<<(unsigned)((G + 127) / 128), 128, 0, g.stream>>>(G, g.bmisc.as<int64_t>(), ndim, n, dp, dx, ds, V, dv, dout);
  g.stats.kernel_launches++;
  if (!dev) {
    CUDA_TRY(cudaMemcpyAsync(out, dout, (size_t)V * G * 8, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)V * G * 8;
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

}  // extern "C"

#include "engine_select.inl"
