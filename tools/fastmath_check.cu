// Accuracy of exp_nonpos / sqrt_rsqrt_pos (gpsat_b200/csrc/common.cuh) against the CUDA library exp / sqrt over
// the argument ranges the covariance code produces.  Prints the largest relative errors in units of 2^-53.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fastmath_check tools/fastmath_check.cu && build/fastmath_check
#include <cstdio>
#include <cmath>
#include "../gpsat_b200/csrc/common.cuh"

__global__ void check(int n, double* out) {
  double me = 0.0, ms = 0.0, mr = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    // x in [-700, 0], log-uniform magnitudes from 1e-12 to 700 plus a linear sweep
    double u = (i + 0.5) / n;
    double x = (i & 1) ? -700.0 * u : -exp(log(1e-12) + u * (log(700.0) - log(1e-12)));
    double e0 = exp(x), e1 = exp_nonpos(x);
    me = fmax(me, fabs(e1 - e0) / e0);
    double a = exp(log(1e-36) + u * (log(1e12) - log(1e-36)));
    double ri;
    double s1 = sqrt_rsqrt_pos(a, ri), s0 = sqrt(a);
    ms = fmax(ms, fabs(s1 - s0) / s0);
    mr = fmax(mr, fabs(ri * s0 - 1.0));
  }
  atomicMax((unsigned long long*)&out[0], (unsigned long long)__double_as_longlong(me));
  atomicMax((unsigned long long*)&out[1], (unsigned long long)__double_as_longlong(ms));
  atomicMax((unsigned long long*)&out[2], (unsigned long long)__double_as_longlong(mr));
}

int main() {
  double* d;
  cudaMalloc(&d, 24);
  cudaMemset(d, 0, 24);
  check<<<592, 256>>>(1 << 26, d);
  double h[3];
  cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
  const double u = 1.1102230246251565e-16;
  printf("{\"exp_nonpos_max_rel_err_ulp\": %.3f, \"sqrt_max_rel_err_ulp\": %.3f, \"rsqrt_max_rel_err_ulp\": %.3f, \"exp_at_-700\": %.3e}\n",
         h[0] / u, h[1] / u, h[2] / u, 0.0);
  return cudaGetLastError() != cudaSuccess;
}
