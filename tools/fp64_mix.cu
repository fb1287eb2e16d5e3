// Do DMMA (FP64 tensor) and DFMA (FP64 vector) share an execution pipe on B200?
// Runs, one 256-thread CTA per SM x 2: (a) DMMA only, (b) DFMA only, (c) warps 0-3 DMMA + warps 4-7 DFMA,
// (d) every warp interleaves 1 DMMA with NF DFMAs.  Prints TFLOP/s per variant as JSON.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_mix tools/fp64_mix.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
constexpr int ITERS = 2048;

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// mode 0: dmma only, 1: dfma only, 2: split by warp, 3: interleaved (NF dfma per dmma)
template <int MODE, int NF>
__global__ void __launch_bounds__(256) k_mix(double* out, double a, double b) {
  double c[8][2], f[16];
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
#pragma unroll
  for (int i = 0; i < 16; ++i) f[i] = threadIdx.x + i;
  const int warp = threadIdx.x >> 5;
  const bool do_mma = MODE == 0 || MODE == 3 || (MODE == 2 && warp < 4);
  const bool do_fma = MODE == 1 || MODE == 3 || (MODE == 2 && warp >= 4);
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        dmma(c[i][0], c[i][1], a, b);
#pragma unroll
        for (int j = 0; j < NF; ++j) f[(i * NF + j) % 16] = fma(f[(i * NF + j) % 16], a, b);
      }
    } else {
      if (do_mma) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma(c[i][0], c[i][1], a, b);
      }
      if (do_fma) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = fma(f[i], a, b);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
#pragma unroll
  for (int i = 0; i < 16; ++i) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K>
double run(K kern, int blocks, double* d, double flops_per_block) {
  kern<<<blocks, 256>>>(d, 1.0000001, 1e-9);
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0);
    kern<<<blocks, 256>>>(d, 1.0000001, 1e-9);
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return flops_per_block * blocks / (best * 1e-3) / 1e12;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int blocks = p.multiProcessorCount * 2 * 8;
  double* d; CK(cudaMalloc(&d, sizeof(double) * blocks * 256));
  const double mma_w = 8.0 * 512 * ITERS;           // flops per warp per kernel: 8 dmma x 256 fma x 2
  const double fma_w = 64.0 * 32 * 2 * ITERS;       // 64 dfma x 32 lanes x 2
  printf("{\"gpu\": \"%s\",\n", p.name);
  printf(" \"dmma_only_tflops\": %.3f,\n", run(k_mix<0, 0>, blocks, d, 8 * mma_w));
  printf(" \"dfma_only_tflops\": %.3f,\n", run(k_mix<1, 0>, blocks, d, 8 * fma_w));
  printf(" \"split_warps_dmma4_dfma4_tflops\": %.3f,\n", run(k_mix<2, 0>, blocks, d, 4 * mma_w + 4 * fma_w));
  printf(" \"interleave_1dmma_2dfma_tflops\": %.3f,\n", run(k_mix<3, 2>, blocks, d, 8 * (mma_w + 8.0 * 2 * 64 * ITERS)));
  printf(" \"interleave_1dmma_4dfma_tflops\": %.3f,\n", run(k_mix<3, 4>, blocks, d, 8 * (mma_w + 8.0 * 4 * 64 * ITERS)));
  printf(" \"interleave_1dmma_8dfma_tflops\": %.3f\n}\n", run(k_mix<3, 8>, blocks, d, 8 * (mma_w + 8.0 * 8 * 64 * ITERS)));
  return 0;
}
