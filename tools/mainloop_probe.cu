// Probe of the shared NT main loop (128 x 128 output tile, both operands k-contiguous, cp.async ->
// shared memory -> DMMA.8x8x4) in several shapes, to see what fraction of the FP64 peak the loop
// structure itself can reach on a B200.  Prints one JSON object.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/mainloop_probe tools/mainloop_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int NB = 128;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N)); }

// THREADS = 32 * WM * WN; warp tile (128 / WM) x (128 / WN) = (8 FM) x (8 FN)
template <int WM, int WN, int KC, int NST>
__global__ void __launch_bounds__(32 * WM * WN, 1) probe_kernel(const double* __restrict__ A, const double* __restrict__ B,
                                                                double* __restrict__ C, int K, int rtiles) {
  constexpr int THREADS = 32 * WM * WN, FM = NB / WM / 8, FN = NB / WN / 8, LDS = KC + 4, STAGE = 2 * NB * LDS;
  extern __shared__ double smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, wm = warp / WN, wn = warp % WN;
  const int lr = lane >> 2, lk = lane & 3;
  const int rt = blockIdx.x % rtiles, ct = (blockIdx.x / rtiles) % rtiles;
  const double* Ag = A + (size_t)rt * NB * K;
  const double* Bg = B + (size_t)ct * NB * K;
  double acc[FM][FN][2];
#pragma unroll
  for (int i = 0; i < FM; ++i)
#pragma unroll
    for (int j = 0; j < FN; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
  auto load = [&](int stage, int slab) {
    double* st = smem + stage * STAGE;
#pragma unroll
    for (int q = 0; q < (NB * KC / 2) / THREADS; ++q) {
      const int id = tid + q * THREADS, row = id / (KC / 2), ch = id % (KC / 2);
      cp_async16(st + row * LDS + ch * 2, Ag + (size_t)row * K + slab * KC + ch * 2);
      cp_async16(st + NB * LDS + row * LDS + ch * 2, Bg + (size_t)row * K + slab * KC + ch * 2);
    }
  };
  const int nslab = K / KC;
#pragma unroll
  for (int s = 0; s < NST - 1; ++s) { if (s < nslab) load(s, s); cp_async_commit(); }
  for (int s = 0; s < nslab; ++s) {
    cp_async_wait<NST - 2>();
    __syncthreads();
    const int nx = s + NST - 1;
    if (nx < nslab) load(nx % NST, nx);
    cp_async_commit();
    const double* st = smem + (s % NST) * STAGE;
    const double* ap = st + (wm * (NB / WM) + lr) * LDS + lk;
    const double* bp = st + NB * LDS + (wn * (NB / WN) + lr) * LDS + lk;
#pragma unroll
    for (int k4 = 0; k4 < KC / 4; ++k4) {
      double a[FM], b[FN];
#pragma unroll
      for (int f = 0; f < FM; ++f) a[f] = ap[f * 8 * LDS + k4 * 4];
#pragma unroll
      for (int f = 0; f < FN; ++f) b[f] = bp[f * 8 * LDS + k4 * 4];
#pragma unroll
      for (int fm = 0; fm < FM; ++fm)
#pragma unroll
        for (int fn = 0; fn < FN; ++fn) dmma884(acc[fm][fn][0], acc[fm][fn][1], a[fm], b[fn]);
    }
  }
  cp_async_wait<0>();
  double* Cg = C + (size_t)blockIdx.x * NB * NB;
#pragma unroll
  for (int fm = 0; fm < FM; ++fm)
#pragma unroll
    for (int fn = 0; fn < FN; ++fn)
      *reinterpret_cast<double2*>(Cg + (wm * (NB / WM) + fm * 8 + lr) * NB + wn * (NB / WN) + fn * 8 + 2 * lk) =
          make_double2(acc[fm][fn][0], acc[fm][fn][1]);
}

template <int WM, int WN, int KC, int NST>
static double run(const char* name, const double* A, const double* B, double* C, int K, int rtiles, int grid, bool first) {
  constexpr int smem = NST * 2 * NB * (KC + 4) * 8;
  auto kern = probe_kernel<WM, WN, KC, NST>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int w = 0; w < 2; ++w) kern<<<grid, 32 * WM * WN, smem>>>(A, B, C, K, rtiles);
  CK(cudaDeviceSynchronize());
  const int reps = 5;
  CK(cudaEventRecord(e0));
  for (int r = 0; r < reps; ++r) kern<<<grid, 32 * WM * WN, smem>>>(A, B, C, K, rtiles);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double tf = (double)reps * grid * 2.0 * NB * NB * K / (ms * 1e-3) / 1e12;
  printf("%s\"%s\": {\"threads\": %d, \"kc\": %d, \"stages\": %d, \"smem_kb\": %.1f, \"tflops\": %.2f}", first ? "" : ", ",
         name, 32 * WM * WN, KC, NST, smem / 1024.0, tf);
  return tf;
}

int main() {
  const int K = 4096, rtiles = 8, grid = 148 * 4;
  double *A, *B, *C;
  CK(cudaMalloc(&A, (size_t)rtiles * NB * K * 8));
  CK(cudaMalloc(&B, (size_t)rtiles * NB * K * 8));
  CK(cudaMalloc(&C, (size_t)grid * NB * NB * 8));
  CK(cudaMemset(A, 0, (size_t)rtiles * NB * K * 8));
  CK(cudaMemset(B, 0, (size_t)rtiles * NB * K * 8));
  printf("{");
  run<2, 4, 16, 3>("w2x4_kc16_s3", A, B, C, K, rtiles, grid, true);      // the shipped loop
  run<2, 4, 16, 4>("w2x4_kc16_s4", A, B, C, K, rtiles, grid, false);
  run<2, 4, 32, 2>("w2x4_kc32_s2", A, B, C, K, rtiles, grid, false);
  run<2, 4, 32, 3>("w2x4_kc32_s3", A, B, C, K, rtiles, grid, false);
  run<4, 2, 16, 3>("w4x2_kc16_s3", A, B, C, K, rtiles, grid, false);
  run<4, 4, 16, 3>("w4x4_kc16_s3", A, B, C, K, rtiles, grid, false);     // 512 threads, 4 warps per scheduler
  run<4, 4, 32, 3>("w4x4_kc32_s3", A, B, C, K, rtiles, grid, false);
  run<4, 4, 16, 4>("w4x4_kc16_s4", A, B, C, K, rtiles, grid, false);
  printf("}\n");
  return 0;
}
