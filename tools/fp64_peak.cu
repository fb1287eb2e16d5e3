// FP64 pipe micro-benchmark for B200 (sm_100a): vector DFMA vs warp-level DMMA shapes.
// Writes one JSON object to stdout. Used once per pool to fix the FP64 roofline denominator
// (MEASURED_PEAKS.json has no FP64 entry).  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int ITERS = 4096;

__global__ void __launch_bounds__(256) k_dfma(double* out, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// m8n8k4: A 1 reg, B 1 reg, C 2 regs. 256 FMA per warp-instruction.
template <int NACC>
__global__ void __launch_bounds__(256) k_dmma884(double* out, double a, double b) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// m16n8k4: A 2 regs, B 1 reg, C 4 regs. 512 FMA.
template <int NACC>
__global__ void __launch_bounds__(256) k_dmma1684(double* out, double a, double b) {
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3]) : "d"(a), "d"(b), "d"(b));
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// m16n8k8: A 4 regs, B 2 regs, C 4 regs. 1024 FMA.
template <int NACC>
__global__ void __launch_bounds__(256) k_dmma1688(double* out, double a, double b) {
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b));
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// m16n8k16: A 8 regs, B 4 regs, C 4 regs. 2048 FMA.
template <int NACC>
__global__ void __launch_bounds__(256) k_dmma16816(double* out, double a, double b) {
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, "
                   "{%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b),
                     "d"(a), "d"(b), "d"(a), "d"(b));
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  return best;
}

int main(int argc, char** argv) {
  int dev = 0; CK(cudaSetDevice(dev));
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
  int sms = p.multiProcessorCount;
  double* out; CK(cudaMalloc(&out, sizeof(double) * 256 * sms * 64));
  printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
  for (int wpb = 1; wpb <= 8; wpb *= 2) {            // warps per SM scheduler sweep (blocks/SM)
    int blocks = sms * wpb;
    double ms, tf;
    ms = time_ms([&] { k_dfma<<<blocks, 256>>>(out, 1.0000001, 1e-9); }, 5);
    tf = 2.0 * 16 * ITERS * 256.0 * blocks / (ms * 1e-3) / 1e12;
    printf(",\n \"dfma_b%d_tflops\": %.3f", wpb, tf);
    ms = time_ms([&] { k_dmma884<8><<<blocks, 256>>>(out, 1.0000001, 1e-9); }, 5);
    tf = 2.0 * 256 * 8 * ITERS * 8.0 * blocks / (ms * 1e-3) / 1e12;
    printf(",\n \"dmma_m8n8k4_b%d_tflops\": %.3f", wpb, tf);
    ms = time_ms([&] { k_dmma1684<8><<<blocks, 256>>>(out, 1.0000001, 1e-9); }, 5);
    tf = 2.0 * 512 * 8 * ITERS * 8.0 * blocks / (ms * 1e-3) / 1e12;
    printf(",\n \"dmma_m16n8k4_b%d_tflops\": %.3f", wpb, tf);
    ms = time_ms([&] { k_dmma1688<8><<<blocks, 256>>>(out, 1.0000001, 1e-9); }, 5);
    tf = 2.0 * 1024 * 8 * ITERS * 8.0 * blocks / (ms * 1e-3) / 1e12;
    printf(",\n \"dmma_m16n8k8_b%d_tflops\": %.3f", wpb, tf);
    ms = time_ms([&] { k_dmma16816<8><<<blocks, 256>>>(out, 1.0000001, 1e-9); }, 5);
    tf = 2.0 * 2048 * 8 * ITERS * 8.0 * blocks / (ms * 1e-3) / 1e12;
    printf(",\n \"dmma_m16n8k16_b%d_tflops\": %.3f", wpb, tf);
  }
  // dependent-chain latency of DMMA 884 (1 warp, 1 accumulator)
  {
    double ms = time_ms([&] { k_dmma884<1><<<1, 32>>>(out, 1.0000001, 1e-9); }, 5);
    printf(",\n \"dmma_m8n8k4_chain_ns_per_op\": %.2f", ms * 1e6 / ITERS);
    ms = time_ms([&] { k_dmma16816<1><<<1, 32>>>(out, 1.0000001, 1e-9); }, 5);
    printf(",\n \"dmma_m16n8k16_chain_ns_per_op\": %.2f", ms * 1e6 / ITERS);
  }
  // sustained: 3 s of m8n8k4 at 4 blocks/SM
  {
    int blocks = sms * 4;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0));
    int n = 0; float ms = 0;
    while (ms < 3000.f) {
      for (int i = 0; i < 20; ++i) k_dmma884<8><<<blocks, 256>>>(out, 1.0000001, 1e-9);
      n += 20;
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      CK(cudaEventElapsedTime(&ms, e0, e1));
    }
    double tf = 2.0 * 256 * 8 * ITERS * 8.0 * blocks * n / (ms * 1e-3) / 1e12;
    printf(",\n \"dmma_m8n8k4_sustained3s_tflops\": %.3f", tf);
    CK(cudaEventRecord(e0));
    n = 0; ms = 0;
    while (ms < 3000.f) {
      for (int i = 0; i < 20; ++i) k_dfma<<<blocks, 256>>>(out, 1.0000001, 1e-9);
      n += 20;
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      CK(cudaEventElapsedTime(&ms, e0, e1));
    }
    tf = 2.0 * 16 * ITERS * 256.0 * blocks * n / (ms * 1e-3) / 1e12;
    printf(",\n \"dfma_sustained3s_tflops\": %.3f", tf);
  }
  printf("\n}\n");
  CK(cudaFree(out));
  return 0;
}
