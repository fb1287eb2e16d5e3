// Probe: throughput of cp.async.bulk (global -> shared, mbarrier completion) per SM as a function of the request size
// and of the number of requests in flight - the pattern of the producer lane of step_kernel's int8 loops.
// One CTA per SM; one thread keeps a ring of R stages of Z bytes full, every stage waits for its own mbarrier before it
// is reissued (no consumer: the wait is the consumer).  W issuing threads (one per warp, each with its own ring) tell a
// per-thread issue limit from a per-SM one.  Source: a private region of `span` bytes per CTA, walked linearly
// and wrapped (span x 148 above / below the 126 MB of L2 = HBM / L2 resident).  Prints GB/s per SM and us per request.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/bulk_copy_probe tools/bulk_copy_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  for (int spin = 0; spin < (1 << 22); ++spin) {
    uint32_t done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return true;
  }
  return false;
}

__global__ void __launch_bounds__(128, 1) copy_kernel(const uint8_t* __restrict__ src, size_t span, int Z, int R, int nreq, int W,
                                                      long long* __restrict__ cycles, int* __restrict__ err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[64];
  if (threadIdx.x == 0) {
    for (int s = 0; s < R * W; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(&bars[s])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0 && w < W) {
    const int per_span = (int)(span / W / Z);
    const uint8_t* base = src + (size_t)blockIdx.x * span + (size_t)w * per_span * Z;
    uint64_t* bars_w = bars + w * R;
    uint8_t* smem_w = smem + (size_t)w * R * Z;
    nreq /= W;
    const long long t0 = clock64();
    bool ok = true;
    for (int q = 0; q < nreq + R && ok; ++q) {
      const int st = q % R;
      if (q >= R) ok = mbar_wait(smem_u32(&bars_w[st]), ((q / R) - 1) & 1);
      if (q < nreq && ok) {
        const uint32_t bar = smem_u32(&bars_w[st]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" :: "r"(bar), "r"(Z) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                     :: "r"(smem_u32(smem_w + (size_t)st * Z)), "l"(base + (size_t)(q % per_span) * Z), "r"(Z), "r"(bar) : "memory");
      }
    }
    if (w == 0) cycles[blockIdx.x] = clock64() - t0;
    if (!ok) *err = 1;
  }
}

int main() {
  const int grid = 148;
  const size_t span_max = (size_t)16 << 20;
  uint8_t* src; CK(cudaMalloc(&src, span_max * grid)); CK(cudaMemset(src, 1, span_max * grid));
  long long* dc; CK(cudaMalloc(&dc, grid * sizeof(long long)));
  int* derr; CK(cudaMalloc(&derr, 4)); CK(cudaMemset(derr, 0, 4));
  CK(cudaFuncSetAttribute(copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 192 * 1024));
  int clk_khz; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  printf("{\"sm_clock_khz\": %d, \"rows\": [", clk_khz);
  bool first = true;
  const size_t spans[2] = {(size_t)512 << 10, span_max};           // 74 MB in total (L2) / 2.4 GB (HBM)
  for (int sp = 0; sp < 2; ++sp)
    for (int Z = 4096; Z <= 65536; Z *= 2)
      for (int R : {1, 2, 3, 6, 12})
      for (int W : {1, 2, 4}) {
        if ((size_t)Z * R * W > 192 * 1024 || R * W > 64 || (W > 1 && (R == 1 || R == 12))) continue;
        const int nreq = (int)(((size_t)8 << 20) / Z);               // 8 MB per CTA
        long long h[grid];
        for (int rep = 0; rep < 2; ++rep) {
          copy_kernel<<<grid, 128, 192 * 1024>>>(src, spans[sp], Z, R, nreq, W, dc, derr);
          CK(cudaDeviceSynchronize());
        }
        CK(cudaMemcpy(h, dc, sizeof(h), cudaMemcpyDeviceToHost));
        double mean = 0; for (int b = 0; b < grid; ++b) mean += (double)h[b]; mean /= grid;
        const double us = mean / (clk_khz * 1e-3), gbs = (double)nreq * Z / (us * 1e-6) / 1e9;
        printf("%s{\"source\": \"%s\", \"bytes\": %d, \"in_flight_per_thread\": %d, \"threads\": %d, \"gbs_per_sm\": %.1f, \"us_per_request\": %.3f}",
               first ? "" : ", ", sp == 0 ? "l2" : "hbm", Z, R, W, gbs, us / nreq);
        first = false;
      }
  int err; CK(cudaMemcpy(&err, derr, 4, cudaMemcpyDeviceToHost));
  printf("], \"timeout\": %d}\n", err);
  return err;
}
