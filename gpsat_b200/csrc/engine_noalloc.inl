// gpsat_local_select_count_noalloc: the counting pass of gpsat_local_select for callers that must not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Included at the end of engine.cu (one translation unit).
//
// The existing entry point keeps its per-(expert, chunk) match counts in a DevBuf that grows with cudaMalloc.  Here
// that scratch is a file-scope device array of fixed size, resident on every GPU that loads the library until the
// process exits.  Host code must not hand the NAME of the array to a kernel (on the host it is a stand-in symbol, not
// a device address): the device address is looked up once per device with cudaGetSymbolAddress, with that device
// current, and that pointer is what the kernels get.
//
// This function calls no cudaMalloc / cudaFree / cudaMallocHost and no library routine that allocates: inputs and
// outputs are the caller's device buffers, the selection rules travel as a kernel argument, and the launches are the
// same select_count_kernel / select_scan_kernel the allocating entry point uses.
#include <mutex>

namespace gpsat {

// 2^20 ints = 4 MiB.  One int per (expert, 1024-point chunk): e.g. 3 000 experts x 350 000 points.
static __device__ __align__(256) int g_noalloc_chunk_counts[GPSAT_NOALLOC_MAX_CHUNK_COUNTS];

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mutex;
static int* g_noalloc_scratch_ptr[NOALLOC_MAX_DEVICES];      // device address of the array, per device (nullptr = not set up)

// Set-up, once per device, with that device current.
static int noalloc_scratch_for(int device, int** out) {
  REQUIRE(device >= 0 && device < NOALLOC_MAX_DEVICES, "device index out of range for the no-alloc scratch table");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_scratch_ptr[device] == nullptr) {
    CUDA_TRY(cudaSetDevice(device));
    int* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_noalloc_chunk_counts));
    REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a NULL scratch address");
    g_noalloc_scratch_ptr[device] = p;
  }
  *out = g_noalloc_scratch_ptr[device];
  return 0;
}

// true when p is device (or managed) memory of `device`
static int noalloc_check_device_ptr(const void* p, int device, const char* what) {
  cudaPointerAttributes at;
  CUDA_TRY(cudaPointerGetAttributes(&at, p));
  if (!((at.type == cudaMemoryTypeDevice && at.device == device) || at.type == cudaMemoryTypeManaged)) {
    g_last_error = std::string("gpsat: ") + what + " must be a device pointer of the engine's device "
                   "(gpsat_local_select_count_noalloc copies nothing and allocates no staging buffer)";
    return 2;
  }
  return 0;
}

}  // namespace gpsat

extern "C" {

int gpsat_local_select_count_noalloc(gpsat_engine* h, int64_t N, int32_t C, const double* data, int64_t E, int32_t CL,
                                     const double* locs, int32_t ncond, const gpsat_select_cond* conds,
                                     int64_t* counts) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(N >= 1 && E >= 1 && C >= 1 && CL >= 1 && E < 65536, "bad sizes (N >= 1, 1 <= E < 65536)");
  REQUIRE(data && locs && counts, "data / locs / counts is NULL");
  const int64_t nch = (N + SEL_CHUNK - 1) / SEL_CHUNK;
  if (E * nch > (int64_t)GPSAT_NOALLOC_MAX_CHUNK_COUNTS) {
    g_last_error = "gpsat: gpsat_local_select_count_noalloc needs E * ceil(N / 1024) = " + std::to_string(E * nch) +
                   " scratch counters, more than the fixed scratch of GPSAT_NOALLOC_MAX_CHUNK_COUNTS = " +
                   std::to_string((int64_t)GPSAT_NOALLOC_MAX_CHUNK_COUNTS) + "; use gpsat_local_select";
    return 2;
  }
  SelectSpec spec;
  if (int rc = select_spec_from(conds, ncond, C, CL, spec)) return rc;
  Engine& g = h->e;
  int* scratch = nullptr;
  if (int rc = noalloc_scratch_for(g.device, &scratch)) return rc;
  if (int rc = noalloc_check_device_ptr(data, g.device, "data")) return rc;
  if (int rc = noalloc_check_device_ptr(locs, g.device, "locs")) return rc;
  if (int rc = noalloc_check_device_ptr(counts, g.device, "counts")) return rc;
  begin_call(g);
  // Nothing in the scratch has to start at zero: select_count_kernel writes every (expert, chunk) counter that
  // select_scan_kernel reads, so what an earlier call left behind is never seen.
  select_count_kernel<<<dim3((unsigned)nch, (unsigned)E), 256, 0, g.stream>>>(N, C, data, CL, locs, spec, (int)nch, scratch);
  select_scan_kernel<<<(unsigned)((E + 127) / 128), 128, 0, g.stream>>>((int)E, (int)nch, scratch, counts);
  g.stats.kernel_launches += 2;
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

}  // extern "C"
