// K^-1 tile products on the int8 tensor path (tcgen05.mma kind::i8, accumulators in TMEM).
//
// FP64 has no tcgen05 kind; the DMMA path peaks at 37 TFLOP/s while the int8 path of the same SM
// sustains 3.5 POPS at N = 128 (tools/i8_fp64_probe.cu, profiles/r01f_i8_*.json).  The product
// U_i U_j^T of lauum_grad_kernel is therefore evaluated by error-free splitting (Ozaki scheme):
//   U[r][k] = 2^(e_r - 6) * sum_{s < S} q_s[r][k] 2^(-7 s),   q_s int8 in [-64, 64],  S = 7: 48 bits below 2^e_r
//   (U_i U_j^T)[r][c] = 2^(e_r + e_c - 12) * sum_g 2^(-7 g) * sum_{s + t = g} sum_k q_s[r][k] q_t[c][k]
// Every slice product is an exact int32 sum; the groups g = 0..S-1 are recombined in int64 and converted
// once.  Pairs with s + t >= S are dropped (below 2^-48 of the row scales): measured error 4.8e-14 of
// sum |a||b| with S = 7 and 3.9e-16 with S = 8 (tools/i8_fp64_probe.cu), against the 1e-7 the gradient is
// checked to; the LML itself never takes this path.  e_r is the exponent of the largest entry of row r of U
// over the columns the product reads, so the planes of a row can be accumulated over the whole k range in
// TMEM and drained once per pass.
//
// TMEM holds four 128 x 128 int32 accumulators (512 columns), so a tile takes two passes over k:
// groups 0..3 (10 slice pairs, planes 0..3), then groups 4..6 (18 pairs, planes 0..6).
// Planes live in HBM per tile as [32-wide slab][plane][k / 16][row][16 bytes]: the inner three are the layout
// the K-major no-swizzle UMMA descriptor reads directly, and the planes a pass needs of one slab
// (0..3 or 0..6) are one contiguous bulk copy per operand.
#pragma once
#include "lauum_kernel.cuh"

namespace gpsat {

constexpr int I8_S = 7;                          // slices per value
constexpr int I8_PLANE_B = NB * NB;              // bytes per plane and tile
constexpr int I8_SLAB_K = 32;                    // k per pipeline stage = one MMA
constexpr int I8_SLAB_B = NB * I8_SLAB_K;        // bytes per plane and slab
constexpr int I8_NST = 3;
constexpr int I8_SP = 8;                                   // plane capacity per operand of a stage
constexpr int I8_STAGE_B = 2 * I8_SP * I8_SLAB_B;          // A planes then B planes: 64 KB
constexpr int I8_PIPE_B = I8_NST * I8_STAGE_B;             // 192 KB, aliased by the Q tile afterwards
static_assert(I8_PIPE_B >= SM_TILE_D * 8, "Q tile must fit in the pipeline stages");
constexpr int LAUUM_I8_SMEM_BYTES = I8_PIPE_B + (SM_XY_D + 64) * 8 + 128;
constexpr int I8_EXP_CLAMP = 400;

__device__ int g_i8_timeout = 0;                 // set when a bounded barrier wait gives up; every product call reads
                                                 // and clears it (end_call in engine_api.inl) and fails if it is set
__constant__ int g_i8_spin_limit = 1 << 22;       // polls before a wait gives up (gpsat_engine_debug_set_i8_spin_limit)
// cycles per phase of lauum_grad_i8_kernel summed over CTAs (thread 0's clock): prologue, pass-0 MMAs, drain 0,
// pass-1 MMAs, drain 1, Q tile store, gradient reduction, tile count.  Built with -DGPSAT_I8_PHASES only.
__device__ unsigned long long g_i8_phase[8];
#ifdef GPSAT_I8_PHASES
#define I8_PHASE(k) do { if (tid == 0) { long long now_ = clock64(); atomicAdd(&g_i8_phase[k], (unsigned long long)(now_ - tphase)); tphase = now_; } } while (0)
#else
#define I8_PHASE(k) do { } while (0)
#endif

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc_kmajor(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;      // distance between the two 16-byte k chunks of one MMA
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;      // distance between 8-row groups
  d |= (uint64_t)1 << 46;                          // sm_100 descriptor version; layout type 0 = no swizzle
  return d;
}
// D = s32, A = B = signed int8, both K-major, M = 128, N = 128
constexpr uint32_t I8_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
               :: "r"(tmem_d), "l"(a), "l"(b), "r"(I8_IDESC), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" :: "r"(bar) : "memory");
}
// bounded: a wait that never completes sets g_i8_timeout and returns false instead of hanging the GPU
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  // limit <= 0 (test hook): the wait completes normally and is then REPORTED as timed out, so the host-side error
  // path runs end to end without abandoning a pipeline that still has bulk copies and MMAs in flight
  const int limit = g_i8_spin_limit;
  const int polls = limit > 0 ? limit : (1 << 22);
  for (int spin = 0; spin < polls; ++spin) {
    uint32_t done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) {
      if (limit <= 0) g_i8_timeout = 1;
      return true;
    }
  }
  g_i8_timeout = 1;
  return false;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
               :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// one lane of a converged warp (elect.sync): inside `if (warp == w)` the compiler then knows that a single lane runs the
// branch and keeps descriptors / addresses in uniform registers, instead of wrapping every tcgen05.mma in an
// elect / broadcast / branch loop as it does under `if (tid == t)`
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
// 32 TMEM lanes x 16 consecutive 32-bit columns; thread l of the warp receives lane (base + l)
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                 "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
               : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
// ties the registers of an earlier tmem_ld16_nowait to the wait, so that no use is scheduled before it
__device__ __forceinline__ void tmem_ld_fence(uint32_t (&v)[16]) {
  asm volatile("" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                    "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]) :: "memory");
}
// exact int32 -> double without a conversion instruction: 2^52 + 2^31 + v, then subtract the offset
__device__ __forceinline__ double i32_to_f64(uint32_t v) {
  return __hiloint2double(0x43300000, (int)(v ^ 0x80000000u)) - 4503601774854144.0;
}
__device__ __forceinline__ double pow2i(int p) { return __longlong_as_double((long long)(p + 1023) << 52); }

// all slice pairs (s, t) of the groups s + t = G0 .. G1-1 of one 32-wide slab into the accumulators g - G0
template <int G0, int G1>
__device__ __forceinline__ void i8_issue_groups(uint32_t tb, uint64_t adesc, uint64_t bdesc, uint32_t acc0) {
#pragma unroll
  for (int g = G0; g < G1; ++g)
#pragma unroll
    for (int s = 0; s <= g; ++s)
      if (s < G1 && g - s < G1)
        umma_i8(tb + (uint32_t)((g - G0) * NB), adesc + (uint64_t)(s * (I8_SLAB_B >> 4)),
                bdesc + (uint64_t)((g - s) * (I8_SLAB_B >> 4)), s == 0 ? acc0 : 1u);
}

// index of tile (j, k), k >= j, in the packed upper block triangle
__device__ __forceinline__ int utile_index(int T, int j, int k) { return j * T - j * (j - 1) / 2 + (k - j); }
__device__ __forceinline__ int lauum_kend(int n) { return ((n + 1 + KC - 1) / KC) * KC; }

// ---- row scales: e_r with max_k |U[r][k]| < 2^e_r over the columns lauum reads -------------------------
// one warp per row
__global__ void __launch_bounds__(256) u_rowmax_kernel(const SlotDesc* __restrict__ slots) {
  const SlotDesc& d = slots[blockIdx.y];
  if (d.expert < 0) return;
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= d.T * NB) return;
  const int kend = lauum_kend(d.n);
  const double* row = d.A + (size_t)r * d.ld;
  double mx = 0.0;
  for (int k = r + lane; k < kend; k += 32) mx = fmax(mx, fabs(row[k]));    // below the diagonal U is zero
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) {
    int e = 0;
    if (mx > 0.0) e = (int)((__double_as_longlong(mx) >> 52) & 0x7ff) - 1022;
    d.uex[r] = max(-I8_EXP_CLAMP, min(I8_EXP_CLAMP, e));
  }
}

// ---- slicing: one CTA per tile (j, k) of the upper block triangle, two threads per row ----------------
__global__ void __launch_bounds__(256) u_slice_kernel(const SlotDesc* __restrict__ slots) {
  const SlotDesc& d = slots[blockIdx.y];
  if (d.expert < 0) return;
  const int T = d.T, t = blockIdx.x;
  if (t >= T * (T + 1) / 2) return;
  // t -> (j, k): row block j, column block k >= j
  int j = 0, rem = t;
  while (rem >= T - j) { rem -= T - j; ++j; }
  const int k = j + rem;
  const int kend = lauum_kend(d.n);
  const int tid = threadIdx.x, r = tid >> 1, h = tid & 1;
  const double sc = pow2i(6 - d.uex[j * NB + r]);
  // slice s: t = x + 1.5 * 2^(52 - 7 s) rounds x to a multiple of 2^(-7 s) (ties to even); the low mantissa
  // byte of t is q_s in two's complement, t - magic is q_s 2^(-7 s) exactly and the remainder x - that is exact
  double magic[I8_S];
#pragma unroll
  for (int s = 0; s < I8_S; ++s) magic[s] = 1.5 * pow2i(52 - 7 * s);
  const double* row = d.A + (size_t)(j * NB + r) * d.ld + k * NB;
  int8_t* planes = d.Uq + (size_t)t * I8_S * I8_PLANE_B;
#pragma unroll 1
  for (int part = 0; part < 2; ++part) {
    const int c0 = h * 64 + part * 32;
    double x[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) {
      const int c = c0 + q;
      const bool live = (k * NB + c < kend) && (j != k || c >= r);
      x[q] = live ? row[c] * sc : 0.0;
    }
#pragma unroll
    for (int s = 0; s < I8_S; ++s) {
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t w[4];
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          uint32_t pk = 0;
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const int q = ch * 16 + q4 * 4 + b;
            const double tq = x[q] + magic[s];
            x[q] -= tq - magic[s];
            pk |= ((uint32_t)__double2loint(tq) & 0xffu) << (8 * b);
          }
          w[q4] = pk;
        }
        const int c = c0 + ch * 16;
        *reinterpret_cast<uint4*>(planes + (size_t)(c >> 5) * (I8_S * I8_SLAB_B) + s * I8_SLAB_B +
                                  ((c >> 4) & 1) * (NB * 16) + r * 16) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
}

// ---- lauum_grad_i8_kernel: same contract as lauum_grad_kernel, tile product on the int8 path -----------
// planes8 != 0: the operands are the planes step_kernel / diag_factor_tile wrote during the sweep (d.Uq8: 8 planes
// per value of which the first 7 are read, one a-priori scale for all rows, the -alpha column stored as zeros, so
// the tile is Ks^-1 alone) instead of the planes of u_slice_kernel (d.Uq, row scales d.uex).
__global__ void __launch_bounds__(GPSAT_THREADS, 1) lauum_grad_i8_kernel(const SlotDesc* __restrict__ slots, int planes8) {
  const SlotDesc& d = slots[blockIdx.y];
  if (d.expert < 0) return;
  if (planes8 && d.Uq8 == nullptr) return;      // precision rule: this evaluation is lauum_grad_kernel's (FP64 mma.sync)
  const int T = d.T;
  const int t = blockIdx.x;
  if (t >= T * (T + 1) / 2) return;
  int i = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
  while (i * (i + 1) / 2 > t) --i;
  while ((i + 1) * (i + 2) / 2 <= t) ++i;
  const int j = t - i * (i + 1) / 2;
  const bool same = i == j;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* pipe = smem_raw;                                   // I8_NST stages, later the Q tile
  double* Cs = reinterpret_cast<double*>(smem_raw);
  double* xr = reinterpret_cast<double*>(smem_raw + I8_PIPE_B);
  double* xc = xr + NB * GPSAT_DMAX;
  double* ar = xc + NB * GPSAT_DMAX;
  double* ac = ar + NB;
  double* red = ac + NB;
  __shared__ __align__(8) uint64_t bars[2 * I8_NST + 1];            // full[NST], empty[NST], accumulators ready
  __shared__ uint32_t tmem_base;
  __shared__ double csd[NB];                                    // 2^e of the column block rows
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = d.n, ld = d.ld;
#ifdef GPSAT_I8_PHASES
  long long tphase = clock64();
#endif
  const int kend = lauum_kend(n);

  if (tid == 0) {
    for (int b = 0; b < 2 * I8_NST + 1; ++b) mbar_init(smem_u32(&bars[b]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base), 512);
  const int8_t* Ubase = planes8 ? d.Uq8 : d.Uq;
  const size_t tile_b = planes8 ? (size_t)8 * I8_PLANE_B : (size_t)I8_S * I8_PLANE_B;
  const size_t slab_b = planes8 ? (size_t)8 * I8_SLAB_B : (size_t)I8_S * I8_SLAB_B;
  // planes read per value: 7 (2^-49 of the scale) unless the a-priori scale may sit far above the entries of U
  // (|U_rr| = 1 / L_rr >= 1 / sqrt(var + noise) against the bound 1 / sqrt(noise)), then all 8
  const int S = (planes8 && d.var > 255.0 * d.noise) ? 8 : I8_S;   // (6 planes in the low-rho tier: gradient error 8e-8, not adopted)
  int eu = 0;
  if (planes8) {      // a-priori scale of the sweep's planes: |U| <= 1 / sqrt(noise) (urow_exp in i8slice.cuh)
    double b2 = 1.0 / d.noise;
    if (!(b2 > 1e-280)) b2 = 1e-280;
    if (!(b2 < 1e280)) b2 = 1e280;
    eu = ((int)((__double_as_longlong(b2) >> 52) & 0x7ff) - 1022 + 1) >> 1;
  }
  if (tid < NB) csd[tid] = pow2i(planes8 ? eu : d.uex[j * NB + tid]);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[I8_NST]), accb = smem_u32(&bars[2 * I8_NST]);
  const uint32_t pipe_s = smem_u32(pipe);
  const uint64_t desc0 = umma_desc_kmajor(pipe_s, NB * 16, 128);
  const int row = (warp & 3) * 32 + lane, chalf = warp >> 2;        // accumulator row, 64-column half
  const int exrow = planes8 ? eu : d.uex[i * NB + row];
  // slabs of the last k block (columns >= kend hold zero planes and are skipped)
  const int last_slabs = (kend - (T - 1) * NB + I8_SLAB_K - 1) / I8_SLAB_K;
  const int nslabs = (T - 1 - i) * (NB / I8_SLAB_K) + last_slabs;

  double acc[64];
#pragma unroll
  for (int q = 0; q < 64; ++q) acc[q] = 0.0;
  I8_PHASE(0);
  int it = 0;                                     // slabs issued so far (producer and MMA thread count alike)
  bool ok = true;
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    const int np = pass == 0 ? 4 : S;                         // planes per operand this pass
    const int g0 = pass * 4, g1 = min(g0 + 4, S);             // groups s + t = g of this pass
    if (warp == 0) {
      // ---- producer: bulk copies of the slab planes ----
      int itp = it;
      if (elect_one())
      for (int kb = i; kb < T && ok; ++kb) {
        const int8_t* ga = Ubase + (size_t)utile_index(T, i, kb) * tile_b;
        const int8_t* gb = Ubase + (size_t)utile_index(T, j, kb) * tile_b;
        const int nsl = kb == T - 1 ? last_slabs : NB / I8_SLAB_K;
        for (int sl = 0; sl < nsl; ++sl, ++itp) {
          const int st = itp % I8_NST;
          if (itp >= I8_NST) ok = ok && mbar_wait(empty0 + 8 * st, ((itp / I8_NST) - 1) & 1);
          if (!ok) break;
          const uint32_t fb = full0 + 8 * st, sbase = pipe_s + st * I8_STAGE_B;
          mbar_expect_tx(fb, (same ? np : 2 * np) * I8_SLAB_B);
          bulk_g2s(sbase, ga + (size_t)sl * slab_b, np * I8_SLAB_B, fb);
          if (!same) bulk_g2s(sbase + I8_SP * I8_SLAB_B, gb + (size_t)sl * slab_b, np * I8_SLAB_B, fb);
        }
      }
    } else if (warp == 1 && elect_one()) {
      // ---- MMA issue: all slice pairs of the pass for every slab (one elected lane of a converged warp, see elect_one) ----
      int itm = it;
      bool first = true;
      for (int kb = i; kb < T && ok; ++kb) {
        const int nsl = kb == T - 1 ? last_slabs : NB / I8_SLAB_K;
        for (int sl = 0; sl < nsl; ++sl, ++itm) {
          const int st = itm % I8_NST;
          ok = ok && mbar_wait(full0 + 8 * st, (itm / I8_NST) & 1);
          if (!ok) break;
          tc_fence_after();
          // descriptors differ only in the start address field (units of 16 bytes, low word)
          const uint64_t adesc = desc0 + (uint64_t)((st * I8_STAGE_B) >> 4);
          const uint64_t bdesc = same ? adesc : adesc + (uint64_t)((I8_SP * I8_SLAB_B) >> 4);
          const uint32_t acc0 = first ? 0u : 1u;
          // straight-line issue per slab: the plane count is resolved outside (a run-time test in front of each
          // tcgen05.mma costs the issuing thread about 20 cycles, see i8_issue_pass in step_kernel.cuh)
          if (pass == 0) i8_issue_groups<0, 4>(tb, adesc, bdesc, acc0);
          else if (S == 8) i8_issue_groups<4, 8>(tb, adesc, bdesc, acc0);
          else i8_issue_groups<4, I8_S>(tb, adesc, bdesc, acc0);
          umma_commit(empty0 + 8 * st);
          first = false;
        }
      }
      umma_commit(accb);
      ok = mbar_wait(accb, pass) && ok;       // only this thread polls; the others sleep in the barrier below
    } else if (pass == 0 && tid >= 64) {
      // warps 2..7 are idle during the loop: they fetch what the gradient reduction needs (coordinates, alpha)
      constexpr int NT = GPSAT_THREADS - 64;
      load_block_coords(d, i * NB, xr, ar, tid - 64, NT);      // ar / ac overwritten with alpha by the same threads
      load_block_coords(d, j * NB, xc, ac, tid - 64, NT);
      for (int r = tid - 64; r < NB; r += NT) {
        int gi = i * NB + r, gj = j * NB + r;
        ar[r] = (gi < n) ? -d.A[(size_t)gi * ld + n] : 0.0;
        ac[r] = (gj < n) ? -d.A[(size_t)gj * ld + n] : 0.0;
      }
    }
    it += nslabs;
    __syncwarp();
    ok = __all_sync(0xffffffffu, ok);         // warp-uniform: tcgen05.ld below is warp-collective
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    I8_PHASE(1 + 2 * pass);
    if (ok) {
      // drain: sum_g G_g 2^(-7 g) in int64 (exact), one conversion per element
      const double rs = pow2i(exrow - 12 - 7 * (g1 - 1));
#pragma unroll
      for (int c = 0; c < 64; c += 16) {
        // all groups of the chunk in flight before one wait (the loads are latency-, not bandwidth-bound)
        uint32_t v[4][16];
#pragma unroll
        for (int g = 0; g < 4; ++g)
          if (g < g1 - g0)
            tmem_ld16_nowait(tb + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(g * NB + chalf * 64 + c), v[g]);
        tmem_ld_wait();
        // Horner over the groups in FP64: every partial sum is an integer below 2^53, so this is exact
        double sum[16];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if (g < g1 - g0) {
            tmem_ld_fence(v[g]);
#pragma unroll
            for (int q = 0; q < 16; ++q) sum[q] = g == 0 ? i32_to_f64(v[g][q]) : fma(sum[q], 128.0, i32_to_f64(v[g][q]));
          }
        }
#pragma unroll
        for (int q = 0; q < 16; ++q) acc[c + q] = fma(sum[q] * rs, csd[chalf * 64 + c + q], acc[c + q]);
      }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    I8_PHASE(2 + 2 * pass);
  }
  // Q tile (Ks^-1 + alpha alpha^T) into shared memory; the pipeline stages are idle now
  {
    double* crow = Cs + row * LDC + chalf * 64;
#pragma unroll
    for (int q = 0; q < 64; q += 2) *reinterpret_cast<double2*>(crow + q) = make_double2(acc[q], acc[q + 1]);
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
  I8_PHASE(5);

  double part[GPSAT_DMAX + 2];
  grad_tile(d, i * NB, j * NB, xr, xc, ar, ac, Cs, tid, part, planes8 ? 0.0 : 1.0);
#pragma unroll
  for (int p = 0; p < GPSAT_DMAX + 2; ++p) {
    double v = part[p];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp * 8 + p] = v;
  }
  __syncthreads();
  if (tid < GPSAT_DMAX + 2) {
    double v = 0.0;
    for (int w8 = 0; w8 < 8; ++w8) v += red[w8 * 8 + tid];
    d.gpart[(size_t)t * (GPSAT_DMAX + 2) + tid] = v;
  }
  I8_PHASE(6);
#ifdef GPSAT_I8_PHASES
  if (tid == 0) atomicAdd(&g_i8_phase[7], 1ull);
#endif
}

// ---- persistent variant: warp-specialised, gradient reduction from registers -----------------------------
// One CTA per SM walks the (slot, tile) list.  Warp 8 streams the slab planes, warp 9 issues the MMAs, warps
// 0..7 drain the accumulators and reduce the tile against dK/dtheta straight from registers, so the
// covariance-derivative epilogue of tile t runs while the tensor pipe works on tile t + 1.
constexpr int I8P_THREADS = 320;
constexpr int I8P_EPI = 256;
constexpr int LAUUM_I8P_SMEM_BYTES = I8_PIPE_B + (SM_XY_D + 64) * 8 + 128;

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;\n" ::: "memory"); }

__device__ __forceinline__ void tmem_ld8_nowait(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_fence8(uint32_t (&v)[8]) {
  asm volatile("" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]) :: "memory");
}

// work item w -> (slot, tile); false when the item is empty
__device__ __forceinline__ bool i8p_item(const SlotDesc* __restrict__ slots, int w, int ntile_max, int& slot, int& t) {
  slot = w / ntile_max;
  t = w - slot * ntile_max;
  const SlotDesc& d = slots[slot];
  return d.expert >= 0 && t < d.T * (d.T + 1) / 2;
}
__device__ __forceinline__ void i8p_tile_ij(int t, int& i, int& j) {
  i = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
  while (i * (i + 1) / 2 > t) --i;
  while ((i + 1) * (i + 2) / 2 <= t) ++i;
  j = t - i * (i + 1) / 2;
}

// Gradient partial sums of one row of the Q tile held in registers: row gi, columns g0c + c0 .. + 63.
// Same arithmetic as grad_tile_t, mapped one row per thread (column data are warp-wide broadcasts).
template <int KIND, int DD>
__device__ __forceinline__ void grad_row_t(const SlotDesc& d, int gi, int g0r, int g0c, int c0, const double* xrow,
                                           double arv, const double* xc, const double* ac, const double (&acc)[64],
                                           double* part) {
  const int n = d.n, D = d.D;
  constexpr int ND = DD > 0 ? DD : GPSAT_DMAX;
  double pd[ND], pk = 0.0, pn = 0.0;
#pragma unroll
  for (int dd = 0; dd < ND; ++dd) pd[dd] = 0.0;
  const bool interior = g0r + NB <= n && g0r > g0c;
  if (gi < n) {
#pragma unroll
    for (int q = 0; q < 64; ++q) {
      const int c = c0 + q, gj = g0c + c;
      if (interior || gj <= gi) {
        const double aa = arv * ac[c];
        const double qv = aa - (acc[q] - aa);          // the tile holds Ks^-1 + alpha alpha^T
        const double wq = (gi == gj) ? qv : 2.0 * qv;
        double sd[ND], r2 = 0.0;
#pragma unroll
        for (int dd = 0; dd < ND; ++dd) {
          const double tdd = (DD > 0 || dd < D) ? xrow[dd] - xc[dd * NB + c] : 0.0;
          sd[dd] = tdd * tdd;
          r2 += sd[dd];
        }
        double ku, gu;
        kunit_gunit_of_r2<KIND>(r2, ku, gu);
        const double wg = wq * gu;
#pragma unroll
        for (int dd = 0; dd < ND; ++dd) pd[dd] = fma(wg, sd[dd], pd[dd]);
        pk = fma(wq, ku, pk);
        if (gi == gj) pn += qv;
      }
    }
  }
#pragma unroll
  for (int dd = 0; dd < GPSAT_DMAX; ++dd) part[dd] = dd < ND ? pd[dd < ND ? dd : 0] : 0.0;
  part[GPSAT_DMAX] = pk;
  part[GPSAT_DMAX + 1] = pn;
}

template <int KIND, int DD>
__global__ void __launch_bounds__(I8P_THREADS, 1) lauum_grad_i8p_kernel(const SlotDesc* __restrict__ slots, int S,
                                                                       int ntile_max) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* xr = reinterpret_cast<double*>(smem_raw + I8_PIPE_B);
  double* xc = xr + NB * GPSAT_DMAX;
  double* ar = xc + NB * GPSAT_DMAX;
  double* ac = ar + NB;
  double* red = ac + NB;
  // full[NST], empty[NST], accumulators ready (MMA -> epilogue), accumulators drained (epilogue -> MMA)
  __shared__ __align__(8) uint64_t bars[2 * I8_NST + 2];
  __shared__ uint32_t tmem_base;
  __shared__ double csd[NB];                                    // 2^e of the column block rows
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int b = 0; b < 2 * I8_NST + 1; ++b) mbar_init(smem_u32(&bars[b]), 1);
    mbar_init(smem_u32(&bars[2 * I8_NST + 1]), I8P_EPI / 32);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[I8_NST]);
  const uint32_t accfull = smem_u32(&bars[2 * I8_NST]), accfree = smem_u32(&bars[2 * I8_NST + 1]);
  const uint32_t pipe_s = smem_u32(smem_raw);
  const int W = S * ntile_max;
  bool ok = true;

  if (warp == 8) {
    // ---- producer ----
    if (lane == 0) {
      int it = 0;
      for (int w = blockIdx.x; w < W && ok; w += gridDim.x) {
        int slot, t, i, j;
        if (!i8p_item(slots, w, ntile_max, slot, t)) continue;
        const SlotDesc& d = slots[slot];
        i8p_tile_ij(t, i, j);
        const int T = d.T;
        const bool same = i == j;
        const int last_slabs = (lauum_kend(d.n) - (T - 1) * NB + I8_SLAB_K - 1) / I8_SLAB_K;
        for (int pass = 0; pass < 2 && ok; ++pass) {
          const int np = pass == 0 ? 4 : I8_S;
          for (int kb = i; kb < T && ok; ++kb) {
            const int8_t* ga = d.Uq + (size_t)utile_index(T, i, kb) * I8_S * I8_PLANE_B;
            const int8_t* gb = d.Uq + (size_t)utile_index(T, j, kb) * I8_S * I8_PLANE_B;
            const int nsl = kb == T - 1 ? last_slabs : NB / I8_SLAB_K;
            for (int sl = 0; sl < nsl; ++sl, ++it) {
              const int st = it % I8_NST;
              if (it >= I8_NST) ok = ok && mbar_wait(empty0 + 8 * st, ((it / I8_NST) - 1) & 1);
              if (!ok) break;
              const uint32_t fb = full0 + 8 * st, sbase = pipe_s + st * I8_STAGE_B;
              mbar_expect_tx(fb, (same ? np : 2 * np) * I8_SLAB_B);
              bulk_g2s(sbase, ga + (size_t)sl * (I8_S * I8_SLAB_B), np * I8_SLAB_B, fb);
              if (!same) bulk_g2s(sbase + I8_S * I8_SLAB_B, gb + (size_t)sl * (I8_S * I8_SLAB_B), np * I8_SLAB_B, fb);
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ---- MMA issue ----
    if (lane == 0) {
      const uint64_t desc0 = umma_desc_kmajor(pipe_s, NB * 16, 128);
      int it = 0, npass = 0;
      for (int w = blockIdx.x; w < W && ok; w += gridDim.x) {
        int slot, t, i, j;
        if (!i8p_item(slots, w, ntile_max, slot, t)) continue;
        const SlotDesc& d = slots[slot];
        i8p_tile_ij(t, i, j);
        const int T = d.T;
        const bool same = i == j;
        const int last_slabs = (lauum_kend(d.n) - (T - 1) * NB + I8_SLAB_K - 1) / I8_SLAB_K;
        const int nslabs = (T - 1 - i) * (NB / I8_SLAB_K) + last_slabs;
        for (int pass = 0; pass < 2 && ok; ++pass, ++npass) {
          if (npass > 0) ok = ok && mbar_wait(accfree, (npass - 1) & 1);      // previous pass drained
          tc_fence_after();
          for (int sl = 0; sl < nslabs && ok; ++sl, ++it) {
            const int st = it % I8_NST;
            ok = ok && mbar_wait(full0 + 8 * st, (it / I8_NST) & 1);
            if (!ok) break;
            tc_fence_after();
            const uint64_t adesc = desc0 + (uint64_t)((st * I8_STAGE_B) >> 4);
            const uint64_t bdesc = same ? adesc : adesc + (uint64_t)((I8_S * I8_SLAB_B) >> 4);
            const uint32_t acc0 = sl == 0 ? 0u : 1u;
            if (pass == 0) {
#pragma unroll
              for (int g = 0; g < 4; ++g)
#pragma unroll
                for (int s = 0; s <= g; ++s)
                  umma_i8(tb + (uint32_t)(g * NB), adesc + (uint64_t)(s * (I8_SLAB_B >> 4)),
                          bdesc + (uint64_t)((g - s) * (I8_SLAB_B >> 4)), s == 0 ? acc0 : 1u);
            } else {
#pragma unroll
              for (int g = 4; g < I8_S; ++g)
#pragma unroll
                for (int s = 0; s <= g; ++s)
                  if (s <= I8_S - 1 && g - s <= I8_S - 1)
                    umma_i8(tb + (uint32_t)((g - 4) * NB), adesc + (uint64_t)(s * (I8_SLAB_B >> 4)),
                            bdesc + (uint64_t)((g - s) * (I8_SLAB_B >> 4)), s == 0 ? acc0 : 1u);
            }
            umma_commit(empty0 + 8 * st);
          }
          umma_commit(accfull);
        }
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue warps 0..7: drain both passes, reduce against dK/dtheta from registers ----
    const int row = (warp & 3) * 32 + lane, chalf = warp >> 2;
    const uint32_t trow = tb + ((uint32_t)((warp & 3) * 32) << 16);
    int npass = 0;
    for (int w = blockIdx.x; w < W; w += gridDim.x) {
      int slot, t, i, j;
      if (!i8p_item(slots, w, ntile_max, slot, t)) continue;
      const SlotDesc& d = slots[slot];
      i8p_tile_ij(t, i, j);
      const int n = d.n, ld = d.ld;
      epi_bar_sync();                                  // the previous tile is done with xr / xc / ar / ac / red
      load_block_coords(d, i * NB, xr, ar, tid);
      load_block_coords(d, j * NB, xc, ac, tid);
      if (tid < NB) csd[tid] = pow2i(d.uex[j * NB + tid]);
      epi_bar_sync();
      if (tid < NB) {
        const int gi = i * NB + tid, gj = j * NB + tid;
        ar[tid] = (gi < n) ? -d.A[(size_t)gi * ld + n] : 0.0;
        ac[tid] = (gj < n) ? -d.A[(size_t)gj * ld + n] : 0.0;
      }
      epi_bar_sync();
      double xrow[GPSAT_DMAX];
#pragma unroll
      for (int dd = 0; dd < GPSAT_DMAX; ++dd) xrow[dd] = dd < d.D ? xr[dd * NB + row] : 0.0;
      const double arv = ar[row];
      const int exrow = d.uex[i * NB + row];
      double acc[64];
#pragma unroll
      for (int q = 0; q < 64; ++q) acc[q] = 0.0;
#pragma unroll 1
      for (int pass = 0; pass < 2; ++pass, ++npass) {
        const int ng = pass == 0 ? 4 : I8_S - 4;
        if (tid == 0) ok = ok && mbar_wait(accfull, npass & 1);     // one poller; the other warps sleep in the barrier
        epi_bar_sync();
        tc_fence_after();
        {
          const double rs = pow2i(exrow - 12 - 7 * (pass * 4 + ng - 1));
#pragma unroll
          for (int c = 0; c < 64; c += 8) {
            uint32_t v[4][8];
#pragma unroll
            for (int g = 0; g < 4; ++g)
              if (g < ng) tmem_ld8_nowait(trow + (uint32_t)(g * NB + chalf * 64 + c), v[g]);
            tmem_ld_wait();
            double sum[8];
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (g < ng) {
                tmem_ld_fence8(v[g]);
#pragma unroll
                for (int q = 0; q < 8; ++q) sum[q] = g == 0 ? i32_to_f64(v[g][q]) : fma(sum[q], 128.0, i32_to_f64(v[g][q]));
              }
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[c + q] = fma(sum[q] * rs, csd[chalf * 64 + c + q], acc[c + q]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(accfree);
      }
      double part[GPSAT_DMAX + 2];
      const int gi = i * NB + row;
      grad_row_t<KIND, DD>(d, gi, i * NB, j * NB, chalf * 64, xrow, arv, xc, ac, acc, part);
#pragma unroll
      for (int p = 0; p < GPSAT_DMAX + 2; ++p) {
        double v = part[p];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[warp * 8 + p] = v;
      }
      epi_bar_sync();
      if (tid < GPSAT_DMAX + 2) {
        double v = 0.0;
        for (int w8 = 0; w8 < 8; ++w8) v += red[w8 * 8 + tid];
        d.gpart[(size_t)t * (GPSAT_DMAX + 2) + tid] = v;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}

}  // namespace gpsat
