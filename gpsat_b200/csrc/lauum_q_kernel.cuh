// lauum_grad_i8q_kernel: persistent, warp-specialised form of lauum_grad_i8_kernel(planes8 = 1) (i8mma.cuh).
//
// The one-CTA-per-tile kernel spends 35 us per tile at c2 size: 18.7 us of int8 MMAs, 5.4 us of drains and 9.3 us of
// FP64 gradient reduction during which the tensor pipe idles.  Here one CTA per SM walks the (slot, tile) list:
//   warp 8  streams the slab planes (cp.async.bulk into a 3-stage mbarrier ring),
//   warp 9  issues the tcgen05.mma instructions and commits the accumulators,
//   warps 0..7 drain the accumulators (tcgen05.ld + exact FP64 Horner), park the Q tile in a per-CTA scratch tile in
//           global memory (135 KB per SM: it stays in L2) and run the SAME rolled grad_tile reduction on it -
// so the MMAs of tile t + 1 run while tile t is reduced.  (The earlier persistent variant, lauum_grad_i8p_kernel, reduced
// straight from registers; its fully unrolled 64-column epilogue was instruction-fetch bound.)  The arithmetic is that
// of lauum_grad_i8_kernel operation for operation, so the gradient partial sums are bitwise identical
// (tests/test_gpu_parity.py::test_persistent_lauum_is_bitwise_equal).
// Measured at c2 (profiles/README.md r02h): 149.6 ms of LAUUM per bench step against 134.4 ms for the one-CTA-per-tile kernel -
// the round trip of the Q tile through L2 (uncoalesced 16-byte row stores, reads at L2 latency) keeps the eight drain /
// reduce warps busy for ~22 us per tile, longer than the 7.5 us of pass-0 MMAs they are meant to hide behind, so the MMA
// warp waits for them.  Not the default; selectable with GPSAT_B200_LAUUM=i8q / GPSAT_LAUUM_I8Q.
#pragma once
#include "i8slice.cuh"

namespace gpsat {

constexpr int I8Q_THREADS = 320;
constexpr int I8Q_EPI = 256;
constexpr int LAUUM_I8Q_SMEM_BYTES = I8_PIPE_B + (SM_XY_D + 64) * 8 + 128;
constexpr size_t I8Q_SCRATCH_D = (size_t)NB * LDC;          // doubles of scratch per CTA

// work item w -> (slot, tile); false when the item is empty or the slot's evaluation is not on the planes
__device__ __forceinline__ bool i8q_item(const SlotDesc* __restrict__ slots, int w, int ntile_max, int& slot, int& t) {
  slot = w / ntile_max;
  t = w - slot * ntile_max;
  const SlotDesc& d = slots[slot];
  return d.expert >= 0 && d.Uq8 != nullptr && t < d.T * (d.T + 1) / 2;
}
// planes read per value (see lauum_grad_i8_kernel): 7, or all 8 when the a-priori scale may sit far above the entries
__device__ __forceinline__ int i8q_planes(const SlotDesc& d) { return d.var > 255.0 * d.noise ? 8 : I8_S; }

__global__ void __launch_bounds__(I8Q_THREADS, 1)
lauum_grad_i8q_kernel(const SlotDesc* __restrict__ slots, int S, int ntile_max, double* __restrict__ qscratch) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* xr = reinterpret_cast<double*>(smem_raw + I8_PIPE_B);
  double* xc = xr + NB * GPSAT_DMAX;
  double* ar = xc + NB * GPSAT_DMAX;
  double* ac = ar + NB;
  double* red = ac + NB;
  // full[NST], empty[NST], accumulators ready (MMA -> drain warps), accumulators drained (drain warps -> MMA)
  __shared__ __align__(8) uint64_t bars[2 * I8_NST + 2];
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int b = 0; b < 2 * I8_NST + 1; ++b) mbar_init(smem_u32(&bars[b]), 1);
    mbar_init(smem_u32(&bars[2 * I8_NST + 1]), I8Q_EPI / 32);
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
  constexpr size_t TILE_B = (size_t)8 * I8_PLANE_B;          // planes of one tile in d.Uq8 (8 per value)
  constexpr size_t SLAB_B = (size_t)8 * I8_SLAB_B;           // planes of one 32-wide slab
  bool ok = true;

  if (warp == 8) {
    // ---- producer ----
    if (elect_one()) {
      int it = 0;
      for (int w = blockIdx.x; w < W && ok; w += gridDim.x) {
        int slot, t, i, j;
        if (!i8q_item(slots, w, ntile_max, slot, t)) continue;
        const SlotDesc& d = slots[slot];
        i8p_tile_ij(t, i, j);
        const int T = d.T, SP = i8q_planes(d);
        const bool same = i == j;
        const int last_slabs = (lauum_kend(d.n) - (T - 1) * NB + I8_SLAB_K - 1) / I8_SLAB_K;
        for (int pass = 0; pass < 2 && ok; ++pass) {
          const int np = pass == 0 ? 4 : SP;
          for (int kb = i; kb < T && ok; ++kb) {
            const int8_t* ga = d.Uq8 + (size_t)utile_index(T, i, kb) * TILE_B;
            const int8_t* gb = d.Uq8 + (size_t)utile_index(T, j, kb) * TILE_B;
            const int nsl = kb == T - 1 ? last_slabs : NB / I8_SLAB_K;
            for (int sl = 0; sl < nsl; ++sl, ++it) {
              const int st = it % I8_NST;
              if (it >= I8_NST) ok = ok && mbar_wait(empty0 + 8 * st, ((it / I8_NST) - 1) & 1);
              if (!ok) break;
              const uint32_t fb = full0 + 8 * st, sbase = pipe_s + st * I8_STAGE_B;
              mbar_expect_tx(fb, (same ? np : 2 * np) * I8_SLAB_B);
              bulk_g2s(sbase, ga + (size_t)sl * SLAB_B, np * I8_SLAB_B, fb);
              if (!same) bulk_g2s(sbase + I8_SP * I8_SLAB_B, gb + (size_t)sl * SLAB_B, np * I8_SLAB_B, fb);
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ---- MMA issue ----
    if (elect_one()) {
      const uint64_t desc0 = umma_desc_kmajor(pipe_s, NB * 16, 128);
      int it = 0, npass = 0;
      for (int w = blockIdx.x; w < W && ok; w += gridDim.x) {
        int slot, t, i, j;
        if (!i8q_item(slots, w, ntile_max, slot, t)) continue;
        const SlotDesc& d = slots[slot];
        i8p_tile_ij(t, i, j);
        const int T = d.T, SP = i8q_planes(d);
        const bool same = i == j;
        const int last_slabs = (lauum_kend(d.n) - (T - 1) * NB + I8_SLAB_K - 1) / I8_SLAB_K;
        const int nslabs = (T - 1 - i) * (NB / I8_SLAB_K) + last_slabs;
        for (int pass = 0; pass < 2 && ok; ++pass, ++npass) {
          if (npass > 0) ok = ok && mbar_wait(accfree, (npass - 1) & 1);      // the previous pass has been drained
          tc_fence_after();
          for (int sl = 0; sl < nslabs && ok; ++sl, ++it) {
            const int st = it % I8_NST;
            ok = ok && mbar_wait(full0 + 8 * st, (it / I8_NST) & 1);
            if (!ok) break;
            tc_fence_after();
            const uint64_t adesc = desc0 + (uint64_t)((st * I8_STAGE_B) >> 4);
            const uint64_t bdesc = same ? adesc : adesc + (uint64_t)((I8_SP * I8_SLAB_B) >> 4);
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
              for (int g = 4; g < I8_SP; ++g)
#pragma unroll
                for (int s = 0; s <= g; ++s)
                  if (g < SP && s < SP && g - s < SP)
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
    // ---- warps 0..7: drain both passes, park the Q tile, reduce it against dK/dtheta ----
    const int row = (warp & 3) * 32 + lane, chalf = warp >> 2;
    const uint32_t trow = tb + ((uint32_t)((warp & 3) * 32) << 16);
    double* Q = qscratch + (size_t)blockIdx.x * I8Q_SCRATCH_D;
    int npass = 0;
    for (int w = blockIdx.x; w < W; w += gridDim.x) {
      int slot, t, i, j;
      if (!i8q_item(slots, w, ntile_max, slot, t)) continue;
      const SlotDesc& d = slots[slot];
      i8p_tile_ij(t, i, j);
      const int n = d.n, ld = d.ld, SP = i8q_planes(d);
      epi_bar_sync();                                  // the previous tile is done with xr / xc / ar / ac / red / Q
      load_block_coords(d, i * NB, xr, ar, tid, I8Q_EPI);
      load_block_coords(d, j * NB, xc, ac, tid, I8Q_EPI);
      epi_bar_sync();
      if (tid < NB) {                                  // ar / ac: alpha of the row / column block (column n of U is -alpha)
        const int gi = i * NB + tid, gj = j * NB + tid;
        ar[tid] = (gi < n) ? -d.A[(size_t)gi * ld + n] : 0.0;
        ac[tid] = (gj < n) ? -d.A[(size_t)gj * ld + n] : 0.0;
      }
      // a-priori scale of the sweep's planes: |U| <= 1 / sqrt(noise) (urow_exp in i8slice.cuh)
      const int eu = urow_exp(d);
      const double cs = pow2i(eu);
      double acc[64];
#pragma unroll
      for (int q = 0; q < 64; ++q) acc[q] = 0.0;
#pragma unroll 1
      for (int pass = 0; pass < 2; ++pass, ++npass) {
        const int ng = pass == 0 ? 4 : SP - 4;
        if (tid == 0) ok = ok && mbar_wait(accfull, npass & 1);     // one poller; the other warps sleep in the barrier
        epi_bar_sync();
        tc_fence_after();
        {
          const double rs = pow2i(eu - 12 - 7 * (pass * 4 + ng - 1));
#pragma unroll
          for (int c = 0; c < 64; c += 16) {
            uint32_t v[4][16];
#pragma unroll
            for (int g = 0; g < 4; ++g)
              if (g < ng) tmem_ld16_nowait(trow + (uint32_t)(g * NB + chalf * 64 + c), v[g]);
            tmem_ld_wait();
            double sum[16];
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (g < ng) {
                tmem_ld_fence(v[g]);
#pragma unroll
                for (int q = 0; q < 16; ++q) sum[q] = g == 0 ? i32_to_f64(v[g][q]) : fma(sum[q], 128.0, i32_to_f64(v[g][q]));
              }
            }
#pragma unroll
            for (int q = 0; q < 16; ++q) acc[c + q] = fma(sum[q] * rs, cs, acc[c + q]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(accfree);           // the MMA warp may overwrite the accumulators
      }
      {
        double* qrow = Q + (size_t)row * LDC + chalf * 64;
#pragma unroll
        for (int q = 0; q < 64; q += 2) *reinterpret_cast<double2*>(qrow + q) = make_double2(acc[q], acc[q + 1]);
      }
      epi_bar_sync();                                  // Q, ar, ac visible to the 256 reducing threads
      double part[GPSAT_DMAX + 2];
      grad_tile(d, i * NB, j * NB, xr, xc, ar, ac, Q, tid, part, 0.0);
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
