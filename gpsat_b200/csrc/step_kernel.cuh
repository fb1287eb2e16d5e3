// step_kernel(i): all off-diagonal work of block column i for every expert in flight.
//   blockIdx.x = slot, blockIdx.y = tile within the step (x fastest: the tile-0 CTAs of all slots,
//   which also factor the next diagonal tile, are dispatched first):
//     [0, TR-1-i)            potrf tile (r, i), r = i + 1 + x
//     [TR-1-i, TR-1-i + i)   trtri tile (j, i), j = x - (TR-1-i)      (only when do_inv)
// Phases of one CTA (256 threads, one 128 x 128 output tile):
//   1. main loop   acc = sum_k A_rk B_ik^T           2 x 4 warp grid, 3-stage cp.async pipeline
//   2. C           = K_ri - acc (potrf, covariance evaluated in a rolled per-column loop) | acc
//   3. out         = C W_ii^T, W_ii lower triangular: strip-mapped DMMA over k <= c only
//   4. (potrf)     S_rr -= out out^T, lower triangle only, strip-mapped; coalesced update
//   5. (tile (i+1, i) only)  S_{i+1,i+1} is now final: factor it in place (diag_factor_tile), so the
//                  serial diagonal factorisation overlaps the other tiles of the step
#pragma once
#include "tiles.cuh"
#include "diag_kernel.cuh"

namespace gpsat {

// acc[strip rows][c] = sum_{k <= c} Cs[row][k] * W[c][k];  W (row-major 128 x 128, global) is
// streamed through NWSTAGE B-only stages.  Each warp reads only its own strip of Cs.
__device__ __forceinline__ void wmult_prefetch(const double* Wg, double* wst, int tid) {
#pragma unroll
  for (int s = 0; s < NWSTAGE - 1; ++s) {
    load_slab(wst + s * NB * LDS_S, Wg + s * KC, NB, tid);
    cp_async_commit();
  }
}

// after wmult_prefetch(Wg, wst)
__device__ __forceinline__ void wmult_strip(AccStrip& acc, const double* Cs, const double* Wg, double* wst,
                                            int tid) {
  const int warp = tid >> 5, lane = tid & 31, lr = lane >> 2, lk = lane & 3;
  const int strip = strip_of_warp(warp);
  constexpr int NSL = NB / KC;
  const double* ap = Cs + (16 * strip + lr) * LDC + lk;
#pragma unroll 1
  for (int s = 0; s < NSL; ++s) {
    cp_async_wait<NWSTAGE - 2>();
    __syncthreads();
    const int nx = s + NWSTAGE - 1;
    if (nx < NSL) load_slab(wst + (nx % NWSTAGE) * NB * LDS_S, Wg + nx * KC, NB, tid);
    cp_async_commit();
    const double* bp = wst + (s % NWSTAGE) * NB * LDS_S + lr * LDS_S + lk;
#pragma unroll
    for (int kk = 0; kk < KC / 4; ++kk) {
      const int kb = s * (KC / 4) + kk;
      mma_strip_k4(acc, ap + kb * 4, LDC, bp + kk * 4, LDS_S, kb >> 1, 16);
    }
  }
  cp_async_wait<0>();
}

__global__ void __launch_bounds__(GPSAT_THREADS, 1)
step_kernel(const SlotDesc* __restrict__ slots, int i, int do_inv) {
  const SlotDesc& d = slots[blockIdx.x];
  if (d.expert < 0 || i >= d.T) return;
  const int npot = d.TR - 1 - i;
  const int ninv = do_inv ? i : 0;
  const int x = blockIdx.y;
  if (x >= npot + ninv) return;
  const bool is_pot = x < npot;
  const int rowblk = is_pot ? (i + 1 + x) : (x - npot);   // r (potrf) or j (trtri)

  extern __shared__ double smem[];
  double* Cs = smem;                       // aliases the pipeline stages
  double* wst = smem + SM_TILE_D;
  double* xr = wst + SM_WST_D;
  double* xc = xr + NB * GPSAT_DMAX;
  double* yr = xc + NB * GPSAT_DMAX;
  double* yc = yr + NB;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int lr = lane >> 2, lk = lane & 3;
  const int strip = strip_of_warp(warp);
  const int ld = d.ld;
  const double* Wg = d.Wd + (size_t)i * NB * NB;

  // ---- 1. main loop (slab loads first, the coordinate loads hide behind them) ----
  {
    Acc acc;
    acc_zero(acc);
    const double* Ag = d.A + (size_t)rowblk * NB * ld;
    const double* Bg = d.A + (size_t)i * NB * ld;
    const int kbeg = is_pot ? 0 : rowblk * NB;
    gemm_nt_prologue(Ag, Bg, ld, kbeg, i * NB, smem, tid);
    if (is_pot) {
      load_block_coords(d, rowblk * NB, xr, yr, tid);
      load_block_coords(d, i * NB, xc, yc, tid);
    }
    gemm_nt_mainloop(acc, Ag, Bg, ld, kbeg, i * NB, smem, tid, !is_pot);   // trtri: U_jj is upper triangular
    __syncthreads();   // coords visible; stages free
    wmult_prefetch(Wg, wst, tid);      // W_ii streams in under the epilogue
    // ---- 2. C to shared memory ----
    acc_to_smem(acc, Cs, tid, is_pot ? -1.0 : 1.0);
  }
  __syncthreads();
  if (is_pot) {
    add_aug_tile(d, rowblk * NB, i * NB, xr, xc, yc, Cs, tid, false);
    __syncthreads();
  }

  // ---- 3. out = C * W_ii^T ----
  AccStrip out;
  acc_zero(out);
  wmult_strip(out, Cs, Wg, wst, tid);
  const double sign = is_pot ? 1.0 : -1.0;
  {
    double* Og = d.A + ((size_t)rowblk * NB + 16 * strip + lr) * ld + (size_t)i * NB + 2 * lk;
#pragma unroll
    for (int rf = 0; rf < 2; ++rf)
#pragma unroll
      for (int cf = 0; cf < 16; ++cf)
        *reinterpret_cast<double2*>(Og + (size_t)rf * 8 * ld + cf * 8) =
            make_double2(sign * out.c[rf][cf][0], sign * out.c[rf][cf][1]);
  }
  if (!is_pot || rowblk >= d.T) return;

  // ---- 4. S_rr -= L_ri L_ri^T  (running Schur complement of the diagonal tile of block row r) ----
  // every warp read only its own strip of Cs in phase 3, so it may overwrite that strip now
  {
    double* Cw = Cs + (16 * strip + lr) * LDC + 2 * lk;
#pragma unroll
    for (int rf = 0; rf < 2; ++rf)
#pragma unroll
      for (int cf = 0; cf < 16; ++cf)
        *reinterpret_cast<double2*>(Cw + rf * 8 * LDC + cf * 8) = make_double2(out.c[rf][cf][0], out.c[rf][cf][1]);
  }
  __syncthreads();
  AccStrip sy;
  acc_zero(sy);
  {
    const double* ap = Cs + (16 * strip + lr) * LDC + lk;
    const double* bp = Cs + lr * LDC + lk;
    const int cf_hi = 2 * strip + 2;
#pragma unroll 2
    for (int kb = 0; kb < NB / 4; ++kb) mma_strip_k4(sy, ap + kb * 4, LDC, bp + kb * 4, LDC, 0, cf_hi);
  }
  __syncthreads();   // all warps done reading L_ri from Cs
  {
    double* Cw = Cs + (16 * strip + lr) * LDC + 2 * lk;
    const int cf_hi = 2 * strip + 2;
#pragma unroll
    for (int rf = 0; rf < 2; ++rf)
#pragma unroll
      for (int cf = 0; cf < 16; ++cf)
        if (cf < cf_hi)
          *reinterpret_cast<double2*>(Cw + rf * 8 * LDC + cf * 8) = make_double2(-sy.c[rf][cf][0], -sy.c[rf][cf][1]);
  }
  __syncthreads();
  // the loops below touch element (r, c) from the same thread: no barrier between them
  if (i == 0) add_aug_tile(d, rowblk * NB, rowblk * NB, xr, xr, yr, Cs, tid, true);
  {
    double* Sg = d.A + (size_t)rowblk * NB * ld + (size_t)rowblk * NB;
    const int c = tid & (NB - 1);
    if (x == 0) {
      // ---- 5. rowblk == i + 1: the Schur complement is complete; factor the next diagonal tile ----
      if (i > 0) {
#pragma unroll 16
        for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB)
          if (c <= r) Cs[r * LDC + c] += Sg[(size_t)r * ld + c];
      }
      __syncthreads();
      diag_factor_tile(d, rowblk, Cs, wst, tid);    // scratch: the W stages + coordinate arrays
      return;
    }
    if (i == 0) {
      for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB)
        if (c <= r) Sg[(size_t)r * ld + c] = Cs[r * LDC + c];
    } else {
      // fire-and-forget reduction at L2 (RED.ADD.F64): no load latency on the CTA's critical path.
      // Only this CTA touches S_rr during step i, so the result does not depend on timing.
#pragma unroll 8
      for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB)
        if (c <= r) atomicAdd(Sg + (size_t)r * ld + c, Cs[r * LDC + c]);
    }
  }
}

}  // namespace gpsat
