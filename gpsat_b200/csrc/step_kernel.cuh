// step_kernel(i): all off-diagonal work of block column i for every expert in flight.
//   x = tile within the step (see the kernel for the CTA -> (slot, x) map):
//     [0, TR-1-i)            potrf tile (r, i), r = i + 1 + x
//     [TR-1-i, TR-1-i + i)   trtri tile (j, i), j = x - (TR-1-i)      (only when do_inv)
// Phases of one CTA (256 threads, one 128 x 128 output tile):
//   1. main loop   acc = sum_k A_rk B_ik^T           2 x 4 warp grid, 3-stage cp.async pipeline
//   2. C           = K_ri - acc (potrf, covariance evaluated in a rolled per-column loop) | acc
//   3. out         = C W_ii^T, W_ii lower triangular: strip-mapped DMMA over k <= c only
//   4. (potrf)     S_rr -= out out^T, lower triangle only, strip-mapped; coalesced update
//   5. (tile (i+1, i) only)  S_{i+1,i+1} is now final: factor it in place (diag_factor_tile), so the
//                  serial diagonal factorisation overlaps the other tiles of the step
//
// int8 tensor path for the potrf main loop (slots with d.Lq != nullptr): every finished tile L_ri is also
// stored as 8 exact int8 slice planes (same error-free splitting and plane layout as i8mma.cuh) with an
// A-PRIORI row scale - row r of the augmented factor satisfies |L_rk| <= |row r| <= sqrt(K_rr), known before
// the factorisation - so that later tiles accumulate sum_k L_rk L_ik^T over the whole k range in TMEM
// (tcgen05.mma kind::i8) and drain once per pass.  Two passes: slice-pair groups g = s + t = 0..3 (10 MMAs per
// 32-wide slab) and 4..7 (26 MMAs); pairs with s + t >= 8 are dropped (2^-56 of the row scales, the level of
// the FP64 accumulation error of the DMMA loop).  The DMMA loop remains for the trtri tiles, for slots without
// planes (sparse / Fourier-feature models, GPSAT_B200_STEP=dmma) and is the reference the tests compare with.
#pragma once
#include "tiles.cuh"
#include "i8slice.cuh"
#include "diag_kernel.cuh"

namespace gpsat {

// cycles per phase of step_kernel summed over CTAs (thread 0's clock), potrf tiles in [0, 16), trtri tiles in
// [16, 32): 0 prologue, 1 main loop (int8: pass 0), 2 drain 0, 3 pass 1, 4 drain 1, 5 C tile store, 6 covariance,
// 7 C W^T, 8 stores + slicing, 9 SYRK + Schur update, 10 diagonal factorisation (11 its panel loop, 12 U sweep, 13 stores +
// slicing), 15 tiles.  -DGPSAT_STEP_PHASES only.  g_step_phase itself is defined in diag_kernel.cuh.
#ifdef GPSAT_STEP_PHASES
#define STEP_PHASE(k) do { if (tid == 0) { long long now_ = clock64(); atomicAdd(&g_step_phase[phbase + (k)], (unsigned long long)(now_ - tphase)); tphase = now_; } } while (0)
#else
#define STEP_PHASE(k) do { } while (0)
#endif

// MMA issue of one pass over k by the calling thread: for every slab all slice pairs (s, t) of the groups s + t = G0 .. G1-1
// (accumulator g - G0), straight-line code per slab - the group range is a template parameter because a run-time test in
// front of each tcgen05.mma costs about 20 cycles of the single issuing thread (measured: 0.1 us per slab).
// HALF: the ring of pass 0 (I8L_NH half stages, i8slice.cuh), else the I8L_NST full stages; each ring has its own barriers
// and counts its slabs from 0.
template <int G0, int G1, bool HALF>
__device__ __forceinline__ bool i8_issue_pass(int nslabs, uint32_t full0, uint32_t empty0, uint32_t accb, int pass,
                                              uint64_t desc0, bool same, uint32_t tb, bool ok) {
  constexpr int NST = HALF ? I8L_NH : I8L_NST;
  for (int sl = 0; sl < nslabs; ++sl) {
    const int st = sl % NST;
    ok = ok && mbar_wait(full0 + 8 * st, (sl / NST) & 1);
    if (!ok) break;
    tc_fence_after();
    const uint64_t adesc = desc0 + (uint64_t)((HALF ? i8l_half_off(st) : (uint32_t)(st * I8L_STAGE_B)) >> 4);
    const uint64_t bdesc = same ? adesc : adesc + (uint64_t)(I8L_OPER_B >> 4);
    const uint32_t acc0 = sl == 0 ? 0u : 1u;
#pragma unroll
    for (int gg = G0; gg < G1; ++gg)
#pragma unroll
      for (int s = 0; s <= gg; ++s)
        umma_i8(tb + (uint32_t)((gg - G0) * NB), adesc + (uint64_t)(s * (I8_SLAB_B >> 4)),
                bdesc + (uint64_t)((gg - s) * (I8_SLAB_B >> 4)), s == 0 ? acc0 : 1u);
    umma_commit(empty0 + 8 * st);
  }
  umma_commit(accb);
  return mbar_wait(accb, pass) && ok;         // only this thread polls; the others sleep in the barrier that follows
}

// Cs = -sum_{k < i NB} L[rowblk][k] L[i][k]^T on the int8 tensor path (see the header comment).  `pipe` is the
// 192 KB area that holds Cs and the W stages afterwards.  All 256 threads call it; ends with a CTA barrier.
// mode I8M_TRTRI: the trtri sum  Cs = +sum_{j NB <= k < i NB} U[rowblk][k] L[i][k]^T  (A planes from d.Uq8).
// mode I8M_DIAG:  Cs = -sum_{k < i NB} L[rowblk][k] L[rowblk][k]^T, the part of the Schur complement of diagonal
// tile rowblk that earlier steps produced (one operand, read once).
// A CTA may call it twice (tile, then diagonal sum): `first` says whether TMEM is to be allocated and the barriers
// are fresh, `last` whether TMEM is released.
enum { I8M_POTRF = 0, I8M_TRTRI = 1, I8M_DIAG = 2 };
__device__ __forceinline__ void potrf_mainloop_i8(const SlotDesc& d, int rowblk, int i, unsigned char* pipe, double* Cs,
                                                  uint64_t* bars, uint32_t* tmem_slot, double* csd, int tid,
                                                  int mode, bool first, bool last, long long& tphase, int phbase,
                                                  double* xr = nullptr, double* xc = nullptr, double* yr = nullptr,
                                                  double* yc = nullptr) {
  const bool is_pot = mode != I8M_TRTRI;
  const bool same = mode == I8M_DIAG;
  const int warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int b = 0; b < I8L_NBAR; ++b) {
      if (!first) asm volatile("mbarrier.inval.shared::cta.b64 [%0];\n" :: "r"(smem_u32(&bars[b])) : "memory");
      mbar_init(smem_u32(&bars[b]), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (first && warp == 0) tmem_alloc(smem_u32(tmem_slot), 512);
  if (tid < NB) csd[tid] = pow2i(lrow_exp(d, (same ? rowblk : i) * NB + tid));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = *tmem_slot;
  const uint32_t full1 = smem_u32(&bars[0]), empty1 = smem_u32(&bars[I8L_NST]), accb = smem_u32(&bars[2 * I8L_NST]);
  const uint32_t fullh = smem_u32(&bars[2 * I8L_NST + 1]), emptyh = smem_u32(&bars[2 * I8L_NST + 1 + I8L_NH]);
  const uint32_t pipe_s = smem_u32(pipe);
  const uint64_t desc0 = umma_desc_kmajor(pipe_s, NB * 16, 128);
  const int row = (warp & 3) * 32 + lane, chalf = warp >> 2;        // accumulator row, 64-column half
  const int exrow = is_pot ? lrow_exp(d, rowblk * NB + row) : urow_exp(d);
  constexpr int SPB = NB / I8_SLAB_K;                               // slabs per 128-block of k
  const int nslabs = (is_pot ? i : i - rowblk) * SPB;
  const int8_t* ga0 = is_pot ? d.Lq + (size_t)ltile_index(rowblk, 0) * I8L_TILE_B
                             : d.Uq8 + (size_t)utile_index(d.T, rowblk, rowblk) * I8L_TILE_B;
  const int8_t* gb0 = same ? ga0 : d.Lq + (size_t)ltile_index(i, is_pot ? 0 : rowblk) * I8L_TILE_B;

  double acc[64];
#pragma unroll
  for (int q = 0; q < 64; ++q) acc[q] = 0.0;
  bool ok = true;
  STEP_PHASE(0);
  const int SG = d.i8_low ? I8L_S - 1 : I8L_S;                // slice-pair groups s + t < SG (low-rho tier: 7, planes 0..6)
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    const int np = pass == 0 ? 4 : SG;                        // planes per operand this pass
    const int ng = pass == 0 ? 4 : SG - 4;                    // groups accumulated (and drained) this pass
    const int nst = pass == 0 ? I8L_NH : I8L_NST;             // pass 0: ring of half stages (two slabs per stage); pass 1 walks the slabs backwards: what pass 0 read last is still in L2, and the int32 sums are exact in any order
    const uint32_t full0 = pass == 0 ? fullh : full1, empty0 = pass == 0 ? emptyh : empty1;
    if (warp == 0) {
      // ---- producer of the A planes: one bulk copy per slab (tiles (r, 0..i-1) are consecutive in the packing).  It also
      // announces the bytes of both operands; the B copies come from a lane of warp 2, because one thread gets a bulk copy
      // out only every 0.285 us whatever its size (tools/bulk_copy_probe.cu), which bounded pass 0 at two requests per slab ----
      if (elect_one())
      for (int sl = 0; sl < nslabs; ++sl) {
        const int st = sl % nst;
        if (sl >= nst) ok = ok && mbar_wait(empty0 + 8 * st, ((sl / nst) - 1) & 1);
        if (!ok) break;
        const uint32_t fb = full0 + 8 * st, sbase = pipe_s + (pass == 0 ? i8l_half_off(st) : (uint32_t)(st * I8L_STAGE_B));
        mbar_expect_tx(fb, (same ? 1 : 2) * np * I8_SLAB_B);
        bulk_g2s(sbase, ga0 + (size_t)(pass == 0 ? sl : nslabs - 1 - sl) * I8L_OPER_B, np * I8_SLAB_B, fb);
      }
    } else if (warp == 2 && !same) {
      // ---- producer of the B planes.  Its complete_tx may reach the full barrier before the A producer's expect_tx: the
      // transaction count is then negative for a moment while the pending arrival keeps the phase open (as with multicast
      // copies issued by a peer CTA).  It cannot run a round ahead: the empty barrier of round k opens after full round k ----
      if (elect_one())
      for (int sl = 0; sl < nslabs; ++sl) {
        const int st = sl % nst;
        if (sl >= nst) ok = ok && mbar_wait(empty0 + 8 * st, ((sl / nst) - 1) & 1);
        if (!ok) break;
        bulk_g2s(pipe_s + (pass == 0 ? i8l_half_off(st) : (uint32_t)(st * I8L_STAGE_B)) + I8L_OPER_B,
                 gb0 + (size_t)(pass == 0 ? sl : nslabs - 1 - sl) * I8L_OPER_B, np * I8_SLAB_B, full0 + 8 * st);
      }
    } else if (warp == 1) {
      // ---- MMA issue: all slice pairs of the pass for every slab ----
      if (elect_one()) {
        if (pass == 0) ok = i8_issue_pass<0, 4, true>(nslabs, full0, empty0, accb, 0, desc0, same, tb, ok);
        else if (SG == I8L_S) ok = i8_issue_pass<4, I8L_S, false>(nslabs, full0, empty0, accb, 1, desc0, same, tb, ok);
        else ok = i8_issue_pass<4, I8L_S - 1, false>(nslabs, full0, empty0, accb, 1, desc0, same, tb, ok);
      }
    } else if (pass == 0 && xr != nullptr && tid >= 96) {
      // warps 3..7 have nothing to do during the loop: they fetch the coordinates the covariance epilogue needs
      load_block_coords(d, rowblk * NB, xr, yr, tid - 96, GPSAT_THREADS - 96);
      load_block_coords(d, i * NB, xc, yc, tid - 96, GPSAT_THREADS - 96);
    }
    __syncwarp();
    ok = __all_sync(0xffffffffu, ok);         // warp-uniform: tcgen05.ld below is warp-collective
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    STEP_PHASE(1 + 2 * pass);
    if (ok) {
      // drain: Horner over the four groups of the pass in FP64 (partial sums are integers below 2^53: exact)
      const double rs = pow2i(exrow - 12 - 7 * (pass * 4 + ng - 1));
#pragma unroll
      for (int c = 0; c < 64; c += 16) {
        uint32_t v[4][16];
#pragma unroll
        for (int g = 0; g < 4; ++g)
          if (g < ng)
            tmem_ld16_nowait(tb + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(g * NB + chalf * 64 + c), v[g]);
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
        for (int q = 0; q < 16; ++q) acc[c + q] = fma(sum[q] * rs, csd[chalf * 64 + c + q], acc[c + q]);
      }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    STEP_PHASE(2 + 2 * pass);
  }
  if (last && warp == 0) tmem_dealloc(tb, 512);
  {
    double* crow = Cs + row * LDC + chalf * 64;
    const double sg = is_pot ? -1.0 : 1.0;
#pragma unroll
    for (int q = 0; q < 64; q += 2) *reinterpret_cast<double2*>(crow + q) = make_double2(sg * acc[q], sg * acc[q + 1]);
  }
  __syncthreads();
  STEP_PHASE(5);
}

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
  // fully unrolled: the first live column fragment of every k block (kb >> 1) becomes a compile-time constant, so
  // the B loads of a k block are issued together instead of one predicated load in front of each pair of DMMAs
#pragma unroll
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
step_kernel(const SlotDesc* __restrict__ slots, int i, int do_inv, int S, int ntile, int diag_sum) {
  // 1-D grid of S * ntile CTAs.  The first S are tile 0 of every slot (the long ones: they also factor the next
  // diagonal tile); the others follow slot by slot, so that the tiles of one slot - which all read block row i
  // of L - run at the same time and share it in L2.
  int slot, x;
  if ((int)blockIdx.x < S) { slot = blockIdx.x; x = 0; }
  else { const int id = blockIdx.x - S; slot = id / (ntile - 1); x = 1 + id - slot * (ntile - 1); }
  const SlotDesc& d = slots[slot];
  if (d.expert < 0 || i >= d.T) return;
  const int npot = d.TR - 1 - i;
  const int ninv = do_inv ? i : 0;
  if (x >= npot + ninv) return;
  const bool is_pot = x < npot;
  const int rowblk = is_pot ? (i + 1 + x) : (x - npot);   // r (potrf) or j (trtri)

  extern __shared__ __align__(128) double smem_step[];
  double* smem = smem_step;
  __shared__ __align__(8) uint64_t bars[I8L_NBAR];
  __shared__ uint32_t tmem_slot;
  __shared__ double csd[NB];
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
#ifdef GPSAT_STEP_PHASES
  long long tphase = clock64();
  const int phbase = is_pot ? 0 : 16;
#else
  long long tphase = 0;
  const int phbase = 0;
#endif

  // ---- 1. main loop (slab loads first, the coordinate loads hide behind them) ----
  const bool use_i8 = d.Lq != nullptr && i > 0 && (is_pot || d.Uq8 != nullptr);
  if (use_i8) {
    // the CTA of tile (i + 1, i) runs the int8 loop a second time for the diagonal tile's Schur complement
    potrf_mainloop_i8(d, rowblk, i, reinterpret_cast<unsigned char*>(smem), Cs, bars, &tmem_slot, csd, tid,
                      is_pot ? I8M_POTRF : I8M_TRTRI, true, !(diag_sum && is_pot && x == 0 && rowblk < d.T), tphase,
                      phbase, is_pot ? xr : nullptr, xc, yr, yc);
    wmult_prefetch(Wg, wst, tid);      // W_ii streams in under the epilogue
  } else {
    Acc acc;
    acc_zero(acc);
    STEP_PHASE(0);
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
    STEP_PHASE(1);
    wmult_prefetch(Wg, wst, tid);      // W_ii streams in under the epilogue
    // ---- 2. C to shared memory ----
    acc_to_smem(acc, Cs, tid, is_pot ? -1.0 : 1.0);
  }
  __syncthreads();
  if (!use_i8) STEP_PHASE(5);
  if (is_pot) {
    add_aug_tile(d, rowblk * NB, i * NB, xr, xc, yc, Cs, tid, false);
    __syncthreads();
  }
  STEP_PHASE(6);

  // ---- 3. out = C * W_ii^T ----
  AccStrip out;
  acc_zero(out);
  wmult_strip(out, Cs, Wg, wst, tid);
  STEP_PHASE(7);
#ifdef GPSAT_STEP_PHASES
  if (tid == 0) atomicAdd(&g_step_phase[phbase + 15], 1ull);
#endif
  const double sign = is_pot ? 1.0 : -1.0;
  // FP64 copy of the tile: read back only by the FP64 mma.sync loops (this slot is not on them in this evaluation), by the
  // FP64 K^-1 kernel / the debug extraction (skip_fp64 = 0 then), and - rows from the observation row n on, column block of
  // column n - by the prediction reduction and the alpha column of the gradient.  Everything else reads the planes.
  const bool dead = d.skip_fp64 && use_i8 && (is_pot ? (rowblk + 1) * NB <= d.n : (d.Uq8 != nullptr && (i + 1) * NB <= d.n));
  if (!dead) {
    double* Og = d.A + ((size_t)rowblk * NB + 16 * strip + lr) * ld + (size_t)i * NB + 2 * lk;
#pragma unroll
    for (int rf = 0; rf < 2; ++rf)
#pragma unroll
      for (int cf = 0; cf < 16; ++cf)
        *reinterpret_cast<double2*>(Og + (size_t)rf * 8 * ld + cf * 8) =
            make_double2(sign * out.c[rf][cf][0], sign * out.c[rf][cf][1]);
  }
  if (!is_pot) {
    if (d.Uq8 == nullptr) return;
    // later trtri steps and the gradient's K^-1 products read U_ji on the int8 path
    double* Cw = Cs + (16 * strip + lr) * LDC + 2 * lk;
#pragma unroll
    for (int rf = 0; rf < 2; ++rf)
#pragma unroll
      for (int cf = 0; cf < 16; ++cf)
        *reinterpret_cast<double2*>(Cw + rf * 8 * LDC + cf * 8) = make_double2(-out.c[rf][cf][0], -out.c[rf][cf][1]);
    __syncthreads();
    slice_tile_u(d, i * NB, Cs, d.Uq8 + (size_t)utile_index(d.T, rowblk, i) * I8L_TILE_B, tid);
    STEP_PHASE(8);
    return;
  }
  const bool syrk = rowblk < d.T;                          // block row has a diagonal tile to update
  const bool slice = d.Lq != nullptr && i + 1 < d.T;       // later steps read this tile on the int8 path
  if (!syrk && !slice) return;

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
  if (slice) slice_tile_l(d, rowblk * NB, Cs, d.Lq + (size_t)ltile_index(rowblk, i) * I8L_TILE_B, tid);
  STEP_PHASE(8);
  // With planes the Schur complement of a diagonal tile is formed in one go by the CTA that factors it (below),
  // so the other tiles of the block row neither update nor touch S_rr.
  const bool diag_i8 = diag_sum && d.Lq != nullptr;
  if (!syrk || (diag_i8 && x != 0)) return;
  if (x == 0 && i > 0 && diag_i8) {
    // ---- 5 (int8 path). rowblk == i + 1: the Schur complement of the next diagonal tile in one go, -sum_{k <= i} L_rk L_rk^T
    // over the planes of block row r INCLUDING the tile this CTA has just sliced (it is the next tile of the packing, so the
    // slab stream simply runs four slabs further) - no FP64 SYRK of the own tile, nothing parked in the slot matrix.  The
    // planes were written through the generic proxy and are read back by bulk copies: fence both ways first.
    __threadfence();
    asm volatile("fence.proxy.async;\n" ::: "memory");
    __syncthreads();
    potrf_mainloop_i8(d, rowblk, i + 1, reinterpret_cast<unsigned char*>(smem), Cs, bars, &tmem_slot, csd, tid, I8M_DIAG,
                      false, true, tphase, phbase);
    add_aug_tile(d, rowblk * NB, rowblk * NB, xr, xr, yr, Cs, tid, true);
    __syncthreads();
    STEP_PHASE(9);
    diag_factor_tile(d, rowblk, Cs, wst, tid);    // scratch: the W stages + coordinate arrays
    STEP_PHASE(10);
    return;
  }
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
      if (i > 0) {                     // (the int8 path has left above)
#pragma unroll 16
        for (int r = tid >> 7; r < NB; r += GPSAT_THREADS / NB)
          if (c <= r) Cs[r * LDC + c] += Sg[(size_t)r * ld + c];
      }
      __syncthreads();
      STEP_PHASE(9);
      diag_factor_tile(d, rowblk, Cs, wst, tid);    // scratch: the W stages + coordinate arrays
      STEP_PHASE(10);
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
  STEP_PHASE(9);
}

}  // namespace gpsat
