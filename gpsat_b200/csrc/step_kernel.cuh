// step_kernel(i): all off-diagonal work of block column i for every expert in flight.
//   blockIdx.y = slot, blockIdx.x = tile within the step:
//     [0, TR-1-i)            potrf tile (r, i), r = i + 1 + x
//     [TR-1-i, TR-1-i + i)   trtri tile (j, i), j = x - (TR-1-i)      (only when do_inv)
#pragma once
#include "tiles.cuh"

namespace gpsat {

// out = Cs(128 x 128, smem, stride LDC) * W^T, W row-major 128 x 128 in global memory, streamed
// through two B-only stages.
__device__ __forceinline__ void gemm_tile_times_Wt(Acc& acc, const double* Cs, const double* Wg,
                                                   double* wst, int tid) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  constexpr int NSL = NB / KC;
  load_slab(wst, Wg, NB, tid);
  cp_async_commit();
  for (int s = 0; s < NSL; ++s) {
    if (s + 1 < NSL) load_slab(wst + ((s + 1) & 1) * NB * LDS_S, Wg + (s + 1) * KC, NB, tid);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    mma_slab(acc, Cs + s * KC, LDC, wst + (s & 1) * NB * LDS_S, LDS_S, KC / 4, wm, wn, lane);
    __syncthreads();
  }
  cp_async_wait<0>();
}

__device__ __forceinline__ void acc_to_smem(const Acc& acc, double* Cs, int tid, double sign) {
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  const int lr = lane >> 2, lk = lane & 3;
#pragma unroll
  for (int fm = 0; fm < 8; ++fm)
#pragma unroll
    for (int fn = 0; fn < 4; ++fn) {
      int r = wm * 64 + fm * 8 + lr, c = wn * 32 + fn * 8 + 2 * lk;
      double2 v = make_double2(sign * acc.c[fm][fn][0], sign * acc.c[fm][fn][1]);
      *reinterpret_cast<double2*>(Cs + r * LDC + c) = v;
    }
}

__global__ void __launch_bounds__(GPSAT_THREADS, 1)
step_kernel(const SlotDesc* __restrict__ slots, int i, int do_inv) {
  const SlotDesc& d = slots[blockIdx.y];
  if (d.expert < 0 || i >= d.T) return;
  const int npot = d.TR - 1 - i;
  const int ninv = do_inv ? i : 0;
  const int x = blockIdx.x;
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
  const int warp = tid >> 5, lane = tid & 31, wm = warp >> 2, wn = warp & 3;
  const int lr = lane >> 2, lk = lane & 3;
  const int ld = d.ld;
  const double* Wg = d.Wd + (size_t)i * NB * NB;

  if (is_pot) {
    load_block_coords(d, rowblk * NB, xr, yr, tid);
    load_block_coords(d, i * NB, xc, yc, tid);
  }

  Acc acc;
  acc_zero(acc);
  const double* Ag = d.A + (size_t)rowblk * NB * ld;
  const double* Bg = d.A + (size_t)i * NB * ld;
  const int kbeg = is_pot ? 0 : rowblk * NB;
  gemm_nt_mainloop(acc, Ag, Bg, ld, kbeg, i * NB, smem, tid);
  __syncthreads();   // coords visible; stages free

  // ---- C = K_ri - acc (potrf) or acc (trtri), to shared memory ----
  if (is_pot) {
#pragma unroll
    for (int fm = 0; fm < 8; ++fm)
#pragma unroll
      for (int fn = 0; fn < 4; ++fn) {
        int r = wm * 64 + fm * 8 + lr, c = wn * 32 + fn * 8 + 2 * lk;
        int gr = rowblk * NB + r, gc = i * NB + c;
        double v0 = tile_value(d, gr, gc, xr + r * GPSAT_DMAX, xc + c * GPSAT_DMAX, yc[c]);
        double v1 = tile_value(d, gr, gc + 1, xr + r * GPSAT_DMAX, xc + (c + 1) * GPSAT_DMAX, yc[c + 1]);
        *reinterpret_cast<double2*>(Cs + r * LDC + c) =
            make_double2(v0 - acc.c[fm][fn][0], v1 - acc.c[fm][fn][1]);
      }
  } else {
    acc_to_smem(acc, Cs, tid, 1.0);
  }
  __syncthreads();

  // ---- out = C * W_ii^T ----
  acc_zero(acc);
  gemm_tile_times_Wt(acc, Cs, Wg, wst, tid);
  const double sign = is_pot ? 1.0 : -1.0;
  {
    double* Og = d.A + (size_t)rowblk * NB * ld + (size_t)i * NB;
#pragma unroll
    for (int fm = 0; fm < 8; ++fm)
#pragma unroll
      for (int fn = 0; fn < 4; ++fn) {
        int r = wm * 64 + fm * 8 + lr, c = wn * 32 + fn * 8 + 2 * lk;
        *reinterpret_cast<double2*>(Og + (size_t)r * ld + c) =
            make_double2(sign * acc.c[fm][fn][0], sign * acc.c[fm][fn][1]);
      }
  }
  if (!is_pot || rowblk >= d.T) return;

  // ---- S_rr -= L_ri L_ri^T  (running Schur complement of the diagonal tile of block row r) ----
  acc_to_smem(acc, Cs, tid, 1.0);     // all warps passed the last barrier of gemm_tile_times_Wt
  __syncthreads();
  acc_zero(acc);
  mma_slab(acc, Cs, LDC, Cs, LDC, NB / 4, wm, wn, lane);
  double* Sg = d.A + (size_t)rowblk * NB * ld + (size_t)rowblk * NB;
#pragma unroll
  for (int fm = 0; fm < 8; ++fm)
#pragma unroll
    for (int fn = 0; fn < 4; ++fn) {
      int r = wm * 64 + fm * 8 + lr, c = wn * 32 + fn * 8 + 2 * lk;
      if (c > r + 1 && (c >> 3) > (r >> 3)) continue;    // strictly upper fragment: never read
      double2 base;
      if (i == 0) {
        int gr = rowblk * NB + r, gc = rowblk * NB + c;
        base.x = tile_value(d, gr, gc, xr + r * GPSAT_DMAX, xr + c * GPSAT_DMAX, yr[c]);
        base.y = tile_value(d, gr, gc + 1, xr + r * GPSAT_DMAX, xr + (c + 1) * GPSAT_DMAX, yr[c + 1]);
      } else {
        base = *reinterpret_cast<const double2*>(Sg + (size_t)r * ld + c);
      }
      base.x -= acc.c[fm][fn][0];
      base.y -= acc.c[fm][fn][1];
      *reinterpret_cast<double2*>(Sg + (size_t)r * ld + c) = base;
    }
}

}  // namespace gpsat
