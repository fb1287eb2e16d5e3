// Error-free int8 slicing of tiles of the factor for the int8 tensor path of step_kernel (see step_kernel.cuh).
#pragma once
#include "i8mma.cuh"

namespace gpsat {

constexpr int I8L_S = 8;                                 // slices per value of L
constexpr int I8L_TILE_B = I8L_S * I8_PLANE_B;           // 128 KB of planes per tile
constexpr int I8L_OPER_B = I8L_S * I8_SLAB_B;            // planes of one operand and slab: 32 KB
constexpr int I8L_STAGE_B = 2 * I8L_OPER_B;
constexpr int I8L_NST = 3;
constexpr int I8L_PIPE_B = I8L_NST * I8L_STAGE_B;        // 192 KB: exactly the C tile + the W stages it aliases
// Pass 0 moves planes 0..3 only, so every stage holds two slabs: a ring of 2 I8L_NST half stages with its own barriers.
// Half stage h: A planes at i8l_half_off(h), B planes I8L_OPER_B behind them.
constexpr int I8L_NH = 2 * I8L_NST;
constexpr int I8L_NBAR = 2 * I8L_NST + 1 + 2 * I8L_NH;   // full[NST], empty[NST], accumulators, full0[NH], empty0[NH]
__host__ __device__ constexpr uint32_t i8l_half_off(int h) {
  return (uint32_t)(h >> 1) * I8L_STAGE_B + (uint32_t)(h & 1) * 4 * I8_SLAB_B;
}
static_assert(I8L_PIPE_B <= (SM_TILE_D + SM_WST_D) * 8, "int8 pipeline must fit in the tile + W stage area");
constexpr int I8L_MAX_T = 256;                           // int32 accumulators: k <= 32768 keeps |G_g| < 2^31

// index of tile (r, k), k < r, in the packed strict lower block triangle
__device__ __forceinline__ int ltile_index(int r, int k) { return r * (r - 1) / 2 + k; }

// A-priori scale exponent of row gr of the augmented factor: |L[gr][k]| < 2^e for every k.
//   observation rows: |row|^2 = K_rr = var + noise;   y row: |L^-1 y|^2 <= |y|^2 / noise;
//   prediction rows: |L^-1 k_*|^2 <= k(x*, x*) = var
__device__ __forceinline__ int lrow_exp(const SlotDesc& d, int gr) {
  double b2 = gr < d.n ? d.var + d.noise : (gr == d.n ? d.ynorm2 / d.noise : d.var);
  if (!(b2 > 1e-280)) b2 = 1e-280;
  if (!(b2 < 1e280)) b2 = 1e280;
  const int e2 = (int)((__double_as_longlong(b2) >> 52) & 0x7ff) - 1022;     // b2 < 2^e2
  return (e2 + 1) >> 1;                                                      // ceil(e2 / 2)
}

// A-priori scale exponent of the rows of U = L^-T: |U_rk| <= |L^-1|_2 = 1 / sqrt(lambda_min(K + s2 I)) <= 1 / sqrt(noise)
__device__ __forceinline__ int urow_exp(const SlotDesc& d) {
  double b2 = 1.0 / d.noise;
  if (!(b2 > 1e-280)) b2 = 1e-280;
  if (!(b2 < 1e280)) b2 = 1e280;
  const int e2 = (int)((__double_as_longlong(b2) >> 52) & 0x7ff) - 1022;
  return (e2 + 1) >> 1;
}

// Split the 128 x 128 tile val(r, c) * 2^(6 - e) into I8L_S exact int8 slice planes (see i8mma.cuh for the
// rounding trick) and store them in the K-major UMMA layout [slab][plane][k / 16][row][16 B].  Two threads per
// row; sc(r) = 2^(6 - e_r).
// ns: planes to produce (I8L_S, or I8L_S - 1 for slots in the low-rho tier, whose consumers read planes 0..6 only - the
// tier is a property of the slot's current evaluation, so producer and consumers always agree).
template <typename FV, typename FS>
__device__ __forceinline__ void slice_tile_f(FV val, FS sc_of_row, int8_t* planes, int tid, int ns = I8L_S) {
  const int r = tid >> 1, h = tid & 1;
  const double sc = sc_of_row(r);
  double magic[I8L_S];
#pragma unroll
  for (int s = 0; s < I8L_S; ++s) magic[s] = 1.5 * pow2i(52 - 7 * s);
#pragma unroll 1
  for (int part = 0; part < 2; ++part) {
    const int c0 = h * 64 + part * 32;
    double x[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) x[q] = val(r, c0 + q) * sc;
#pragma unroll
    for (int s = 0; s < I8L_S; ++s) {
      if (s >= ns) break;
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
        *reinterpret_cast<uint4*>(planes + (size_t)(c >> 5) * I8L_OPER_B + s * I8_SLAB_B +
                                  ((c >> 4) & 1) * (NB * 16) + r * 16) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
}

// The tile held in Cs ([128][LDC]) as a tile of L (per-row scales of rows g0r ..).
__device__ __forceinline__ void slice_tile_l(const SlotDesc& d, int g0r, const double* Cs, int8_t* planes, int tid) {
  slice_tile_f([&](int r, int c) { return Cs[r * LDC + c]; },
               [&](int r) { return pow2i(6 - lrow_exp(d, g0r + r)); }, planes, tid, d.i8_low ? I8L_S - 1 : I8L_S);
}

// The tile held in Cs as tile (.., column block at g0c) of U: one scale; columns >= n (the -alpha column of the
// augmented inverse and the padding) are stored as zeros, so products of planes give Ks^-1 without alpha alpha^T.
__device__ __forceinline__ void slice_tile_u(const SlotDesc& d, int g0c, const double* Cs, int8_t* planes, int tid) {
  const double sc = pow2i(6 - urow_exp(d));
  const int ncol = d.n - g0c;
  slice_tile_f([&](int r, int c) { return c < ncol ? Cs[r * LDC + c] : 0.0; }, [&](int) { return sc; }, planes, tid,
               d.i8_low ? I8L_S - 1 : I8L_S);
}

// Diagonal tile U_ii = W_ii^T after diag_factor_tile: strict upper part in Cs, diagonal in wdiag, zero below;
// rows / columns >= ncol (augmentation) are stored as zeros.
__device__ __forceinline__ void slice_tile_udiag(const SlotDesc& d, int ncol, const double* Cs, const double* wdiag,
                                                 int8_t* planes, int tid) {
  const double sc = pow2i(6 - urow_exp(d));
  slice_tile_f([&](int r, int c) { return c >= ncol ? 0.0 : (c > r ? Cs[r * LDC + c] : (c == r ? wdiag[r] : 0.0)); },
               [&](int) { return sc; }, planes, tid, d.i8_low ? I8L_S - 1 : I8L_S);
}

}  // namespace gpsat
