// Batched local data selection: for every expert location, the rows of a point
// table that satisfy all conditions of a `local_select` list - the reference evaluates them per
// expert on the host (DataLoader.local_data_select, GPSat/dataloader.py:2354-2447: column-wise
// comparisons against location + val, KD-tree ball queries for multi-column entries) and the
// prediction-location radius filter (GPSat/prediction_locations.py:18-43).  Here one launch covers
// all experts: E x N predicate evaluations, row order preserved (stable compaction).
#pragma once
#include "common.cuh"

namespace gpsat {

constexpr int SEL_MAX_COND = 8;
constexpr int SEL_CHUNK = 1024;          // points per CTA (4 per thread)

struct SelectCond {
  int kind;          // 0: data[col0] <op> loc[loc0] + val;  1: sum_k (data[col_k] - loc[loc_k])^2 <op> val^2
  int op;            // 0 >=, 1 >, 2 ==, 3 <, 4 <=
  int ncol;
  int col[3];        // columns of the point table
  int loc[3];        // columns of the expert-location table
  double val;
};

struct SelectSpec {
  int ncond;
  SelectCond c[SEL_MAX_COND];
};

__device__ __forceinline__ bool sel_cmp(int op, double a, double b) {
  switch (op) {
    case 0: return a >= b;
    case 1: return a > b;
    case 2: return a == b;
    case 3: return a < b;
    default: return a <= b;
  }
}

__device__ __forceinline__ bool sel_match(const SelectSpec& s, const double* __restrict__ row,
                                          const double* __restrict__ loc) {
  bool ok = true;
  for (int q = 0; q < s.ncond; ++q) {
    const SelectCond& c = s.c[q];
    if (c.kind == 0) {
      ok = ok && sel_cmp(c.op, row[c.col[0]], loc[c.loc[0]] + c.val);
    } else {
      double d2 = 0.0;
      for (int k = 0; k < c.ncol; ++k) {
        const double t = row[c.col[k]] - loc[c.loc[k]];
        d2 = __dadd_rn(d2, __dmul_rn(t, t));   // no fma contraction: the host's sum of squares, bit for bit
      }
      ok = ok && sel_cmp(c.op, d2, c.val * c.val);
    }
  }
  return ok;
}

// pass 1: matches per (expert, chunk)
__global__ void __launch_bounds__(256) select_count_kernel(int64_t N, int C, const double* __restrict__ data, int CL,
                                                           const double* __restrict__ locs, SelectSpec spec,
                                                           int nch, int* __restrict__ cnt) {
  const int e = blockIdx.y, ch = blockIdx.x;
  const double* loc = locs + (size_t)e * CL;
  int mine = 0;
  for (int j = 0; j < SEL_CHUNK / 256; ++j) {
    const int64_t i = (int64_t)ch * SEL_CHUNK + j * 256 + threadIdx.x;
    if (i < N && sel_match(spec, data + (size_t)i * C, loc)) ++mine;
  }
  __shared__ int red[256];
  red[threadIdx.x] = mine;
  __syncthreads();
  for (int h = 128; h > 0; h >>= 1) {
    if (threadIdx.x < h) red[threadIdx.x] += red[threadIdx.x + h];
    __syncthreads();
  }
  if (threadIdx.x == 0) cnt[(size_t)e * nch + ch] = red[0];
}

// exclusive scan of the chunk counts of every expert (in place) and the expert totals
__global__ void select_scan_kernel(int E, int nch, int* __restrict__ cnt, int64_t* __restrict__ counts) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  int run = 0;
  for (int ch = 0; ch < nch; ++ch) {
    const int c = cnt[(size_t)e * nch + ch];
    cnt[(size_t)e * nch + ch] = run;
    run += c;
  }
  counts[e] = run;
}

// pass 2: row indices in increasing order at idx[off[e] + ...]
__global__ void __launch_bounds__(256) select_fill_kernel(int64_t N, int C, const double* __restrict__ data, int CL,
                                                          const double* __restrict__ locs, SelectSpec spec, int nch,
                                                          const int* __restrict__ base, const int64_t* __restrict__ off,
                                                          int64_t* __restrict__ idx) {
  const int e = blockIdx.y, ch = blockIdx.x;
  const double* loc = locs + (size_t)e * CL;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __shared__ int wsum[8];
  int64_t pos = off[e] + base[(size_t)e * nch + ch];
  for (int j = 0; j < SEL_CHUNK / 256; ++j) {
    const int64_t i = (int64_t)ch * SEL_CHUNK + j * 256 + threadIdx.x;
    const bool m = i < N && sel_match(spec, data + (size_t)i * C, loc);
    const unsigned bal = __ballot_sync(0xffffffffu, m);
    if (lane == 0) wsum[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < 8; ++w) { if (w < warp) before += wsum[w]; total += wsum[w]; }
    if (m) idx[pos + before + __popc(bal & ((1u << lane) - 1u))] = i;
    pos += total;
    __syncthreads();
  }
}

}  // namespace gpsat
