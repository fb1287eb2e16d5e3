// Host side of the engine + the per-expert optimiser kernel + the C ABI (include/gpsat_b200.h).
//
// Scheduling model ("continuous batching" of experts): S slots are in flight on the GPU.  One
// tick = one batched evaluation of (lml, grad) for every occupied slot (a left-looking blocked
// sweep, see tiles.cuh) followed by lbfgs_kernel, which advances every expert's private L-BFGS
// state machine (lbfgs.cuh) by one evaluation, retires converged experts and refills their
// slots from the queue of pending experts - all on the device.  The host only replays the
// launch sequence until the device reports that no slot is occupied.  This replaces the
// reference's serial loop `for idx in range(len(xprt_locs))` (GPSat/local_experts.py:930) with
// its per-expert scipy.optimize.minimize call (GPSat/models/gpflow_models.py:317-321).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <numeric>

#include "../../include/gpsat_b200.h"
#include "common.cuh"
#include "lbfgs.cuh"
#include "tiles.cuh"
#include "diag_kernel.cuh"
#include "lauum_kernel.cuh"
#include "i8mma.cuh"
#include "step_kernel.cuh"
#include "lauum_q_kernel.cuh"

namespace gpsat {

static thread_local std::string g_last_error;

#define CUDA_TRY(expr)                                                                          \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      g_last_error = std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + ":" + \
                     std::to_string(__LINE__) + ")";                                            \
      return 1;                                                                                 \
    }                                                                                           \
  } while (0)

#define REQUIRE(cond, msg)                                      \
  do {                                                          \
    if (!(cond)) { g_last_error = std::string("gpsat: ") + msg; return 2; } \
  } while (0)

constexpr int PG = GPSAT_DMAX + 2;     // stride of per-tile gradient partials
constexpr double LOG2PI = 1.8378770664093453;

// ---- per-expert problem description on the device ------------------------------------------
struct BatchDev {
  int kind, D, P;
  int64_t E;
  const int64_t* off_n;     // device copies
  const int64_t* off_m;
  const double* X;
  const double* y;
  const double* Xs;
  const int* order;         // experts sorted by n, descending
  const double* ynorm2;     // [E] |y|^2 per expert
  double i8_rho_max;        // precision rule: int8-sliced k-sums only while (var + noise) / noise <= this (0 = always)
  double i8_rho_low;        // ... and with one slice-pair group fewer while (var + noise) / noise <= this (0 = never)
  int skip_fp64;            // see SlotDesc::skip_fp64 (set by the host after engine_alloc_slots)
  double i8_tier_guard;     // fit: the low-rho tier only while |stp g'd| > guard * max(|f|, 1) (lbfgs_kernel)
};

struct FitDev {
  double* theta;            // [E][P] in: start, out: optimum
  const int* tkind;         // [E][P]
  const double* ta;
  const double* tb;
  const unsigned char* trainable;
  const int* pkind;         // [E][P] prior family, nullptr = no priors
  const double* pa;
  const double* pb;
  double* objective;        // [E] minimised value at the optimum: -(lml + log prior)
  double* lml;              // [E] log marginal likelihood at the optimum (may be nullptr)
  int* n_evals;
  int* n_iters;
  int* status;
  LbfgsState* states;       // [S]
  int* counters;            // [0] next expert in queue, [1] occupied slots, [2] max T, [3] evals,
                            // [4] occupied slots whose next evaluation the precision rule runs on FP64 mma.sync
  LbfgsOptions opt;
};

// Slot-private scratch carved from the workspace; constant for the lifetime of a call.
struct SlotMem {
  double* A; double* Wd; double* Ld; double* logdet; double* quad; double* gpart; int* notpd;
  int8_t* Uq; int* uex; int8_t* Lq; int8_t* Uq8;
};

// |y|^2 per expert in a fixed summation order (one CTA of 128 threads per expert)
__global__ void __launch_bounds__(128) ynorm_kernel(const int64_t* __restrict__ off_n, const double* __restrict__ y,
                                                    double* __restrict__ out) {
  __shared__ double red[4];
  const int e = blockIdx.x, tid = threadIdx.x;
  const double* ye = y + off_n[e];
  const int n = (int)(off_n[e + 1] - off_n[e]);
  double s = 0.0;
  for (int k = tid; k < n; k += 128) s = fma(ye[k], ye[k], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((tid & 31) == 0) red[tid >> 5] = s;
  __syncthreads();
  if (tid == 0) out[e] = (red[0] + red[1]) + (red[2] + red[3]);
}

}  // namespace gpsat
#include "sgpr_kernels.cuh"
#include "vff_kernels.cuh"
#include "post_kernels.cuh"
#include "select_kernels.cuh"
namespace gpsat {

// ---- transforms (gpflow.utilities.positive / tfp.bijectors.Sigmoid) -------------------------
__device__ __forceinline__ double dev_sigmoid(double x) {
  double e = exp(-fabs(x));
  return x >= 0.0 ? 1.0 / (1.0 + e) : e / (1.0 + e);
}
__device__ __forceinline__ double dev_softplus(double x) {
  double l = log1p(exp(-fabs(x)));
  return x > 0.0 ? x + l : l;
}
__device__ __forceinline__ double transform_fwd(int kind, double a, double b, double xi) {
  if (kind == GPSAT_TRANSFORM_SIGMOID) { double s = dev_sigmoid(xi); return b * s + a * (1.0 - s); }
  return a + dev_softplus(xi);
}
__device__ __forceinline__ double transform_inv(int kind, double a, double b, double th) {
  if (kind == GPSAT_TRANSFORM_SIGMOID) return log(th - a) - log(b - th);
  double y = th - a;
  return y + log(-expm1(-y));
}
__device__ __forceinline__ double transform_dfwd(int kind, double a, double b, double xi) {
  double s = dev_sigmoid(xi);
  return kind == GPSAT_TRANSFORM_SIGMOID ? (b - a) * s * (1.0 - s) : s;
}

// log prior density of one trainable parameter in GPflow's convention (gpflow.Parameter.
// log_prior_density with prior_on = constrained): log p(theta) + log |d theta / d xi|.
// Returns the value and adds its derivative w.r.t. xi to *dxi.
__device__ __forceinline__ double log_prior_xi(int pk, double pa, double pb, int tk, double ta, double tb,
                                               double xi, double th, double* dxi) {
  if (pk == GPSAT_PRIOR_NONE) return 0.0;
  const double HALF_LOG2PI = 0.9189385332046727;
  double lp, dlp;
  if (pk == GPSAT_PRIOR_NORMAL) {
    double z = (th - pa) / pb;
    lp = -0.5 * z * z - log(pb) - HALF_LOG2PI;
    dlp = -z / pb;
  } else if (pk == GPSAT_PRIOR_LOGNORMAL) {
    double lt = log(th), z = (lt - pa) / pb;
    lp = -lt - log(pb) - HALF_LOG2PI - 0.5 * z * z;
    dlp = -(1.0 + z / pb) / th;
  } else {
    lp = pa * log(pb) - lgamma(pa) + (pa - 1.0) * log(th) - pb * th;
    dlp = (pa - 1.0) / th - pb;
  }
  double s = dev_sigmoid(xi);
  double ldj, dldj;
  if (tk == GPSAT_TRANSFORM_SIGMOID) { ldj = log(tb - ta) + log(s) + log1p(-s); dldj = 1.0 - 2.0 * s; }
  else { ldj = -dev_softplus(-xi); dldj = 1.0 - s; }
  *dxi += dlp * transform_dfwd(tk, ta, tb, xi) + dldj;
  return lp + ldj;
}

__device__ __forceinline__ void slot_set_expert(SlotDesc& d, const SlotMem& mem, const BatchDev& b, int e,
                                                int with_pred, int ld) {
  d.expert = e;
  d.kind = b.kind;
  d.D = b.D;
  int n = (int)(b.off_n[e + 1] - b.off_n[e]);
  int m = (with_pred && b.off_m) ? (int)(b.off_m[e + 1] - b.off_m[e]) : 0;
  d.n = n; d.m = m;
  d.T = (n + 1 + GPSAT_NB - 1) / GPSAT_NB;
  d.TR = (n + 1 + m + GPSAT_NB - 1) / GPSAT_NB;
  d.ld = ld;
  d.X = b.X + b.off_n[e] * b.D;
  d.y = b.y + b.off_n[e];
  d.Xs = (m > 0) ? b.Xs + b.off_m[e] * b.D : nullptr;
  d.A = mem.A; d.Wd = mem.Wd; d.Ld = mem.Ld; d.logdet = mem.logdet; d.quad = mem.quad;
  d.gpart = mem.gpart; d.notpd = mem.notpd;
  d.pre = nullptr; d.pre_s = nullptr; d.ldpre = 0;
  d.Uq = mem.Uq; d.uex = mem.uex;
  d.Lq = d.Lq_mem = mem.Lq; d.Uq8 = d.Uq8_mem = mem.Uq8; d.ynorm2 = b.ynorm2 ? b.ynorm2[e] : 0.0;
  d.skip_fp64 = b.skip_fp64;
  *d.notpd = 0;
  *d.quad = 0.0;
}

// Precision rule of the int8-sliced k-sums (i8slice.cuh): their truncation is 2^-56 of A-PRIORI row bounds
// (|L_rk| <= sqrt(var + noise), |U_rk| <= 1 / sqrt(noise), |L^-1 y| <= |y| / sqrt(noise)), so against the entries
// themselves the accuracy falls with rho = (var + noise) / noise.  Evaluations with rho above rho_max take the FP64
// mma.sync loops instead (the FP64 tiles are always stored, so the choice is free per slot and per evaluation).
// The threshold comes from the measured envelope, profiles/r02_precision_envelope.md.
// Low-rho tier: the same bound argument read the other way - while rho <= rho_low the a-priori scales sit within a few
// bits of the entries, so the last group of slice pairs (s + t = 7: 2^-49 of the scale product) is below what the
// north star asks by the margin measured in profiles/r02_low_tier_envelope.md, and the sweep's k-sums run with 28 instead
// of 36 slice pairs per slab (the gradient's K^-1 products keep their 7 planes: with 6 the gradient error reached 8e-8).
// c2 / north-star / c3 experts end at rho = 3 .. 35, median 11 (tools/rho_hist.py).
__device__ __forceinline__ void slot_set_theta(SlotDesc& d, const double* theta, double rho_max, double rho_low) {
  for (int k = 0; k < d.D; ++k) d.inv_ls[k] = 1.0 / theta[k];
  d.var = theta[d.D];
  d.noise = theta[d.D + 1];
  const bool ok = !(rho_max > 0.0) || (d.var + d.noise <= rho_max * d.noise);
  d.Lq = ok ? d.Lq_mem : nullptr;
  d.Uq8 = ok ? d.Uq8_mem : nullptr;
  // the squared-exponential kernel's matrices are worse conditioned at the same rho (gradient of the n = 2000, rho = 31
  // sweep case: 4e-8 with the tier, 8e-10 without; 7e-11 at rho = 11), so its tier ends at 3/8 of the threshold (12)
  const double lim = d.kind == GPSAT_RBF ? 0.375 * rho_low : rho_low;
  d.i8_low = (lim > 0.0 && d.var + d.noise <= lim * d.noise) ? 1 : 0;
}

// lml and d lml / d theta from the sweep's by-products (fixed summation order => reproducible)
__device__ __forceinline__ double slot_finish_eval(const SlotDesc& d, const double* theta, double* gtheta,
                                                   bool with_grad) {
  double logdet = 0.0;
  for (int t = 0; t < d.T; ++t) logdet += d.logdet[t];
  double lml = -0.5 * (*d.quad) - logdet - 0.5 * d.n * LOG2PI;
  if (*d.notpd) lml = -INFINITY;
  if (with_grad) {
    double s[PG];
    for (int p = 0; p < PG; ++p) s[p] = 0.0;
    int ntile = d.T * (d.T + 1) / 2;
    for (int t = 0; t < ntile; ++t)
      for (int p = 0; p < PG; ++p) s[p] += d.gpart[(size_t)t * PG + p];
    for (int k = 0; k < d.D; ++k) gtheta[k] = 0.5 * d.var * s[k] / theta[k];
    gtheta[d.D] = 0.5 * s[GPSAT_DMAX];
    gtheta[d.D + 1] = 0.5 * s[GPSAT_DMAX + 1];
  }
  return lml;
}

// Start the optimisation of expert e in slot s.
__device__ void fit_begin_expert(SlotDesc& d, const SlotMem& mem, LbfgsState& st, const BatchDev& b,
                                 const FitDev& f, int e, int ld) {
  slot_set_expert(d, mem, b, e, 0, ld);
  const int P = b.P;
  double x0[GPSAT_PMAX];
  int np = 0;
  for (int p = 0; p < P; ++p) {
    size_t q = (size_t)e * P + p;
    if (f.trainable[q]) x0[np++] = transform_inv(f.tkind[q], f.ta[q], f.tb[q], f.theta[q]);
  }
  lbfgs_init(&st, np, x0);
  slot_set_theta(d, f.theta + (size_t)e * P, b.i8_rho_max, b.i8_rho_low);
}

__global__ void fit_init_kernel(SlotDesc* slots, const SlotMem* mems, BatchDev b, FitDev f, int S, int ld) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  if (s < b.E) {
    fit_begin_expert(slots[s], mems[s], f.states[s], b, f, b.order[s], ld);
    atomicAdd(&f.counters[1], 1);
    atomicMax(&f.counters[2], slots[s].T);
    if (slots[s].Uq8_mem != nullptr && slots[s].Uq8 == nullptr) atomicAdd(&f.counters[4], 1);
  } else {
    slots[s].expert = -1;
  }
}

// theta at the optimiser's current trial point (fixed parameters keep their input value)
__device__ __forceinline__ void fit_theta_trial(const FitDev& f, const LbfgsState& st, int e, int P, double* theta) {
  int k = 0;
  for (int p = 0; p < P; ++p) {
    size_t q = (size_t)e * P + p;
    theta[p] = f.trainable[q] ? transform_fwd(f.tkind[q], f.ta[q], f.tb[q], st.xtrial[k++]) : f.theta[q];
  }
}

// Feed one evaluation (val = lml or elbo at theta, gtheta = its gradient w.r.t. theta) to the
// optimiser of expert e: chain rule to xi, priors, L-BFGS step.  When the optimiser stops the
// per-expert outputs are written (theta at the accepted iterate, objective, ...).  Returns the
// lbfgs_step code.
__device__ int fit_advance(const FitDev& f, LbfgsState& st, int e, int P, double val, const double* theta,
                           const double* gtheta) {
  double gxi[GPSAT_PMAX];
  {
    int k = 0;
    for (int p = 0; p < P; ++p) {
      size_t q = (size_t)e * P + p;
      if (f.trainable[q]) {
        gxi[k] = -gtheta[p] * transform_dfwd(f.tkind[q], f.ta[q], f.tb[q], st.xtrial[k]);
        ++k;
      }
    }
  }
  double logprior = 0.0;
  if (f.pkind) {
    int k = 0;
    for (int p = 0; p < P; ++p) {
      size_t q = (size_t)e * P + p;
      if (f.trainable[q]) {
        double dxi = 0.0;
        logprior += log_prior_xi(f.pkind[q], f.pa[q], f.pb[q], f.tkind[q], f.ta[q], f.tb[q], st.xtrial[k],
                                 theta[p], &dxi);
        gxi[k] -= dxi;
        ++k;
      }
    }
  }
  double fval = isfinite(val) ? -val - logprior : INFINITY;
  atomicAdd(&f.counters[3], 1);
  int rc = lbfgs_step(&st, fval, gxi, &f.opt);
  if (rc == LBFGS_CONTINUE) return rc;
  // retire: theta at the accepted iterate
  int k = 0;
  for (int p = 0; p < P; ++p) {
    size_t q = (size_t)e * P + p;
    if (f.trainable[q]) f.theta[q] = transform_fwd(f.tkind[q], f.ta[q], f.tb[q], st.x[k++]);
  }
  f.objective[e] = st.f;
  if (f.lml) {
    double lp = 0.0;
    if (f.pkind) {
      int kk = 0;
      for (int p = 0; p < P; ++p) {
        size_t q = (size_t)e * P + p;
        if (f.trainable[q]) {
          double dxi = 0.0;
          lp += log_prior_xi(f.pkind[q], f.pa[q], f.pb[q], f.tkind[q], f.ta[q], f.tb[q], st.x[kk], f.theta[q], &dxi);
          ++kk;
        }
      }
    }
    f.lml[e] = -st.f - lp;
  }
  f.n_evals[e] = st.nfev;
  f.n_iters[e] = st.iter;
  int status = GPSAT_STATUS_OK;
  if (rc == LBFGS_STOP_MAXITER || rc == LBFGS_STOP_MAXFUN) status = GPSAT_STATUS_MAXITER;
  else if (rc == LBFGS_ABNORMAL) status = GPSAT_STATUS_LINESEARCH;
  else if (rc == LBFGS_NOT_FINITE) status = GPSAT_STATUS_NOT_PD;
  f.status[e] = status;
  return rc;
}

// One optimiser step for every occupied slot (one thread per slot).
__global__ void lbfgs_kernel(SlotDesc* slots, const SlotMem* mems, BatchDev b, FitDev f, int S, int ld) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  SlotDesc& d = slots[s];
  if (d.expert < 0) return;
  const int P = b.P;
  LbfgsState& st = f.states[s];
  int e = d.expert;
  double theta[GPSAT_PMAX], gtheta[GPSAT_PMAX];
  fit_theta_trial(f, st, e, P, theta);
  double lml = slot_finish_eval(d, theta, gtheta, true);
  int rc = fit_advance(f, st, e, P, lml, theta, gtheta);
  if (rc != LBFGS_CONTINUE) {
    int idx = atomicAdd(&f.counters[0], 1);
    if (idx < b.E) fit_begin_expert(d, mems[s], st, b, f, b.order[idx], ld);
    else d.expert = -1;
  } else {
    fit_theta_trial(f, st, e, P, theta);
    // The low-rho tier carries noise of ~1e-10 |f| into the objective.  Far from the optimum that is nothing; in the
    // last iterations, where the line search tests decreases of a few ftol |f|, it costs spurious trial steps (c3 expert
    // 64: 49 instead of 25 evaluations to the same end point).  So the tier is offered only while the decrease the
    // line search expects from this trial, |stp g'd|, is above i8_tier_guard * max(|f|, 1): the evaluations that decide
    // convergence always run with all slice pairs.
    const bool far = fabs(st.stp * st.gd) > b.i8_tier_guard * fmax(fabs(st.f), 1.0);
    slot_set_theta(d, theta, b.i8_rho_max, far ? b.i8_rho_low : 0.0);
    *d.notpd = 0;
    *d.quad = 0.0;
  }
  if (d.expert >= 0) {
    atomicAdd(&f.counters[1], 1);
    atomicMax(&f.counters[2], d.T);
    if (d.Uq8_mem != nullptr && d.Uq8 == nullptr) atomicAdd(&f.counters[4], 1);
  }
}

// Plain evaluation passes (lml_grad / predict): experts order[base .. base+S) occupy the slots.
__global__ void eval_init_kernel(SlotDesc* slots, const SlotMem* mems, BatchDev b, const double* theta,
                                 int base, int S, int with_pred, int ld) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  if (base + s < b.E) {
    int e = b.order[base + s];
    slot_set_expert(slots[s], mems[s], b, e, with_pred, ld);
    slot_set_theta(slots[s], theta + (size_t)e * b.P, b.i8_rho_max, b.i8_rho_low);
  } else {
    slots[s].expert = -1;
  }
}

__global__ void eval_finish_kernel(const SlotDesc* slots, BatchDev b, const double* theta, double* lml,
                                   double* grad, int* status, int S) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const SlotDesc& d = slots[s];
  if (d.expert < 0) return;
  int e = d.expert;
  double gtheta[GPSAT_PMAX];
  double v = slot_finish_eval(d, theta + (size_t)e * b.P, gtheta, grad != nullptr);
  if (lml) lml[e] = v;
  if (grad) for (int p = 0; p < b.P; ++p) grad[(size_t)e * b.P + p] = gtheta[p];
  if (status) status[e] = (*d.notpd) ? GPSAT_STATUS_NOT_PD : GPSAT_STATUS_OK;
}

// Copy the dense augmented factor of slot 0 out (test hook).
__global__ void debug_extract_kernel(const SlotDesc* slots, double* Lout, double* Uout) {
  const SlotDesc& d = slots[0];
  const int N = d.n + 1;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < (size_t)N * N;
       idx += (size_t)gridDim.x * blockDim.x) {
    int r = (int)(idx / N), c = (int)(idx % N);
    int tr = r / GPSAT_NB, tc = c / GPSAT_NB;
    double lv = 0.0, uv = 0.0;
    if (tr == tc) {
      lv = d.Ld[(size_t)tr * GPSAT_NB * GPSAT_NB + (size_t)(r % GPSAT_NB) * GPSAT_NB + (c % GPSAT_NB)];
      uv = d.A[(size_t)r * d.ld + c];
    } else if (tr > tc) {
      lv = d.A[(size_t)r * d.ld + c];
    } else {
      uv = d.A[(size_t)r * d.ld + c];
    }
    Lout[idx] = lv;
    if (Uout) Uout[idx] = uv;
  }
}

#include "sgpr_fit.cuh"
#include "vff_fit.cuh"

}  // namespace gpsat

#include "engine_host.inl"
#include "engine_noalloc.inl"
