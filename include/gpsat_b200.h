/* gpsat_b200.h - C ABI of the B200-native local-expert GP engine (libgpsat_b200.so).
 *
 * Drop-in boundary for the per-expert hot path of akpetty/GPSat.  The reference is pure Python
 * and has no FFI of its own; the calls below replace, in batch over E experts, what one
 * GPSat model instance does per expert through GPflow/TensorFlow:
 *
 *   gpsat_gpr_lml_grad  <- GPflowGPRModel.get_objective_function_value
 *                          (GPSat/models/gpflow_models.py:334-337) and the loss+gradient closure
 *                          gpflow.optimizers.Scipy evaluates (gpflow_models.py:317-321)
 *   gpsat_gpr_fit       <- GPflowGPRModel.optimise_parameters (gpflow_models.py:290-329) for the
 *                          whole loop body of LocalExpertOI.run (GPSat/local_experts.py:1126-1132)
 *   gpsat_gpr_predict   <- GPflowGPRModel.predict (gpflow_models.py:186-273; predict_y/predict_f
 *                          at :229-230), called at local_experts.py:1157-1164
 *   gpsat_sgpr_*        <- GPflowSGPRModel (gpflow_models.py:666-901)
 *
 * Conventions
 *   - all floating point data is float64, C-contiguous; offsets are int64; flags are int32/uint8
 *   - experts are described raggedly: expert e owns rows off_n[e] .. off_n[e+1] of X / y and
 *     rows off_m[e] .. off_m[e+1] of Xs and of the prediction outputs
 *   - X, Xs are the *scaled* coordinates (already divided by coords_scale,
 *     GPSat/models/base_model.py:243) and y the centred/scaled observations (:244-245)
 *   - theta layout per expert: [lengthscale_1..lengthscale_D, kernel_variance, likelihood_variance],
 *     P = D + 2 values, constrained space
 *   - `mem` says where the data arrays live: GPSAT_MEM_HOST (the engine copies in and out on its
 *     own stream) or GPSAT_MEM_DEVICE (device pointers of the engine's device; nothing is copied).
 *     off_n / off_m are always host pointers.
 *   - every function returns 0 on success, non-zero on API misuse (2), a CUDA error (1) or an internal
 *     pipeline time-out (3; the outputs of that call are invalid); gpsat_last_error() gives the text.
 *     Numerical trouble is reported per expert in status[].
 *   - a handle is bound to one device and one stream and is not thread-safe; use one handle per
 *     host thread / GPU rank.  No global mutable state besides the last-error string.
 */
#ifndef GPSAT_B200_H
#define GPSAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPSAT_MEM_HOST 0
#define GPSAT_MEM_DEVICE 1

/* kernel ids (gpflow.kernels names accepted by GPflowGPRModel(kernel=...), gpflow_models.py:72-75) */
#define GPSAT_KERNEL_MATERN12 0
#define GPSAT_KERNEL_MATERN32 1
#define GPSAT_KERNEL_MATERN52 2
#define GPSAT_KERNEL_RBF 3

/* parameter transforms: theta = a + softplus(xi)            (gpflow.utilities.positive(lower=a))
 *                       theta = a + (b - a) * sigmoid(xi)   (tfp.bijectors.Sigmoid(low=a, high=b),
 *                                                            gpflow_models.py:490-494)            */
#define GPSAT_TRANSFORM_SOFTPLUS 0
#define GPSAT_TRANSFORM_SIGMOID 1

/* per-expert status */
#define GPSAT_STATUS_OK 0            /* converged (SciPy success == True) / evaluated            */
#define GPSAT_STATUS_NOT_PD 1        /* K + s2 I not positive definite at the (starting) point    */
#define GPSAT_STATUS_LINESEARCH 2    /* ABNORMAL_TERMINATION_IN_LNSRCH (success == False)         */
#define GPSAT_STATUS_MAXITER 3       /* maxiter / maxfun reached (success == False)               */
#define GPSAT_STATUS_BAD_INPUT 4     /* non-finite input                                          */

#define GPSAT_MAX_DIMS 6

/* prior families (extension; the reference never sets a prior).  GPflow convention:
 * training_loss = -(lml + sum_p [log p(theta_p) + log |d theta_p / d xi_p|]) for parameters that
 * carry a prior on their constrained value. */
#define GPSAT_PRIOR_NONE 0
#define GPSAT_PRIOR_NORMAL 1         /* a = mean, b = standard deviation  */
#define GPSAT_PRIOR_LOGNORMAL 2      /* a = mu,   b = sigma (of log theta) */
#define GPSAT_PRIOR_GAMMA 3          /* a = shape (concentration), b = rate */

typedef struct gpsat_engine gpsat_engine;

typedef struct {
  int32_t kernel;          /* GPSAT_KERNEL_*                       */
  int32_t D;               /* input dimensions, 1..GPSAT_MAX_DIMS  */
  int64_t E;               /* number of experts                    */
  const int64_t* off_n;    /* [E+1] host                           */
  const int64_t* off_m;    /* [E+1] host, may be NULL (no prediction rows) */
  const double* X;         /* [off_n[E]][D]                        */
  const double* y;         /* [off_n[E]]                           */
  const double* Xs;        /* [off_m[E]][D], may be NULL           */
} gpsat_batch;

typedef struct {
  double pgtol;            /* SciPy gtol,   default 1e-5                     */
  double ftol;             /* SciPy ftol,   default 2.220446049250313e-09    */
  int32_t maxiter;         /* default 10000 (gpflow_models.py:291)           */
  int32_t maxfun;          /* default 15000                                  */
  int32_t maxls;           /* default 20                                     */
  int32_t max_slots;       /* experts in flight per GPU, 0 = automatic       */
} gpsat_fit_options;

typedef struct {
  const int32_t* kind;     /* [E][P] GPSAT_PRIOR_*; same memory space as the other per-expert arrays */
  const double* a;         /* [E][P] */
  const double* b;         /* [E][P] */
} gpsat_priors;

typedef struct {
  int64_t kernel_launches; /* kernels launched by the last call      */
  int64_t ticks;           /* batched objective evaluations (sweeps) */
  int64_t evals;           /* per-expert objective evaluations       */
  int64_t slots;           /* experts in flight                      */
  double h2d_bytes;
  double d2h_bytes;
  double gpu_ms;           /* CUDA-event time of the last call, copies included */
  double sweep_flops;      /* sum over evaluations of n^3 (+ n^2 m for predict) */
  /* filled when profiling is on (gpsat_engine_set_profile): CUDA-event time per kernel type,
   * summed over the launches of the last call, and the launch counts */
  double ms_diag, ms_step, ms_lauum, ms_opt, ms_reduce;
  int64_t n_diag, n_step, n_lauum, n_opt, n_reduce;
} gpsat_stats;

int gpsat_version(void);
const char* gpsat_last_error(void);

/* Name of CUDA device `device` (what the reference stores in run_details.device through
 * BaseGPRModel.gpu_name, GPSat/models/base_model.py:279-300).  Returns 0 and a NUL-terminated
 * string in buf, non-zero if there is no such device. */
int gpsat_device_name(int device, char* buf, size_t len);

int gpsat_engine_create(int device, gpsat_engine** out);
int gpsat_engine_destroy(gpsat_engine* eng);
int gpsat_engine_stats(const gpsat_engine* eng, gpsat_stats* out);
/* cudaStream_t the engine launches on (so callers can record their own events on it). */
void* gpsat_engine_stream(const gpsat_engine* eng);
/* on != 0: bracket every kernel launch with CUDA events and report per-kernel times in stats. */
int gpsat_engine_set_profile(gpsat_engine* eng, int on);
/* How the K^-1 tile products of the gradient (the LAUUM piece of the sweep, the inverse GPflow's
 * autodiff forms implicitly at gpflow_models.py:317-321) are evaluated:
 *   GPSAT_LAUUM_I8 (default) / GPSAT_LAUUM_I8Q  error-free int8 slicing on the tcgen05 tensor path
 *   GPSAT_LAUUM_DMMA FP64 mma.sync
 * The environment variable GPSAT_B200_LAUUM=dmma selects the second at engine creation. */
#define GPSAT_LAUUM_DMMA 0
#define GPSAT_LAUUM_I8 1             /* default: one CTA per tile */
#define GPSAT_LAUUM_I8Q 2            /* persistent, warp-specialised; same arithmetic, bitwise the same results (slower so far) */
int gpsat_engine_set_lauum_path(gpsat_engine* eng, int path);
/* How the potrf tiles of the sweep (the Cholesky GPflow runs inside GPR.log_marginal_likelihood / predict_f,
 * gpflow_models.py:229-230, 337) accumulate sum_k L_rk L_ik^T:
 *   GPSAT_STEP_DMMA FP64 mma.sync
 *   GPSAT_STEP_I8   8 exact int8 slices per value with a-priori row scales on the tcgen05 tensor path
 *                   (error at the level of the FP64 accumulation error)
 * The environment variable GPSAT_B200_STEP=i8|dmma selects the path at engine creation. */
#define GPSAT_STEP_DMMA 0
#define GPSAT_STEP_I8 1
int gpsat_engine_set_step_path(gpsat_engine* eng, int path);
/* Precision rule of the int8-sliced arithmetic.  The slices are taken against a-priori row bounds
 * (sqrt(var + noise) for L, 1 / sqrt(noise) for U = L^-T), so relative to the entries themselves the accuracy
 * falls with rho = (kernel_variance + likelihood_variance) / likelihood_variance.  Every objective evaluation
 * whose rho exceeds rho_max runs its k-sums and K^-1 products as FP64 mma.sync instead - chosen per expert and
 * per evaluation on the device, because the optimiser moves theta.  rho_max = 0 turns the rule off (int8-sliced
 * at any conditioning; for measurements only).  Default: gpsat_engine_precision_rule_default(), from the envelope
 * measured against an extended-precision oracle (profiles/r02_precision_envelope.md); the environment variable
 * GPSAT_B200_RHO_MAX overrides it at engine creation.  GPflow's own floor of the likelihood variance is 1e-6
 * (gpflow_models.py:404-411), i.e. rho up to 1e7 is reachable through the reference's API. */
int gpsat_engine_set_precision_rule(gpsat_engine* eng, double rho_max);
double gpsat_engine_precision_rule_default(void);
/* Low-rho tier of the same rule (no counterpart in the reference): while rho <= rho_low the a-priori scales sit within
 * a few bits of the entries, and the k-sums of the Cholesky / inverse sweep drop their last group of slice pairs
 * (2^-49 of the scale product; 28 instead of 36 int8 MMAs per 32-wide slab).  The squared-exponential kernel uses
 * 3/8 of rho_low (its matrices are worse conditioned at equal rho).  rho_low = 0 turns the tier off.  Default gpsat_engine_precision_low_tier_default() from the same envelope
 * measurement; the environment variable GPSAT_B200_RHO_LOW overrides it at engine creation. */
int gpsat_engine_set_precision_low_tier(gpsat_engine* eng, double rho_low);
double gpsat_engine_precision_low_tier_default(void);
/* A bounded barrier wait of the int8 tensor pipeline that gives up makes the call that ran it FAIL (non-zero
 * return, gpsat_last_error() says so; per-expert outputs of that call must not be used).  out = number of calls
 * that failed that way since the engine was created. */
int gpsat_engine_i8_timeouts(gpsat_engine* eng, int* out);
/* Test hook: number of polls after which such a wait gives up (default 2^22, i.e. seconds).  polls <= 0: every
 * wait completes normally and is then reported as timed out (exercises the failure path of the calls without
 * abandoning a pipeline that has copies in flight).  Process-wide. */
int gpsat_engine_debug_set_i8_spin_limit(gpsat_engine* eng, int polls);
/* Cycle counts per phase of the int8-path gradient kernel summed over its CTAs (all zero unless the library was
 * built with -DGPSAT_I8_PHASES; a development aid): prologue, pass-0 MMAs, drain 0, pass-1 MMAs, drain 1,
 * Q tile store, gradient reduction, tiles. */
int gpsat_engine_i8_phases(gpsat_engine* eng, uint64_t* out8, int reset);
/* Same for step_kernel (library built with -DGPSAT_STEP_PHASES): potrf tiles in out32[0..15], trtri tiles in
 * out32[16..31]; see g_step_phase in gpsat_b200/csrc/step_kernel.cuh for the phase list. */
int gpsat_engine_step_phases(gpsat_engine* eng, uint64_t* out32, int reset);
void gpsat_fit_options_default(gpsat_fit_options* opt);

/* lml[e] = log marginal likelihood, grad[e][P] = d lml / d theta (constrained space). */
int gpsat_gpr_lml_grad(gpsat_engine* eng, const gpsat_batch* batch, int mem, const double* theta,
                       double* lml, double* grad, int32_t* status);

/* Batched L-BFGS in unconstrained space.  theta_io: start values in, optimum out.
 * tkind/ta/tb: transform per parameter; trainable: 1 = optimise, 0 = keep fixed.
 * priors: may be NULL (no priors).  objective: the minimised value -(lml + log prior) at the
 * optimum, i.e. gpflow's training_loss; lml: the log marginal likelihood alone (may be NULL);
 * n_evals / n_iters: SciPy nfev / nit equivalents. */
int gpsat_gpr_fit(gpsat_engine* eng, const gpsat_batch* batch, int mem, double* theta_io,
                  const int32_t* tkind, const double* ta, const double* tb, const uint8_t* trainable,
                  const gpsat_fit_options* opt, const gpsat_priors* priors, double* objective, double* lml,
                  int32_t* n_evals, int32_t* n_iters, int32_t* status);

/* Predictive mean / variance at batch->Xs for fixed theta: fmean = f*, fvar = f*_var,
 * yvar = f*_var + likelihood variance; lml[e] as a by-product of the same factorisation. */
int gpsat_gpr_predict(gpsat_engine* eng, const gpsat_batch* batch, int mem, const double* theta,
                      double* fmean, double* fvar, double* yvar, double* lml, int32_t* status);

/* Same, plus the full posterior covariance of every expert's prediction points
 * (GPR.predict_f(full_cov=True), gpflow_models.py:230, 245-263): expert e owns a dense row-major
 * m_e x m_e block at fcov + sum_{e' < e} m_e'^2.  y_cov = f_cov + likelihood_variance * I. */
int gpsat_gpr_predict_cov(gpsat_engine* eng, const gpsat_batch* batch, int mem, const double* theta,
                          double* fmean, double* fvar, double* yvar, double* fcov, double* lml,
                          int32_t* status);

/* ---- sparse GPR (Titsias 2009): GPflowSGPRModel, GPSat/models/gpflow_models.py:666-901 ----------
 * Inducing points are ragged like the observations: expert e owns rows off_z[e] .. off_z[e+1] of
 * Z ([sum M_e][D], divided by coords_scale like X).  off_z is always a HOST array; Z and theta follow
 * `mem`.  theta layout as for GPR: [lengthscales (D), kernel variance, likelihood variance].
 *
 * gpsat_sgpr_elbo: elbo[e] = gpflow.models.SGPR.elbo() (jitter 1e-6 on Kuu) — the value behind
 * GPflowSGPRModel.get_objective_function_value (gpflow_models.py:860-862, which reports +elbo).
 * grad (may be NULL): d elbo / d theta [E][P] in constrained space with Z fixed — the gradient
 * TensorFlow autodiff hands to gpflow.optimizers.Scipy (gpflow_models.py:317-321 inherited).
 * status[e] = 1 when Kuu + jitter I or B is not positive definite (elbo = -inf). */
int gpsat_sgpr_elbo(gpsat_engine* eng, const gpsat_batch* batch, const int64_t* off_z, const double* Z,
                    int mem, const double* theta, double* elbo, double* grad, int32_t* status);

/* gpsat_sgpr_elbo plus the gradient with respect to the inducing points: gradz [sum M_e][D] in the layout of Z
 * (follows `mem`), d elbo / d Z in the scaled coordinates Z is given in.  This is what TensorFlow differentiates once
 * GPflowSGPRModel.optimise_parameters(train_inducing_points=True) has made inducing_variable.Z trainable
 * (gpflow_models.py:864-899, set_trainable at :896-897); grad and gradz must both be non-NULL. */
int gpsat_sgpr_elbo_gradz(gpsat_engine* eng, const gpsat_batch* batch, const int64_t* off_z, const double* Z,
                          int mem, const double* theta, double* elbo, double* grad, double* gradz,
                          int32_t* status);

/* SGPR.predict_f / predict_y (full_cov = False) at batch->Xs, as GPflowSGPRModel.predict uses them
 * (gpflow_models.py:229-230 inherited); elbo (may be NULL) comes out of the same factorisation. */
int gpsat_sgpr_predict(gpsat_engine* eng, const gpsat_batch* batch, const int64_t* off_z, const double* Z,
                       int mem, const double* theta, double* fmean, double* fvar, double* yvar,
                       double* elbo, int32_t* status);

/* Same, plus the full posterior covariance of every expert's prediction points (SGPR.predict_f(full_cov=True), reached
 * through GPflowSGPRModel.predict(full_cov=True), gpflow_models.py:230, 245-263): expert e owns a dense row-major
 * m_e x m_e block at fcov + sum_{e' < e} m_e'^2.  y_cov = f_cov + likelihood_variance * I. */
int gpsat_sgpr_predict_cov(gpsat_engine* eng, const gpsat_batch* batch, const int64_t* off_z, const double* Z,
                           int mem, const double* theta, double* fmean, double* fvar, double* yvar, double* fcov,
                           double* elbo, int32_t* status);

/* Replaces GPflowSGPRModel.optimise_parameters (gpflow_models.py:290-329 inherited: L-BFGS-B on
 * training_loss = -elbo, inducing points fixed, :896-897) for all experts.  Arguments as
 * gpsat_gpr_fit; objective[e] = -(elbo + log prior) at the optimum, elbo[e] (may be NULL) = elbo. */
int gpsat_sgpr_fit(gpsat_engine* eng, const gpsat_batch* batch, const int64_t* off_z, const double* Z,
                   int mem, double* theta_io, const int32_t* tkind, const double* ta, const double* tb,
                   const uint8_t* trainable, const gpsat_fit_options* opt, const gpsat_priors* priors,
                   double* objective, double* elbo, int32_t* n_evals, int32_t* n_iters, int32_t* status);

/* ---- variational Fourier features: GPflowVFFModel -> GPR_kron, GPSat/vff.py:564-675,
 * GPSat/models/vff_model.py:14-264 -------------------------------------------------------------
 * One 1-D Matern kernel per input dimension (batch->kernel in {MATERN12, MATERN32, MATERN52},
 * D <= 3), cos / sin inducing features with frequencies 2 pi k / (b - a), k = 0 .. nfreq - 1
 * (ms = arange(num_inducing_features), vff_model.py:201-204) on the domain [a, b] of each
 * dimension (scaled coordinates, vff_model.py:180-199).  The arrays of gpsat_vff_domain are HOST
 * arrays [E][D].  theta layout: [lengthscale_1..D, variance_1..D, likelihood variance], P = 2 D + 1
 * (each 1-D kernel has its own variance, get_kernel_variance is their product, vff_model.py:217-218).
 * Observations must lie strictly inside (a, b) (vff.py:602-603; checked for host batches).
 * Prediction points outside [a, b] use the kernel's edge features (vff.py:472-497); Matern52 has
 * none (:498-511) - the caller must keep them inside. */
typedef struct gpsat_vff_domain {
  const double* a;        /* [E][D] lower edge of the feature domain */
  const double* b;        /* [E][D] upper edge */
  const int32_t* nfreq;   /* [E][D] number of frequencies per dimension (2 nfreq - 1 <= 128 features) */
} gpsat_vff_domain;

/* elbo[e] = GPR_kron.elbo() (vff.py:612-643; GPflowVFFModel.get_objective_function_value reports it
 * as is, vff_model.py:209-211); grad (may be NULL) = d elbo / d theta [E][P]. */
int gpsat_vff_elbo(gpsat_engine* eng, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem,
                   const double* theta, double* elbo, double* grad, int32_t* status);

/* GPR_kron.predict_f (vff.py:645-675, full_cov = False) and predict_y at batch->Xs. */
int gpsat_vff_predict(gpsat_engine* eng, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem,
                      const double* theta, double* fmean, double* fvar, double* yvar, double* elbo,
                      int32_t* status);

/* The inherited optimise_parameters (gpflow_models.py:290-329) on -elbo; arguments as gpsat_gpr_fit
 * with P = 2 D + 1. */
int gpsat_vff_fit(gpsat_engine* eng, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem,
                  double* theta_io, const int32_t* tkind, const double* ta, const double* tb,
                  const uint8_t* trainable, const gpsat_fit_options* opt, const gpsat_priors* priors,
                  double* objective, double* elbo, int32_t* n_evals, int32_t* n_iters, int32_t* status);

/* ---- post-processing reductions (the "next" rows of the local-expert workflow) ------------------
 * gaussian_2d_weight, the kernel of smooth_hyperparameters (GPSat/postprocessing.py:22-52, 96-395):
 * for F fields given at the same n positions (vals: [F][n]) and n0 reference positions (out: [F][n0]),
 *   out[f][i] = sum_j w_ij vals[f][j] / sum_j w_ij over non-NaN vals[f][j],
 *   w_ij = exp(-(((x_j - x0_i) / lx)^2 + ((y_j - y0_i) / ly)^2) / 2);  NaN when the weights sum to 0. */
int gpsat_gaussian_smooth(gpsat_engine* eng, int mem, int64_t n0, const double* x0, const double* y0,
                          int64_t n, int64_t F, const double* x, const double* y, double lx, double ly,
                          const double* vals, double* out);

/* glue_local_predictions_1d / _2d (GPSat/postprocessing.py:462-593): rows sorted by prediction
 * location, group g = rows seg[g] .. seg[g+1) (seg: HOST array [G+1]); pred, xprt: [ndim][n]
 * prediction / expert coordinates, sigma [n] = inference_radius / R, vals [V][n]:
 *   w_i = prod_d normal_pdf(pred[d][i]; xprt[d][i], sigma[i]),  out[v][g] = sum w vals[v] / sum w. */
int gpsat_glue_predictions(gpsat_engine* eng, int mem, int64_t n, int64_t G, const int64_t* seg, int ndim,
                           const double* pred, const double* xprt, const double* sigma, int64_t V,
                           const double* vals, double* out);

/* ---- batched local selection (DataLoader.local_data_select, GPSat/dataloader.py:2354-2447, and the
 * max_dist filter of PredictionLocations, GPSat/prediction_locations.py:18-43) for all experts ------
 * data [N][C]: point table (row-major, host); locs [E][CL]: expert locations (host).  A row is kept
 * for expert e when every condition holds:
 *   kind 0:  data[col[0]] <op> locs[e][loc_col[0]] + val          (op: 0 >=, 1 >, 2 ==, 3 <, 4 <=)
 *   kind 1:  sum_k (data[col[k]] - locs[e][loc_col[k]])^2 <op> val^2   (ball query: op 4; strict: op 3)
 * Two calls: with idx = off = NULL the matches per expert are counted into counts[E]; with
 * off[E+1] (exclusive prefix sums of counts) and idx[off[E]] the matching row indices are written in
 * increasing order per expert.  All arrays are HOST arrays. */
typedef struct gpsat_select_cond {
  int32_t kind, op, ncol;
  int32_t col[3];
  int32_t loc_col[3];
  double val;
} gpsat_select_cond;

int gpsat_local_select(gpsat_engine* eng, int64_t N, int32_t C, const double* data, int64_t E, int32_t CL,
                       const double* locs, int32_t ncond, const gpsat_select_cond* conds, int64_t* counts,
                       const int64_t* off, int64_t* idx);

/* ---- allocation-free entry point -------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * The counting pass of gpsat_local_select (counts[e] = number of rows of the point table that satisfy every
 * condition for expert e; same conditions, same kernels, same results) without any allocation: the function calls
 * no cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine.
 *   - data [N][C], locs [E][CL] and counts [E] are DEVICE pointers of the engine's device, owned by the caller
 *     (nothing is copied or staged; host pointers are rejected); conds is a host array as for gpsat_local_select.
 *   - the per-(expert, 1024-point chunk) counters live in a fixed device array inside the library:
 *     E * ceil(N / 1024) must not exceed GPSAT_NOALLOC_MAX_CHUNK_COUNTS (2^20 counters = 4 MiB, resident on every
 *     GPU that loads the library until the process exits); larger problems are rejected with return code 2 and a
 *     message - use gpsat_local_select for those.
 *   - every call, from every engine and thread on a device, shares that one array, and it keeps the last call's
 *     contents (harmless: each call overwrites every counter it reads, nothing needs to start at zero).  Calls
 *     that overlap in time on the same GPU RACE on it: serialise them (one call at a time per device).
 *   - N >= 1, 1 <= E < 65536.  The function returns after the engine's stream has drained, like the other calls. */
#define GPSAT_NOALLOC_MAX_CHUNK_COUNTS (1 << 20)
int gpsat_local_select_count_noalloc(gpsat_engine* eng, int64_t N, int32_t C, const double* data, int64_t E,
                                     int32_t CL, const double* locs, int32_t ncond, const gpsat_select_cond* conds,
                                     int64_t* counts);

/* Test hook (single expert, host memory): dense factor of the augmented matrix.
 * Lout [(n+1)*(n+1)] row-major lower: rows < n = chol(K + s2 I), row n = (L^-1 y)^T, 1.
 * Uout [(n+1)*(n+1)] row-major upper: L_aug^-T (may be NULL). */
int gpsat_gpr_factor_debug(gpsat_engine* eng, const gpsat_batch* batch, const double* theta,
                           double* Lout, double* Uout);

#ifdef __cplusplus
}
#endif
#endif /* GPSAT_B200_H */
