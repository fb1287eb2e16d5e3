// Slot management and the optimiser tick of the sparse-GPR path (included by engine.cu, inside
// namespace gpsat, after the exact-GPR fit kernels whose L-BFGS plumbing it shares).
// Replaces GPflowSGPRModel.optimise_parameters (GPSat/models/gpflow_models.py:290-329 inherited:
// gpflow.optimizers.Scipy().minimize(training_loss = -elbo), inducing points fixed :896-897).

__global__ void sgpr_eval_init_kernel(SgprDesc* descs, SlotDesc* sk, SlotDesc* sb, double* ws, SgprLay L, BatchDev b,
                                      SgprIn in, const double* theta, int base, int S, int with_pred) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  if (base + s < b.E) {
    const int e = b.order[base + s];
    sgpr_set_expert(descs[s], sk[s], sb[s], ws + L.total * s, L, b, in, e, with_pred);
    sgpr_set_theta(descs[s], sk[s], sb[s], theta + (size_t)e * b.P);
  } else {
    descs[s].expert = -1; sk[s].expert = -1; sb[s].expert = -1;
  }
}

__global__ void sgpr_eval_finish_kernel(const SgprDesc* descs, const SlotDesc* sk, const SlotDesc* sb, BatchDev b,
                                        double* elbo, double* grad, int* status, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const SgprDesc& d = descs[s];
  if (d.expert < 0) return;
  double gtheta[GPSAT_PMAX];
  const double v = sgpr_finish_eval(d, sk[s], sb[s], gtheta, grad != nullptr);
  if (elbo) elbo[d.expert] = v;
  if (grad) for (int p = 0; p < b.P; ++p) grad[(size_t)d.expert * b.P + p] = gtheta[p];
  if (status) status[d.expert] = isfinite(v) ? GPSAT_STATUS_OK : GPSAT_STATUS_NOT_PD;
}

__device__ void sgpr_fit_begin_expert(SgprDesc& d, SlotDesc& k, SlotDesc& bs, double* base, const SgprLay& L,
                                      LbfgsState& st, const BatchDev& b, const SgprIn& in, const FitDev& f, int e) {
  sgpr_set_expert(d, k, bs, base, L, b, in, e, 0);
  const int P = b.P;
  double x0[GPSAT_PMAX];
  int np = 0;
  for (int p = 0; p < P; ++p) {
    const size_t q = (size_t)e * P + p;
    if (f.trainable[q]) x0[np++] = transform_inv(f.tkind[q], f.ta[q], f.tb[q], f.theta[q]);
  }
  lbfgs_init(&st, np, x0);
  sgpr_set_theta(d, k, bs, f.theta + (size_t)e * P);
}

__global__ void sgpr_fit_init_kernel(SgprDesc* descs, SlotDesc* sk, SlotDesc* sb, double* ws, SgprLay L, BatchDev b,
                                     SgprIn in, FitDev f, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  if (s < b.E) {
    sgpr_fit_begin_expert(descs[s], sk[s], sb[s], ws + L.total * s, L, f.states[s], b, in, f, b.order[s]);
    atomicAdd(&f.counters[1], 1);
  } else {
    descs[s].expert = -1; sk[s].expert = -1; sb[s].expert = -1;
  }
}

// One optimiser step for every occupied slot (one thread per slot); the minimised value is -elbo.
__global__ void sgpr_lbfgs_kernel(SgprDesc* descs, SlotDesc* sk, SlotDesc* sb, double* ws, SgprLay L, BatchDev b,
                                  SgprIn in, FitDev f, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  SgprDesc& d = descs[s];
  if (d.expert < 0) return;
  const int P = b.P, e = d.expert;
  LbfgsState& st = f.states[s];
  double theta[GPSAT_PMAX], gtheta[GPSAT_PMAX];
  fit_theta_trial(f, st, e, P, theta);
  const double elbo = sgpr_finish_eval(d, sk[s], sb[s], gtheta, true);
  const int rc = fit_advance(f, st, e, P, elbo, theta, gtheta);
  if (rc != LBFGS_CONTINUE) {
    const int idx = atomicAdd(&f.counters[0], 1);
    if (idx < b.E) sgpr_fit_begin_expert(d, sk[s], sb[s], ws + L.total * s, L, st, b, in, f, b.order[idx]);
    else { d.expert = -1; sk[s].expert = -1; sb[s].expert = -1; }
  } else {
    fit_theta_trial(f, st, e, P, theta);
    sgpr_set_theta(d, sk[s], sb[s], theta);
  }
  if (d.expert >= 0) atomicAdd(&f.counters[1], 1);
}
