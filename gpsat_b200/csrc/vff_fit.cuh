// Slot management and the optimiser tick of the VFF path (included by engine.cu inside namespace
// gpsat).  Replaces GPflowVFFModel's inherited optimise_parameters / get_objective_function_value /
// predict (GPSat/models/vff_model.py:209-211, gpflow_models.py:186-329) for a batch of experts.

struct VffLay {
  size_t feat, featS, kuf, part, phi, pm, kust, dimw, dims, gp, a, side, total, tile_d;
  int Tmax, lgd;
};

struct VffIn {
  const double* a;       // [E][D] device
  const double* b;       // [E][D]
  const int* nfreq;      // [E][D]
};

__device__ __forceinline__ void vff_set_expert(VffDesc& d, SgprDesc& sd, SlotDesc& s, double* base, const VffLay& L,
                                               const BatchDev& bd, const VffIn& in, int e, int with_pred) {
  const int D = bd.D;
  d.expert = e; d.fresh = 1;
  d.n = (int)(bd.off_n[e + 1] - bd.off_n[e]);
  d.m = (with_pred && bd.off_m) ? (int)(bd.off_m[e + 1] - bd.off_m[e]) : 0;
  d.D = D; d.kind = bd.kind;
  int M = 1;
  for (int dd = 0; dd < D; ++dd) {
    d.nf[dd] = in.nfreq[(size_t)e * D + dd];
    d.Md[dd] = 2 * d.nf[dd] - 1;
    d.a[dd] = in.a[(size_t)e * D + dd];
    d.b[dd] = in.b[(size_t)e * D + dd];
    M *= d.Md[dd];
  }
  for (int dd = D; dd < VFF_DMAX; ++dd) { d.nf[dd] = 1; d.Md[dd] = 1; d.a[dd] = 0.0; d.b[dd] = 1.0; }
  d.M = M;
  d.npad = (d.n + NB - 1) / NB * NB;
  d.mpad = (max(d.m, 1) + NB - 1) / NB * NB;
  d.Mp = (M + 1 + NB - 1) / NB * NB;
  d.X = bd.X + bd.off_n[e] * D; d.y = bd.y + bd.off_n[e];
  d.Xs = d.m > 0 ? bd.Xs + bd.off_m[e] * D : nullptr;
  d.feat = base + L.feat; d.featS = base + L.featS; d.Kuf = base + L.kuf; d.part = base + L.part;
  d.Phi = base + L.phi; d.Pm = base + L.pm; d.KusT = base + L.kust; d.dimw = base + L.dimw;
  d.dims = base + L.dims; d.gp = base + L.gp;
  // view of the static SYRK for the sparse-path kernels: Phi' = Kuf Kuf^T
  sd.expert = e; sd.n = d.n; sd.m = 0; sd.M = M; sd.D = D; sd.kind = d.kind;
  sd.npad = d.npad; sd.mpad = d.mpad; sd.Mp = d.Mp;
  sd.Aq = d.Kuf; sd.part = d.part; sd.Bm = d.Phi;
  // the slot that factors P'
  double* side = base + L.side;
  s.expert = e;
  s.n = M; s.m = d.m;
  s.T = (M + 1 + NB - 1) / NB;
  s.TR = (M + 1 + d.m + NB - 1) / NB;
  s.ld = s.T * NB;
  s.kind = GPSAT_PRECOMPUTED; s.D = 0;
  s.X = nullptr; s.Xs = nullptr;
  s.y = d.Pm + (size_t)M * d.Mp;
  s.A = base + L.a;
  s.Wd = side; s.Ld = side + L.tile_d; s.logdet = side + 2 * L.tile_d;
  s.quad = s.logdet + L.lgd;
  s.gpart = nullptr;
  s.notpd = reinterpret_cast<int*>(s.quad + 2);
  s.pre = d.Pm; s.pre_s = d.KusT; s.ldpre = d.Mp;
  s.Uq = nullptr; s.uex = nullptr; s.Lq = nullptr; s.Uq8 = nullptr; s.ynorm2 = 0.0;
  s.var = 0.0; s.noise = 0.0;
}

__device__ __forceinline__ void vff_set_theta(VffDesc& d, SlotDesc& s, const double* theta) {
  for (int dd = 0; dd < d.D; ++dd) { d.ell[dd] = theta[dd]; d.var[dd] = theta[d.D + dd]; }
  for (int dd = d.D; dd < VFF_DMAX; ++dd) { d.ell[dd] = 1.0; d.var[dd] = 1.0; }
  d.noise = theta[2 * d.D];
  *s.quad = 0.0; *s.notpd = 0;
}

// after a tick: the static quantities of every occupied slot exist
__device__ __forceinline__ void vff_mark_built(VffDesc& d, SgprDesc& sd) { d.fresh = 0; sd.expert = -1; }

__global__ void vff_eval_init_kernel(VffDesc* descs, SgprDesc* sds, SlotDesc* slots, double* ws, VffLay L, BatchDev b,
                                     VffIn in, const double* theta, int base, int S, int with_pred) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  if (base + s < b.E) {
    const int e = b.order[base + s];
    vff_set_expert(descs[s], sds[s], slots[s], ws + L.total * s, L, b, in, e, with_pred);
    vff_set_theta(descs[s], slots[s], theta + (size_t)e * b.P);
  } else {
    descs[s].expert = -1; sds[s].expert = -1; slots[s].expert = -1;
  }
}

__global__ void vff_eval_finish_kernel(const VffDesc* descs, const SlotDesc* slots, BatchDev b, double* elbo,
                                       double* grad, int* status, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  const VffDesc& d = descs[s];
  if (d.expert < 0) return;
  double gtheta[GPSAT_PMAX];
  const double v = vff_finish_eval(d, slots[s], gtheta, grad != nullptr);
  if (elbo) elbo[d.expert] = v;
  if (grad) for (int p = 0; p < b.P; ++p) grad[(size_t)d.expert * b.P + p] = gtheta[p];
  if (status) status[d.expert] = isfinite(v) ? GPSAT_STATUS_OK : GPSAT_STATUS_NOT_PD;
}

__device__ void vff_fit_begin_expert(VffDesc& d, SgprDesc& sd, SlotDesc& s, double* base, const VffLay& L,
                                     LbfgsState& st, const BatchDev& b, const VffIn& in, const FitDev& f, int e) {
  vff_set_expert(d, sd, s, base, L, b, in, e, 0);
  const int P = b.P;
  double x0[GPSAT_PMAX];
  int np = 0;
  for (int p = 0; p < P; ++p) {
    const size_t q = (size_t)e * P + p;
    if (f.trainable[q]) x0[np++] = transform_inv(f.tkind[q], f.ta[q], f.tb[q], f.theta[q]);
  }
  lbfgs_init(&st, np, x0);
  vff_set_theta(d, s, f.theta + (size_t)e * P);
}

__global__ void vff_fit_init_kernel(VffDesc* descs, SgprDesc* sds, SlotDesc* slots, double* ws, VffLay L, BatchDev b,
                                    VffIn in, FitDev f, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  if (s < b.E) {
    vff_fit_begin_expert(descs[s], sds[s], slots[s], ws + L.total * s, L, f.states[s], b, in, f, b.order[s]);
    atomicAdd(&f.counters[1], 1);
  } else {
    descs[s].expert = -1; sds[s].expert = -1; slots[s].expert = -1;
  }
}

// One optimiser step for every occupied slot (one thread per slot); the minimised value is -elbo.
__global__ void vff_lbfgs_kernel(VffDesc* descs, SgprDesc* sds, SlotDesc* slots, double* ws, VffLay L, BatchDev b,
                                 VffIn in, FitDev f, int S) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  VffDesc& d = descs[s];
  if (d.expert < 0) return;
  const int P = b.P, e = d.expert;
  LbfgsState& st = f.states[s];
  double theta[GPSAT_PMAX], gtheta[GPSAT_PMAX];
  fit_theta_trial(f, st, e, P, theta);
  const double elbo = vff_finish_eval(d, slots[s], gtheta, true);
  const int rc = fit_advance(f, st, e, P, elbo, theta, gtheta);
  if (rc != LBFGS_CONTINUE) {
    const int idx = atomicAdd(&f.counters[0], 1);
    if (idx < b.E) vff_fit_begin_expert(d, sds[s], slots[s], ws + L.total * s, L, st, b, in, f, b.order[idx]);
    else { d.expert = -1; sds[s].expert = -1; slots[s].expert = -1; }
  } else {
    fit_theta_trial(f, st, e, P, theta);
    vff_set_theta(d, slots[s], theta);
    vff_mark_built(d, sds[s]);
  }
  if (d.expert >= 0) atomicAdd(&f.counters[1], 1);
}
