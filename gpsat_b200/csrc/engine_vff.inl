// Host orchestration of the VFF path (GPflowVFFModel -> GPR_kron, GPSat/vff.py:564-675).
// Included at the end of engine_sgpr.inl (global namespace, using namespace gpsat).

static VffLay vff_layout(int D, int npad, int mpad, int Mp, int TRmax, int nsplit, bool with_pred) {
  VffLay L{};
  const size_t mm = (size_t)Mp * Mp;
  const int MT = Mp / NB;
  L.Tmax = MT;
  L.lgd = align_up(MT, 2);
  L.tile_d = (size_t)MT * NB * NB;
  size_t o = 0;
  auto take = [&](size_t& field, size_t n) { field = o; o += (n + 1) & ~(size_t)1; };
  take(L.feat, (size_t)D * VFF_MD_MAX * npad);
  take(L.featS, with_pred ? (size_t)D * VFF_MD_MAX * mpad : 0);
  take(L.kuf, (size_t)Mp * npad);
  take(L.part, (size_t)nsplit * mm);
  take(L.phi, mm); take(L.pm, mm);
  take(L.kust, with_pred ? (size_t)mpad * Mp : 0);
  take(L.dimw, (size_t)D * VFF_NMAT * VFF_MD_MAX * VFF_MD_MAX);
  take(L.dims, 4 * VFF_DMAX);
  take(L.gp, (size_t)MT * (MT + 1) / 2 * VFF_NPART);
  take(L.a, (size_t)TRmax * NB * Mp);
  take(L.side, 2 * L.tile_d + L.lgd + 4);
  L.total = o;
  return L;
}

struct VffCall {
  BatchDev bd;
  VffIn in;
  VffLay L;
  int S, nsplit, npad, mpad, Mp, TRmax, D;
  double* ws;
  VffDesc* descs;
  SgprDesc* sds;
  SlotDesc* slots;
};

static int vff_setup(Engine& g, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem, bool with_pred,
                     int max_slots, VffCall& c) {
  const int64_t E = batch->E;
  const int D = batch->D;
  REQUIRE(batch->kernel == GPSAT_MATERN12 || batch->kernel == GPSAT_MATERN32 || batch->kernel == GPSAT_MATERN52,
          "VFF is defined for Matern12 / Matern32 / Matern52 only (vff.py:567-568)");
  REQUIRE(D <= VFF_DMAX, "VFF supports at most 3 input dimensions");
  REQUIRE(dom && dom->a && dom->b && dom->nfreq, "vff domain is NULL");
  int64_t Mmax = 0, MRmax = 0;
  const bool host = mem == GPSAT_MEM_HOST;
  if (engine_load_batch(g, batch, mem, with_pred, c.bd)) return 1;
  for (int64_t e = 0; e < E; ++e) {
    int64_t M = 1;
    for (int dd = 0; dd < D; ++dd) {
      const int nf = dom->nfreq[e * D + dd];
      REQUIRE(nf >= 1 && 2 * nf - 1 <= VFF_MD_MAX, "nfreq out of range [1, 128]");
      REQUIRE(dom->a[e * D + dd] < dom->b[e * D + dd], "vff domain needs a < b");
      M *= 2 * nf - 1;
    }
    REQUIRE(M < (1 << 15), "too many inducing features (product over dimensions must stay below 32768)");
    Mmax = std::max(Mmax, M);
    MRmax = std::max(MRmax, M + 1 + g.hm[e]);
    if (host) {
      // GPR_kron.__init__ asserts the data lie strictly inside the domain (vff.py:602-603)
      const double* X = batch->X + batch->off_n[e] * D;
      for (int64_t i = 0; i < g.hn[e]; ++i)
        for (int dd = 0; dd < D; ++dd)
          REQUIRE(X[i * D + dd] > dom->a[e * D + dd] && X[i * D + dd] < dom->b[e * D + dd],
                  "observation outside the VFF domain (a, b)");
    }
  }
  c.bd.P = 2 * D + 1;
  c.D = D;
  int64_t nmax = 0, mmax = 0;
  for (int64_t e = 0; e < E; ++e) { nmax = std::max(nmax, g.hn[e]); mmax = std::max(mmax, g.hm[e]); }
  c.npad = align_up((int)nmax, NB);
  c.mpad = align_up((int)std::max<int64_t>(mmax, 1), NB);
  c.Mp = align_up((int)Mmax + 1, NB);
  c.TRmax = (int)((MRmax + NB - 1) / NB);
  // domain arrays: [a | b] doubles then nfreq ints, in g.bmisc
  const size_t ED = (size_t)E * D;
  if (g.bmisc.ensure(ED * 20)) return 1;
  char* mb = g.bmisc.as<char>();
  CUDA_TRY(cudaMemcpyAsync(mb, dom->a, ED * 8, cudaMemcpyHostToDevice, g.stream));
  CUDA_TRY(cudaMemcpyAsync(mb + ED * 8, dom->b, ED * 8, cudaMemcpyHostToDevice, g.stream));
  CUDA_TRY(cudaMemcpyAsync(mb + ED * 16, dom->nfreq, ED * 4, cudaMemcpyHostToDevice, g.stream));
  g.stats.h2d_bytes += (double)ED * 20;
  c.in.a = reinterpret_cast<const double*>(mb);
  c.in.b = reinterpret_cast<const double*>(mb + ED * 8);
  c.in.nfreq = reinterpret_cast<const int*>(mb + ED * 16);
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  const int MT = c.Mp / NB, ntl = MT * (MT + 1) / 2;
  int S = (int)std::min<int64_t>(E, max_slots > 0 ? max_slots : 148);
  for (;;) {
    c.nsplit = (c.npad + SGPR_KCHUNK - 1) / SGPR_KCHUNK;
    c.L = vff_layout(D, c.npad, c.mpad, c.Mp, c.TRmax, c.nsplit, with_pred);
    if (c.L.total * 8 * S <= (size_t)((free_b + g.bsg.cap) * 0.8) || S == 1) break;
    S = std::max(1, S / 2);
  }
  c.S = S;
  if (g.bsg.ensure(c.L.total * 8 * S) || g.bslots.ensure(sizeof(SlotDesc) * S) ||
      g.bmems.ensure((sizeof(VffDesc) + sizeof(SgprDesc)) * S))
    return 1;
  c.ws = g.bsg.as<double>();
  c.slots = g.bslots.as<SlotDesc>();
  c.descs = g.bmems.as<VffDesc>();
  c.sds = reinterpret_cast<SgprDesc*>(c.descs + S);
  return 0;
}

static int vff_launch_eval(Engine& g, const VffCall& c, bool with_grad, bool with_pred, double* dfm, double* dfv,
                           double* dyv) {
  const int S = c.S, MT = c.Mp / NB, ntl = MT * (MT + 1) / 2, NT = c.npad / NB, PT = c.mpad / NB, D = c.D;
  cudaStream_t st = g.stream;
  // static quantities of the experts that just entered a slot (no-ops for the others)
  vff_feat_kernel<<<dim3((c.npad + 255) / 256, D, S), 256, 0, st>>>(c.descs, 0);
  vff_kuf_build_kernel<<<dim3(NT, MT, S), GPSAT_THREADS, 0, st>>>(c.descs, 0);
  sgpr_gemm_kernel<<<dim3(ntl * c.nsplit, S), GPSAT_THREADS, SGPR_GEMM_SMEM_BYTES, st>>>(c.sds, SGPR_GEMM_SYRK, c.nsplit, 0);
  sgpr_reduce_kernel<<<dim3(std::min(256, (c.Mp * c.Mp + 255) / 256), S), 256, 0, st>>>(c.sds, c.nsplit, 0);
  // this evaluation
  vff_dim_kernel<<<dim3(D, S), 128, VFF_DIM_SMEM_BYTES, st>>>(c.descs);
  g.stats.kernel_launches += 5;
  if (with_pred) {
    vff_feat_kernel<<<dim3((c.mpad + 255) / 256, D, S), 256, 0, st>>>(c.descs, 1);
    vff_kuf_build_kernel<<<dim3(PT, MT, S), GPSAT_THREADS, 0, st>>>(c.descs, 1);
    g.stats.kernel_launches += 2;
  }
  vff_build_p_kernel<<<dim3(MT, MT, S), GPSAT_THREADS, 0, st>>>(c.descs);
  if (launch_sweep(g, S, MT, with_pred ? c.TRmax : MT, with_grad, c.slots, false)) return 1;
  vff_grad_kernel<<<dim3(ntl, S), GPSAT_THREADS, VFF_GRAD_SMEM_BYTES, st>>>(c.descs, c.slots, with_grad ? 1 : 0);
  g.stats.kernel_launches += 2;
  if (with_pred) {
    vff_predict_reduce_kernel<<<dim3((unsigned)((c.mpad * 32 + 255) / 256), S), 256, 0, st>>>(c.descs, c.slots,
                                                                                            c.bd.off_m, dfm, dfv, dyv);
    g.stats.kernel_launches++;
  }
  g.stats.ticks++;
  return 0;
}

static int64_t vff_M(const gpsat_vff_domain* dom, int64_t e, int D) {
  int64_t M = 1;
  for (int dd = 0; dd < D; ++dd) M *= 2 * dom->nfreq[e * D + dd] - 1;
  return M;
}

static int vff_eval_impl(gpsat_engine* h, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem,
                         const double* theta, bool with_pred, double* fmean, double* fvar, double* yvar,
                         double* elbo, double* grad, int32_t* status) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, with_pred)) return 2;
  REQUIRE(theta != nullptr, "theta is NULL");
  REQUIRE(!with_pred || (fmean && fvar && yvar), "prediction output is NULL");
  REQUIRE(with_pred || elbo, "elbo is NULL");
  Engine& g = h->e;
  begin_call(g);
  const int64_t E = batch->E;
  if (E == 0) return end_call(g);
  const bool dev = mem == GPSAT_MEM_DEVICE, with_grad = grad != nullptr;
  VffCall c;
  if (int rc = vff_setup(g, batch, dom, mem, with_pred, 0, c)) return rc;
  const int P = c.bd.P;
  const int64_t Mtot = with_pred ? batch->off_m[E] : 0;
  const size_t EP = (size_t)E * P;
  size_t o_theta = 0, o_elbo = EP * 8, o_grad = o_elbo + (size_t)E * 8, o_status = o_grad + EP * 8;
  size_t o_pred = (o_status + (size_t)E * 4 + 15) & ~(size_t)15;
  if (g.bout.ensure(o_pred + (size_t)std::max<int64_t>(Mtot, 1) * 3 * 8)) return 1;
  char* base = g.bout.as<char>();
  const void* dtheta = nullptr;
  if (stage_in(g, mem, theta, EP * 8, base + o_theta, &dtheta)) return 1;
  double* delbo = elbo ? (dev ? elbo : reinterpret_cast<double*>(base + o_elbo)) : nullptr;
  double* dgrad = grad ? (dev ? grad : reinterpret_cast<double*>(base + o_grad)) : nullptr;
  int* dstatus = status ? (dev ? status : reinterpret_cast<int*>(base + o_status)) : nullptr;
  double* dfm = dev ? fmean : reinterpret_cast<double*>(base + o_pred);
  double* dfv = dev ? fvar : dfm + Mtot;
  double* dyv = dev ? yvar : dfv + Mtot;
  const int S = c.S;
  for (int64_t b0 = 0; b0 < E; b0 += S) {
    vff_eval_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(c.descs, c.sds, c.slots, c.ws, c.L, c.bd, c.in,
                                                                (const double*)dtheta, (int)b0, S, with_pred ? 1 : 0);
    if (vff_launch_eval(g, c, with_grad, with_pred, dfm, dfv, dyv)) return 1;
    vff_eval_finish_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(c.descs, c.slots, c.bd, delbo, dgrad, dstatus, S);
    g.stats.kernel_launches += 2;
  }
  for (int64_t e = 0; e < E; ++e) {
    const double M = (double)vff_M(dom, e, batch->D), n = (double)g.hn[e];
    g.stats.sweep_flops += M * M * n + (with_grad ? M * M * M : M * M * M / 3.0) + M * M * (double)g.hm[e];
  }
  g.stats.evals = E; g.stats.slots = S;
  if (!dev) {
    if (with_pred && Mtot > 0) {
      CUDA_TRY(cudaMemcpyAsync(fmean, dfm, (size_t)Mtot * 8, cudaMemcpyDeviceToHost, g.stream));
      CUDA_TRY(cudaMemcpyAsync(fvar, dfv, (size_t)Mtot * 8, cudaMemcpyDeviceToHost, g.stream));
      CUDA_TRY(cudaMemcpyAsync(yvar, dyv, (size_t)Mtot * 8, cudaMemcpyDeviceToHost, g.stream));
    }
    if (elbo) CUDA_TRY(cudaMemcpyAsync(elbo, delbo, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
    if (grad) CUDA_TRY(cudaMemcpyAsync(grad, dgrad, EP * 8, cudaMemcpyDeviceToHost, g.stream));
    if (status) CUDA_TRY(cudaMemcpyAsync(status, dstatus, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)Mtot * 24 + (double)E * 12 + (grad ? (double)EP * 8 : 0.0);
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

extern "C" {

int gpsat_vff_elbo(gpsat_engine* h, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem,
                   const double* theta, double* elbo, double* grad, int32_t* status) {
  return vff_eval_impl(h, batch, dom, mem, theta, false, nullptr, nullptr, nullptr, elbo, grad, status);
}

int gpsat_vff_predict(gpsat_engine* h, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem,
                      const double* theta, double* fmean, double* fvar, double* yvar, double* elbo, int32_t* status) {
  return vff_eval_impl(h, batch, dom, mem, theta, true, fmean, fvar, yvar, elbo, nullptr, status);
}

int gpsat_vff_fit(gpsat_engine* h, const gpsat_batch* batch, const gpsat_vff_domain* dom, int mem, double* theta_io,
                  const int32_t* tkind, const double* ta, const double* tb, const uint8_t* trainable,
                  const gpsat_fit_options* opt_in, const gpsat_priors* priors, double* objective, double* elbo,
                  int32_t* n_evals, int32_t* n_iters, int32_t* status) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, false)) return 2;
  REQUIRE(theta_io && tkind && ta && tb && trainable, "theta / transform argument is NULL");
  REQUIRE(objective && n_evals && n_iters && status, "output argument is NULL");
  REQUIRE(!priors || (priors->kind && priors->a && priors->b), "priors has a NULL member");
  gpsat_fit_options opt;
  if (opt_in) opt = *opt_in; else gpsat_fit_options_default(&opt);
  REQUIRE(opt.maxls > 0 && opt.maxiter >= 0 && opt.maxfun >= 0, "bad fit options");
  Engine& g = h->e;
  begin_call(g);
  const int64_t E = batch->E;
  if (E == 0) return end_call(g);
  VffCall c;
  if (int rc = vff_setup(g, batch, dom, mem, false, opt.max_slots, c)) return rc;
  const int S = c.S, P = c.bd.P;
  FitDev fd;
  if (fit_stage(g, mem, E, P, S, theta_io, tkind, ta, tb, trainable, opt, priors, objective, elbo, n_evals, n_iters,
                status, fd))
    return 1;
  int init_cnt[4] = {(int)std::min<int64_t>(S, E), 0, 0, 0};
  CUDA_TRY(cudaMemcpyAsync(fd.counters, init_cnt, sizeof(init_cnt), cudaMemcpyHostToDevice, g.stream));
  vff_fit_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(c.descs, c.sds, c.slots, c.ws, c.L, c.bd, c.in, fd, S);
  g.stats.kernel_launches++;
  int evals = 0;
  for (;;) {
    CUDA_TRY(cudaMemcpyAsync(g.h_counters, fd.counters, 4 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    evals = g.h_counters[3];
    if (g.h_counters[1] == 0) break;
    CUDA_TRY(cudaMemsetAsync(fd.counters + 1, 0, 2 * sizeof(int), g.stream));
    if (vff_launch_eval(g, c, true, false, nullptr, nullptr, nullptr)) return 1;
    {
      ProfScope ps(g, PROF_OPT);
      vff_lbfgs_kernel<<<(S + 63) / 64, 64, 0, g.stream>>>(c.descs, c.sds, c.slots, c.ws, c.L, c.bd, c.in, fd, S);
    }
    g.stats.kernel_launches++;
  }
  g.stats.evals = evals; g.stats.slots = S;
  if (fit_fetch(g, mem, E, P, fd, theta_io, objective, elbo, n_evals, n_iters, status)) return 1;
  CUDA_TRY(cudaGetLastError());
  if (int rc = end_call(g)) return rc;
  if (mem != GPSAT_MEM_DEVICE)
    for (int64_t e = 0; e < E; ++e) {
      const double M = (double)vff_M(dom, e, batch->D);
      g.stats.sweep_flops += M * M * (double)g.hn[e] + (double)n_evals[e] * M * M * M;
    }
  return 0;
}

}  // extern "C"

#include "engine_post.inl"
