// Host orchestration of the sparse-GPR (SGPR) path: ELBO (+ gradient) at fixed theta, prediction,
// and the batched optimisation.  Replaces GPflowSGPRModel.get_objective_function_value / predict /
// optimise_parameters (GPSat/models/gpflow_models.py:860-862, :229-230 and :290-329 inherited) for
// a batch of experts.  Included at the end of engine_fit.inl (global namespace, using namespace gpsat).

static SgprLay sgpr_layout(int npad, int mpad, int Mp, int nsplit, bool with_grad, bool with_pred, bool with_z = false) {
  SgprLay L{};
  const size_t mm = (size_t)Mp * Mp, nm = (size_t)npad * Mp, pm = (size_t)mpad * Mp;
  const int MT = Mp / NB;
  L.Tmax = MT;
  L.lgd = align_up(MT, 2);
  L.tile_d = (size_t)MT * NB * NB;
  const size_t side = 2 * L.tile_d + L.lgd + 4;     // Wd, Ld, logdet, quad, notpd
  size_t o = 0;
  auto take = [&](size_t& field, size_t n) { field = o; o += (n + 1) & ~(size_t)1; };
  take(L.ct, nm); take(L.aq, nm); take(L.at, with_grad ? nm : 0);
  take(L.w, mm); take(L.u, with_grad ? mm : 0);
  take(L.wb, (with_grad || with_pred) ? mm : 0); take(L.ub, with_grad ? mm : 0);
  take(L.bm, mm); take(L.binv, with_grad ? mm : 0); take(L.q, with_grad ? mm : 0); take(L.tt, with_grad ? mm : 0);
  take(L.part, (size_t)nsplit * mm);
  take(L.gam, with_grad ? Mp : 0); take(L.bid, with_grad ? Mp : 0); take(L.res, with_grad ? npad : 0);
  take(L.rrp, with_grad ? npad / SGPR_RES_ROWS : 0);
  take(L.gpc, with_grad ? (size_t)(npad / NB) * MT * PG : 0); take(L.gpk, with_grad ? (size_t)MT * MT * PG : 0);
  take(L.sums, SGPR_NSUMS);
  L.with_z = with_z ? 1 : 0;
  take(L.zpc, with_z ? (size_t)(npad / NB) * MT * 2 * NB * GPSAT_DMAX : 0);
  take(L.zpk, with_z ? (size_t)MT * MT * 2 * NB * GPSAT_DMAX : 0);
  take(L.cts, with_pred ? pm : 0); take(L.t1, with_pred ? pm : 0); take(L.t2, with_pred ? pm : 0);
  take(L.ak, mm); take(L.ab, mm);
  take(L.side_k, side); take(L.side_b, side);
  L.total = o;
  return L;
}

// Everything one SGPR call needs on the device.
struct SgprCall {
  BatchDev bd;
  SgprIn in;
  SgprLay L;
  int S, nsplit, npad, mpad, Mp;
  double* ws;
  SgprDesc* descs;
  SlotDesc *sk, *sb;
};

static int sgpr_setup(Engine& g, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                      bool with_grad, bool with_pred, int max_slots, SgprCall& c, bool with_z = false) {
  const int64_t E = batch->E;
  const int D = batch->D;
  int Mmax = 0;
  for (int64_t e = 0; e < E; ++e) {
    const int64_t Me = off_z[e + 1] - off_z[e];
    REQUIRE(Me >= 1 && Me < (1 << 16), "every expert needs between 1 and 65535 inducing points");
    Mmax = std::max<int>(Mmax, (int)Me);
  }
  const bool dev = mem == GPSAT_MEM_DEVICE;
  if (engine_load_batch(g, batch, mem, with_pred, c.bd)) return 1;
  const int64_t NZ = off_z[E];
  int64_t nmax = 0, mmax = 0;
  for (int64_t e = 0; e < E; ++e) { nmax = std::max(nmax, g.hn[e]); mmax = std::max(mmax, g.hm[e]); }
  c.npad = align_up((int)nmax, NB);
  c.mpad = align_up((int)std::max<int64_t>(mmax, 1), NB);
  c.Mp = align_up(Mmax + 1, NB);
  // [off_z | zeros] in g.bmisc, Z in g.bZ
  const size_t o_zero = (size_t)(E + 1) * 8;
  if (g.bmisc.ensure(o_zero + (size_t)c.Mp * 8)) return 1;
  CUDA_TRY(cudaMemcpyAsync(g.bmisc.p, off_z, (size_t)(E + 1) * 8, cudaMemcpyHostToDevice, g.stream));
  CUDA_TRY(cudaMemsetAsync(g.bmisc.as<char>() + o_zero, 0, (size_t)c.Mp * 8, g.stream));
  c.in.off_z = g.bmisc.as<int64_t>();
  c.in.zero = reinterpret_cast<const double*>(g.bmisc.as<char>() + o_zero);
  c.in.Z = Z;
  if (!dev) {
    if (g.bZ.ensure((size_t)NZ * D * 8)) return 1;
    CUDA_TRY(cudaMemcpyAsync(g.bZ.p, Z, (size_t)NZ * D * 8, cudaMemcpyHostToDevice, g.stream));
    c.in.Z = g.bZ.as<double>();
  }
  g.stats.h2d_bytes += (double)(E + 1) * 8 + (dev ? 0.0 : (double)NZ * D * 8);
  // wave size and split-K factor of the SYRK: enough CTAs to fill the machine a few times over
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  const int MT = c.Mp / NB, ntile_l = MT * (MT + 1) / 2;
  int S = (int)std::min<int64_t>(E, max_slots > 0 ? max_slots : 148);
  for (;;) {
    c.nsplit = (c.npad + SGPR_KCHUNK - 1) / SGPR_KCHUNK;
    c.L = sgpr_layout(c.npad, c.mpad, c.Mp, c.nsplit, with_grad, with_pred, with_z);
    if (c.L.total * 8 * S <= (size_t)((free_b + g.bsg.cap) * 0.8) || S == 1) break;
    S = std::max(1, S / 2);
  }
  c.S = S;
  if (g.bsg.ensure(c.L.total * 8 * S) || g.bslots.ensure(sizeof(SlotDesc) * 2 * S) ||
      g.bmems.ensure(sizeof(SgprDesc) * S))
    return 1;
  c.ws = g.bsg.as<double>();
  c.sk = g.bslots.as<SlotDesc>();
  c.sb = c.sk + S;
  c.descs = g.bmems.as<SgprDesc>();
  return 0;
}

// One evaluation of every occupied slot: ELBO by-products, optionally the gradient partial sums
// and / or the prediction reductions.
static int sgpr_launch_eval(Engine& g, const SgprCall& c, bool with_grad, bool with_pred, double* dfm, double* dfv,
                            double* dyv, const int64_t* doffc = nullptr, double* dcov = nullptr, int64_t mmax = 0,
                            double* dgradz = nullptr) {
  const int S = c.S, MT = c.Mp / NB, ntl = MT * (MT + 1) / 2, NT = c.npad / NB, PT = c.mpad / NB;
  const unsigned T32 = c.Mp / 32;
  cudaStream_t st = g.stream;
  const SgprDesc* dd = c.descs;
  auto gemm = [&](int mode, int tiles, int nsplit) {
    sgpr_gemm_kernel<<<dim3(tiles * nsplit, S), GPSAT_THREADS, SGPR_GEMM_SMEM_BYTES, st>>>(dd, mode, nsplit,
                                                                                          with_grad ? 1 : 0);
    g.stats.kernel_launches++;
  };
  kuf_build_kernel<<<dim3(NT, MT, S), GPSAT_THREADS, 0, st>>>(dd, 0);
  if (launch_sweep(g, S, MT, MT, true, c.sk, false)) return 1;
  transpose_w_kernel<<<dim3(T32, T32, S), 256, 0, st>>>(dd, c.sk, 0, with_grad ? 1 : 0);
  gemm(SGPR_GEMM_A, MT * NT, 1);
  gemm(SGPR_GEMM_SYRK, ntl, c.nsplit);
  sgpr_reduce_kernel<<<dim3(std::min(256, (c.Mp * c.Mp + 255) / 256), S), 256, 0, st>>>(dd, c.nsplit, 1);
  if (launch_sweep(g, S, MT, MT, with_grad || with_pred, c.sb, false)) return 1;
  g.stats.kernel_launches += 3;
  if (with_grad || with_pred) {
    transpose_w_kernel<<<dim3(T32, T32, S), 256, 0, st>>>(dd, c.sb, 1, with_grad ? 1 : 0);
    g.stats.kernel_launches++;
  }
  if (with_grad) {
    sgpr_gamma_kernel<<<dim3((c.Mp * 32 + 255) / 256, S), 256, 0, st>>>(dd, c.sb);
    sgpr_resid_kernel<<<dim3(c.npad / SGPR_RES_ROWS, S), 256, 0, st>>>(dd);
    gemm(SGPR_GEMM_V, NT * MT, 1);
    gemm(SGPR_GEMM_H, NT * MT, 1);
    gemm(SGPR_GEMM_GC, NT * MT, 1);
    gemm(SGPR_GEMM_BINV, ntl, 1);
    sgpr_q_kernel<<<dim3(std::min(256, (c.Mp * c.Mp + 255) / 256), S), 256, 0, st>>>(dd);
    gemm(SGPR_GEMM_TT, MT * MT, 1);
    gemm(SGPR_GEMM_GK, MT * MT, 1);
    g.stats.kernel_launches += 3;
    if (dgradz) {
      sgpr_gradz_kernel<<<dim3((unsigned)((c.Mp * GPSAT_DMAX + 255) / 256), S), 256, 0, st>>>(dd, c.in.off_z, dgradz);
      g.stats.kernel_launches++;
    }
  }
  sgpr_sums_kernel<<<S, 256, 0, st>>>(dd, with_grad ? 1 : 0);
  g.stats.kernel_launches++;
  if (with_pred) {
    kuf_build_kernel<<<dim3(PT, MT, S), GPSAT_THREADS, 0, st>>>(dd, 1);
    gemm(SGPR_GEMM_T1, PT * MT, 1);
    gemm(SGPR_GEMM_T2, PT * MT, 1);
    sgpr_predict_reduce_kernel<<<dim3((unsigned)((c.mpad * 32 + 255) / 256), S), 256, 0, st>>>(dd, c.sb, c.bd.off_m, dfm,
                                                                                             dfv, dyv);
    g.stats.kernel_launches += 2;
    if (dcov && mmax > 0) {
      const unsigned nt = (unsigned)((mmax + 63) / 64);
      sgpr_predict_cov_kernel<<<dim3(nt, nt, S), 256, 0, st>>>(dd, doffc, dcov);
      g.stats.kernel_launches++;
    }
  }
  g.stats.ticks++;
  return 0;
}

static double sgpr_flops(const Engine& g, const int64_t* off_z, int64_t e, bool with_grad, bool with_pred) {
  const double n = (double)g.hn[e], M = (double)(off_z[e + 1] - off_z[e]);
  double f = 2.0 * M * M * M / 3.0 + 2.0 * n * M * M;
  if (with_grad) f += 3.0 * n * M * M + 4.0 * M * M * M;
  if (with_pred) f += 2.0 * (double)g.hm[e] * M * M;
  return f;
}

static int sgpr_eval_impl(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                          const double* theta, bool with_pred, double* fmean, double* fvar, double* yvar,
                          double* elbo, double* grad, int32_t* status, double* fcov = nullptr, double* gradz = nullptr) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, with_pred)) return 2;
  REQUIRE(theta != nullptr && off_z != nullptr, "theta / off_z is NULL");
  REQUIRE(!with_pred || (fmean && fvar && yvar), "prediction output is NULL");
  REQUIRE(with_pred || elbo, "elbo is NULL");
  Engine& g = h->e;
  begin_call(g);
  const int64_t E = batch->E;
  if (E == 0) return end_call(g);
  REQUIRE(Z != nullptr, "Z is NULL");
  const int P = batch->D + 2;
  const bool dev = mem == GPSAT_MEM_DEVICE, with_grad = grad != nullptr;
  SgprCall c;
  REQUIRE(!gradz || (with_grad && !with_pred), "gradz needs grad and no prediction");
  if (int rc = sgpr_setup(g, batch, off_z, Z, mem, with_grad, with_pred, 0, c, gradz != nullptr)) return rc;
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
  // full covariance blocks: expert e owns m_e x m_e doubles at off_c[e] (as gpsat_gpr_predict_cov)
  double* dcov = nullptr;
  int64_t* doffc = nullptr;
  int64_t ctotal = 0, mmax = 0;
  if (fcov && with_pred) {
    std::vector<int64_t> offc(E + 1, 0);
    for (int64_t e = 0; e < E; ++e) { offc[e + 1] = offc[e] + g.hm[e] * g.hm[e]; mmax = std::max(mmax, g.hm[e]); }
    ctotal = offc[E];
    // [off_c | blocks (host callers)] in g.bcov: g.bmisc already holds off_z and the zero observations (sgpr_setup)
    const size_t o_blocks = ((size_t)(E + 1) * 8 + 255) & ~(size_t)255;
    if (g.bcov.ensure(o_blocks + (dev ? 0 : (size_t)std::max<int64_t>(ctotal, 1) * 8))) return 1;
    doffc = g.bcov.as<int64_t>();
    CUDA_TRY(cudaMemcpyAsync(doffc, offc.data(), (E + 1) * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));    // offc is a stack-lifetime source
    dcov = dev ? fcov : reinterpret_cast<double*>(g.bcov.as<char>() + o_blocks);
  }
  // d elbo / d Z: [off_z[E]][D] in the layout of Z; staged in g.bcov for host callers (no covariance blocks in this mode)
  double* dgradz = nullptr;
  const size_t zbytes = (size_t)off_z[E] * batch->D * 8;
  if (gradz) {
    if (dev) dgradz = gradz;
    else {
      if (g.bcov.ensure(std::max<size_t>(zbytes, 8))) return 1;
      dgradz = g.bcov.as<double>();
    }
  }
  for (int64_t b0 = 0; b0 < E; b0 += S) {
    sgpr_eval_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(c.descs, c.sk, c.sb, c.ws, c.L, c.bd, c.in,
                                                                 (const double*)dtheta, (int)b0, S, with_pred ? 1 : 0);
    if (sgpr_launch_eval(g, c, with_grad, with_pred, dfm, dfv, dyv, doffc, dcov, mmax, dgradz)) return 1;
    sgpr_eval_finish_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(c.descs, c.sk, c.sb, c.bd, delbo, dgrad, dstatus, S);
    g.stats.kernel_launches += 2;
  }
  for (int64_t e = 0; e < E; ++e) g.stats.sweep_flops += sgpr_flops(g, off_z, e, with_grad, with_pred);
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
    if (dcov && ctotal > 0) CUDA_TRY(cudaMemcpyAsync(fcov, dcov, (size_t)ctotal * 8, cudaMemcpyDeviceToHost, g.stream));
    if (gradz && zbytes > 0) {
      CUDA_TRY(cudaMemcpyAsync(gradz, dgradz, zbytes, cudaMemcpyDeviceToHost, g.stream));
      g.stats.d2h_bytes += (double)zbytes;
    }
    g.stats.d2h_bytes += (double)Mtot * 24 + (double)E * 12 + (grad ? (double)EP * 8 : 0.0) + (dcov ? (double)ctotal * 8 : 0.0);
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

extern "C" {

int gpsat_sgpr_elbo(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                    const double* theta, double* elbo, double* grad, int32_t* status) {
  return sgpr_eval_impl(h, batch, off_z, Z, mem, theta, false, nullptr, nullptr, nullptr, elbo, grad, status);
}

int gpsat_sgpr_elbo_gradz(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                          const double* theta, double* elbo, double* grad, double* gradz, int32_t* status) {
  REQUIRE(grad != nullptr && gradz != nullptr, "grad / gradz is NULL");
  return sgpr_eval_impl(h, batch, off_z, Z, mem, theta, false, nullptr, nullptr, nullptr, elbo, grad, status, nullptr, gradz);
}

int gpsat_sgpr_predict(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                       const double* theta, double* fmean, double* fvar, double* yvar, double* elbo, int32_t* status) {
  return sgpr_eval_impl(h, batch, off_z, Z, mem, theta, true, fmean, fvar, yvar, elbo, nullptr, status);
}

int gpsat_sgpr_predict_cov(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                           const double* theta, double* fmean, double* fvar, double* yvar, double* fcov, double* elbo,
                           int32_t* status) {
  REQUIRE(fcov != nullptr, "fcov is NULL");
  return sgpr_eval_impl(h, batch, off_z, Z, mem, theta, true, fmean, fvar, yvar, elbo, nullptr, status, fcov);
}

int gpsat_sgpr_fit(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                   double* theta_io, const int32_t* tkind, const double* ta, const double* tb,
                   const uint8_t* trainable, const gpsat_fit_options* opt_in, const gpsat_priors* priors,
                   double* objective, double* elbo, int32_t* n_evals, int32_t* n_iters, int32_t* status) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, false)) return 2;
  REQUIRE(off_z != nullptr, "off_z is NULL");
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
  REQUIRE(Z != nullptr, "Z is NULL");
  const int P = batch->D + 2;
  SgprCall c;
  if (int rc = sgpr_setup(g, batch, off_z, Z, mem, true, false, opt.max_slots, c)) return rc;
  const int S = c.S;
  FitDev fd;
  if (fit_stage(g, mem, E, P, S, theta_io, tkind, ta, tb, trainable, opt, priors, objective, elbo, n_evals, n_iters,
                status, fd))
    return 1;
  // counters: [0] next queue position, [1] occupied, [2] unused, [3] evaluations
  int init_cnt[4] = {(int)std::min<int64_t>(S, E), 0, 0, 0};
  CUDA_TRY(cudaMemcpyAsync(fd.counters, init_cnt, sizeof(init_cnt), cudaMemcpyHostToDevice, g.stream));
  sgpr_fit_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(c.descs, c.sk, c.sb, c.ws, c.L, c.bd, c.in, fd, S);
  g.stats.kernel_launches++;
  int evals = 0;
  for (;;) {
    CUDA_TRY(cudaMemcpyAsync(g.h_counters, fd.counters, 4 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    evals = g.h_counters[3];
    if (g.h_counters[1] == 0) break;
    CUDA_TRY(cudaMemsetAsync(fd.counters + 1, 0, 2 * sizeof(int), g.stream));
    if (sgpr_launch_eval(g, c, true, false, nullptr, nullptr, nullptr)) return 1;
    {
      ProfScope ps(g, PROF_OPT);
      sgpr_lbfgs_kernel<<<(S + 63) / 64, 64, 0, g.stream>>>(c.descs, c.sk, c.sb, c.ws, c.L, c.bd, c.in, fd, S);
    }
    g.stats.kernel_launches++;
  }
  g.stats.evals = evals; g.stats.slots = S;
  if (fit_fetch(g, mem, E, P, fd, theta_io, objective, elbo, n_evals, n_iters, status)) return 1;
  CUDA_TRY(cudaGetLastError());
  if (int rc = end_call(g)) return rc;
  if (mem != GPSAT_MEM_DEVICE)
    for (int64_t e = 0; e < E; ++e) g.stats.sweep_flops += (double)n_evals[e] * sgpr_flops(g, off_z, e, true, false);
  return 0;
}

}  // extern "C"

#include "engine_vff.inl"
