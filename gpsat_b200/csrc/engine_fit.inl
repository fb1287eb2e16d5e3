// gpsat_gpr_fit and the factor test hook.  Included at the end of engine_api.inl.
// Stage the per-expert optimiser inputs on the device (or adopt device pointers) and point the
// outputs at device memory.  Uses g.bout and g.bstates.
static int fit_stage(Engine& g, int mem, int64_t E, int P, int S, double* theta_io, const int32_t* tkind,
                     const double* ta, const double* tb, const uint8_t* trainable, const gpsat_fit_options& opt,
                     const gpsat_priors* priors, double* objective, double* lml, int32_t* n_evals, int32_t* n_iters,
                     int32_t* status, FitDev& fd) {
  // per-expert arrays on the device: [theta | ta | tb | pa | pb | lml | obj] doubles,
  // [tkind | pkind | nev | nit | status] ints, trainable bytes, counters
  const size_t EP = (size_t)E * P;
  size_t o_theta = 0, o_ta = o_theta + EP * 8, o_tb = o_ta + EP * 8, o_pa = o_tb + EP * 8, o_pb = o_pa + EP * 8;
  size_t o_lml = o_pb + EP * 8, o_obj = o_lml + (size_t)E * 8;
  size_t o_tk = o_obj + (size_t)E * 8, o_pk = o_tk + EP * 4, o_nev = o_pk + EP * 4, o_nit = o_nev + (size_t)E * 4;
  size_t o_st = o_nit + (size_t)E * 4, o_tr = o_st + (size_t)E * 4, o_cnt = (o_tr + EP + 15) & ~(size_t)15;
  if (g.bout.ensure(o_cnt + 32) || g.bstates.ensure(sizeof(LbfgsState) * S)) return 1;
  char* base = g.bout.as<char>();
  const bool dev = mem == GPSAT_MEM_DEVICE;
  const void* p = nullptr;
  if (dev) fd.theta = theta_io;
  else {
    CUDA_TRY(cudaMemcpyAsync(base + o_theta, theta_io, EP * 8, cudaMemcpyHostToDevice, g.stream));
    g.stats.h2d_bytes += (double)EP * 8;
    fd.theta = reinterpret_cast<double*>(base + o_theta);
  }
  if (stage_in(g, mem, ta, EP * 8, base + o_ta, &p)) return 1; fd.ta = (const double*)p;
  if (stage_in(g, mem, tb, EP * 8, base + o_tb, &p)) return 1; fd.tb = (const double*)p;
  if (stage_in(g, mem, tkind, EP * 4, base + o_tk, &p)) return 1; fd.tkind = (const int*)p;
  if (stage_in(g, mem, trainable, EP, base + o_tr, &p)) return 1; fd.trainable = (const unsigned char*)p;
  fd.pkind = nullptr; fd.pa = nullptr; fd.pb = nullptr;
  if (priors) {
    if (stage_in(g, mem, priors->kind, EP * 4, base + o_pk, &p)) return 1; fd.pkind = (const int*)p;
    if (stage_in(g, mem, priors->a, EP * 8, base + o_pa, &p)) return 1; fd.pa = (const double*)p;
    if (stage_in(g, mem, priors->b, EP * 8, base + o_pb, &p)) return 1; fd.pb = (const double*)p;
  }
  fd.objective = dev ? objective : reinterpret_cast<double*>(base + o_obj);
  fd.lml = lml ? (dev ? lml : reinterpret_cast<double*>(base + o_lml)) : nullptr;
  fd.n_evals = dev ? n_evals : reinterpret_cast<int*>(base + o_nev);
  fd.n_iters = dev ? n_iters : reinterpret_cast<int*>(base + o_nit);
  fd.status = dev ? status : reinterpret_cast<int*>(base + o_st);
  fd.states = g.bstates.as<LbfgsState>();
  fd.counters = reinterpret_cast<int*>(base + o_cnt);
  fd.opt = LbfgsOptions{opt.pgtol, opt.ftol, opt.maxiter, opt.maxfun, opt.maxls};

  return 0;
}

static int fit_fetch(Engine& g, int mem, int64_t E, int P, const FitDev& fd, double* theta_io, double* objective,
                     double* lml, int32_t* n_evals, int32_t* n_iters, int32_t* status) {
  const bool dev = mem == GPSAT_MEM_DEVICE;
  const size_t EP = (size_t)E * P;
  if (!dev) {
    CUDA_TRY(cudaMemcpyAsync(theta_io, fd.theta, EP * 8, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyAsync(objective, fd.objective, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
    if (lml) CUDA_TRY(cudaMemcpyAsync(lml, fd.lml, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyAsync(n_evals, fd.n_evals, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyAsync(n_iters, fd.n_iters, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyAsync(status, fd.status, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)EP * 8 + (double)E * (20 + (lml ? 8 : 0));
  }
  return 0;
}

extern "C" {

int gpsat_gpr_fit(gpsat_engine* h, const gpsat_batch* batch, int mem, double* theta_io, const int32_t* tkind,
                  const double* ta, const double* tb, const uint8_t* trainable, const gpsat_fit_options* opt_in,
                  const gpsat_priors* priors, double* objective, double* lml, int32_t* n_evals, int32_t* n_iters,
                  int32_t* status) {
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
  const int P = batch->D + 2;
  BatchDev bd;
  if (engine_load_batch(g, batch, mem, false, bd)) return 1;
  const int S = choose_slots(g, E, g.Tmax, g.Tmax, opt.max_slots);
  int ld = 0;
  if (engine_alloc_slots(g, S, g.Tmax, g.Tmax, ld)) return 1;
  bd.skip_fp64 = g.lauum_planes8 ? 1 : 0;

  FitDev fd;
  if (fit_stage(g, mem, E, P, S, theta_io, tkind, ta, tb, trainable, opt, priors, objective, lml, n_evals, n_iters,
                status, fd))
    return 1;
  const bool dev = mem == GPSAT_MEM_DEVICE;
  const size_t EP = (size_t)E * P;

  SlotDesc* slots = g.bslots.as<SlotDesc>();
  const SlotMem* mems = g.bmems.as<SlotMem>();
  // counters: [0] next queue position, [1] occupied, [2] max T, [3] evaluations
  int init_cnt[5] = {(int)std::min<int64_t>(S, E), 0, 0, 0, 0};
  CUDA_TRY(cudaMemcpyAsync(fd.counters, init_cnt, sizeof(init_cnt), cudaMemcpyHostToDevice, g.stream));
  fit_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(slots, mems, bd, fd, S, ld);
  g.stats.kernel_launches++;
  int evals = 0;
  for (;;) {
    CUDA_TRY(cudaMemcpyAsync(g.h_counters, fd.counters, 5 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyFromSymbolAsync(g.h_counters + 6, g_i8_timeout, sizeof(int), 0, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    const int occupied = g.h_counters[1], Tcur = g.h_counters[2], dmma_slots = g.h_counters[4];
    evals = g.h_counters[3];
    if (occupied == 0 || g.h_counters[6] != 0) break;      // done, or an int8 pipeline time-out (end_call reports it)
    CUDA_TRY(cudaMemsetAsync(fd.counters + 1, 0, 2 * sizeof(int), g.stream));
    CUDA_TRY(cudaMemsetAsync(fd.counters + 4, 0, sizeof(int), g.stream));
    if (launch_sweep(g, S, Tcur, Tcur, true, nullptr, true, dmma_slots)) return 1;
    {
      ProfScope ps(g, PROF_OPT);
      lbfgs_kernel<<<(S + 63) / 64, 64, 0, g.stream>>>(slots, mems, bd, fd, S, ld);
    }
    g.stats.kernel_launches++;
    g.stats.ticks++;
  }
  g.stats.evals = evals; g.stats.slots = S;
  if (fit_fetch(g, mem, E, P, fd, theta_io, objective, lml, n_evals, n_iters, status)) return 1;
  CUDA_TRY(cudaGetLastError());
  if (int rc = end_call(g)) return rc;
  // flops: n^3 per evaluation (needs per-expert counts, which are on the host only in HOST mode)
  if (!dev) for (int64_t e = 0; e < E; ++e) g.stats.sweep_flops += (double)n_evals[e] * (double)g.hn[e] * g.hn[e] * g.hn[e];
  return 0;
}

int gpsat_gpr_factor_debug(gpsat_engine* h, const gpsat_batch* batch, const double* theta, double* Lout,
                           double* Uout) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, false)) return 2;
  REQUIRE(batch->E == 1, "factor_debug takes exactly one expert");
  REQUIRE(theta && Lout, "theta / Lout is NULL");
  Engine& g = h->e;
  begin_call(g);
  const int P = batch->D + 2;
  BatchDev bd;
  if (engine_load_batch(g, batch, GPSAT_MEM_HOST, false, bd)) return 1;
  int ld = 0;
  if (engine_alloc_slots(g, 1, g.Tmax, g.Tmax, ld)) return 1;
  const size_t N = (size_t)g.hn[0] + 1;
  if (g.bout.ensure(P * 8 + 2 * N * N * 8)) return 1;
  double* dth = g.bout.as<double>();
  double* dL = dth + P;
  double* dU = dL + N * N;
  CUDA_TRY(cudaMemcpyAsync(dth, theta, P * 8, cudaMemcpyHostToDevice, g.stream));
  eval_init_kernel<<<1, 128, 0, g.stream>>>(g.bslots.as<SlotDesc>(), g.bmems.as<SlotMem>(), bd, dth, 0, 1, 0, ld);
  if (launch_sweep(g, 1, g.Tmax, g.Tmax, true)) return 1;
  debug_extract_kernel<<<256, 256, 0, g.stream>>>(g.bslots.as<SlotDesc>(), dL, Uout ? dU : nullptr);
  CUDA_TRY(cudaMemcpyAsync(Lout, dL, N * N * 8, cudaMemcpyDeviceToHost, g.stream));
  if (Uout) CUDA_TRY(cudaMemcpyAsync(Uout, dU, N * N * 8, cudaMemcpyDeviceToHost, g.stream));
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

}  // extern "C"

#include "engine_sgpr.inl"
