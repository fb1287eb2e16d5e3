// C ABI entry points (include/gpsat_b200.h).  Included at the end of engine_host.inl.
using namespace gpsat;

struct gpsat_engine { Engine e; };

extern "C" {

int gpsat_version(void) { return 100; }
const char* gpsat_last_error(void) { return g_last_error.c_str(); }

void* gpsat_engine_stream(const gpsat_engine* h) { return h ? (void*)h->e.stream : nullptr; }

int gpsat_engine_set_profile(gpsat_engine* h, int on) {
  REQUIRE(h != nullptr, "engine is NULL");
  h->e.profile = on != 0;
  return 0;
}

int gpsat_engine_set_lauum_path(gpsat_engine* h, int path) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(path == GPSAT_LAUUM_DMMA || path == GPSAT_LAUUM_I8 || path == GPSAT_LAUUM_I8Q, "unknown lauum path");
  h->e.i8_lauum = path != GPSAT_LAUUM_DMMA;
  h->e.i8_lauum_q = path == GPSAT_LAUUM_I8Q;
  return 0;
}

int gpsat_engine_set_step_path(gpsat_engine* h, int path) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(path == GPSAT_STEP_DMMA || path == GPSAT_STEP_I8, "unknown step path");
  h->e.i8_step = path == GPSAT_STEP_I8;
  return 0;
}

int gpsat_engine_i8_phases(gpsat_engine* h, uint64_t* out8, int reset) {
  REQUIRE(h != nullptr && out8 != nullptr, "NULL argument");
  CUDA_TRY(cudaStreamSynchronize(h->e.stream));
  CUDA_TRY(cudaMemcpyFromSymbol(out8, g_i8_phase, 8 * sizeof(uint64_t)));
  if (reset) {
    uint64_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    CUDA_TRY(cudaMemcpyToSymbol(g_i8_phase, z, sizeof(z)));
  }
  return 0;
}

int gpsat_engine_step_phases(gpsat_engine* h, uint64_t* out32, int reset) {
  REQUIRE(h != nullptr && out32 != nullptr, "NULL argument");
  CUDA_TRY(cudaStreamSynchronize(h->e.stream));
  CUDA_TRY(cudaMemcpyFromSymbol(out32, g_step_phase, 32 * sizeof(uint64_t)));
  if (reset) {
    uint64_t z[32] = {};
    CUDA_TRY(cudaMemcpyToSymbol(g_step_phase, z, sizeof(z)));
  }
  return 0;
}

int gpsat_engine_i8_timeouts(gpsat_engine* h, int* out) {
  REQUIRE(h != nullptr && out != nullptr, "NULL argument");
  *out = (int)h->e.i8_timeouts_seen;
  return 0;
}

int gpsat_engine_debug_set_i8_spin_limit(gpsat_engine* h, int polls) {
  REQUIRE(h != nullptr, "engine is NULL");
  CUDA_TRY(cudaSetDevice(h->e.device));
  CUDA_TRY(cudaStreamSynchronize(h->e.stream));
  CUDA_TRY(cudaMemcpyToSymbol(g_i8_spin_limit, &polls, sizeof(int)));
  return 0;
}

int gpsat_engine_set_precision_rule(gpsat_engine* h, double rho_max) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(rho_max >= 0.0, "rho_max must be >= 0 (0 = int8-sliced arithmetic at any conditioning)");
  h->e.i8_rho_max = rho_max;
  return 0;
}

double gpsat_engine_precision_rule_default(void) { return I8_RHO_MAX_DEFAULT; }

int gpsat_engine_set_precision_low_tier(gpsat_engine* h, double rho_low) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(rho_low >= 0.0, "rho_low must be >= 0 (0 = every int8-sliced k-sum keeps all its slice pairs)");
  h->e.i8_rho_low = rho_low;
  return 0;
}

double gpsat_engine_precision_low_tier_default(void) { return I8_RHO_LOW_DEFAULT; }

void gpsat_fit_options_default(gpsat_fit_options* o) {
  o->pgtol = 1e-5;
  o->ftol = 2.220446049250313e-09;
  o->maxiter = 10000;
  o->maxfun = 15000;
  o->maxls = 20;
  o->max_slots = 0;
}

int gpsat_device_name(int device, char* buf, size_t len) {
  REQUIRE(buf != nullptr && len > 0, "buf is NULL");
  int count = 0;
  CUDA_TRY(cudaGetDeviceCount(&count));
  REQUIRE(device >= 0 && device < count, "no such CUDA device");
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  snprintf(buf, len, "%s", prop.name);
  return 0;
}

int gpsat_engine_create(int device, gpsat_engine** out) {
  REQUIRE(out != nullptr, "out is NULL");
  int count = 0;
  CUDA_TRY(cudaGetDeviceCount(&count));
  REQUIRE(device >= 0 && device < count, "no such CUDA device");
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  REQUIRE(prop.major == 10, "libgpsat_b200 is built for sm_100a (B200) only");
  gpsat_engine* h = new gpsat_engine();
  h->e.device = device;
  CUDA_TRY(cudaStreamCreateWithFlags(&h->e.stream, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreate(&h->e.ev0));
  CUDA_TRY(cudaEventCreate(&h->e.ev1));
  CUDA_TRY(cudaMallocHost(&h->e.h_counters, 8 * sizeof(int)));
  if (engine_setup_kernels()) { delete h; return 1; }
  if (const char* lp = getenv("GPSAT_B200_LAUUM")) {
    h->e.i8_lauum = strcmp(lp, "dmma") != 0;
    h->e.i8_persistent = strcmp(lp, "i8p") == 0;
    h->e.i8_lauum_q = strcmp(lp, "i8q") == 0;
  }
  if (const char* sp = getenv("GPSAT_B200_STEP")) h->e.i8_step = strcmp(sp, "dmma") != 0;
  if (const char* dp = getenv("GPSAT_B200_DIAGSUM")) h->e.i8_diag_sum = strcmp(dp, "0") != 0;
  if (const char* rp = getenv("GPSAT_B200_RHO_MAX")) h->e.i8_rho_max = std::max(0.0, atof(rp));
  if (const char* rp = getenv("GPSAT_B200_RHO_LOW")) h->e.i8_rho_low = std::max(0.0, atof(rp));
  if (const char* rp = getenv("GPSAT_B200_TIER_GUARD")) h->e.i8_tier_guard = std::max(0.0, atof(rp));
  h->e.num_sms = prop.multiProcessorCount;
  *out = h;
  return 0;
}

int gpsat_engine_destroy(gpsat_engine* h) {
  if (!h) return 0;
  Engine& g = h->e;
  cudaSetDevice(g.device);
  cudaStreamSynchronize(g.stream);
  for (cudaEvent_t ev : g.evpool) cudaEventDestroy(ev);
  for (DevBuf* b : {&g.bX, &g.by, &g.bXs, &g.boffn, &g.boffm, &g.border, &g.btheta, &g.bmisc, &g.bslots,
                    &g.bmems, &g.bA, &g.bside, &g.bstates, &g.bout, &g.bcov, &g.bZ, &g.bsg, &g.bUq, &g.bLq, &g.bUq8, &g.bynorm, &g.bQ})
    b->release();
  if (g.h_counters) cudaFreeHost(g.h_counters);
  cudaEventDestroy(g.ev0); cudaEventDestroy(g.ev1);
  cudaStreamDestroy(g.stream);
  delete h;
  return 0;
}

int gpsat_engine_stats(const gpsat_engine* h, gpsat_stats* out) {
  REQUIRE(h && out, "NULL argument");
  *out = h->e.stats;
  return 0;
}

// ---- helpers for host<->device staging of small per-expert arrays ---------------------------
static int stage_in(Engine& g, int mem, const void* src, size_t bytes, void* dev_dst, const void** out) {
  if (mem == GPSAT_MEM_DEVICE) { *out = src; return 0; }
  CUDA_TRY(cudaMemcpyAsync(dev_dst, src, bytes, cudaMemcpyHostToDevice, g.stream));
  g.stats.h2d_bytes += (double)bytes;
  *out = dev_dst;
  return 0;
}

static void begin_call(Engine& g) {
  cudaSetDevice(g.device);
  g.stats = gpsat_stats{};
  g.evtype.clear();
  g.evcount.clear();
  cudaEventRecord(g.ev0, g.stream);
}
// Every product call ends here.  A bounded barrier wait of the int8 pipeline that gave up (mbar_wait in i8mma.cuh)
// left unfinished accumulators in some tile: the call fails instead of returning them.
static int end_call(Engine& g) {
  CUDA_TRY(cudaMemcpyFromSymbolAsync(g.h_counters + 6, g_i8_timeout, sizeof(int), 0, cudaMemcpyDeviceToHost, g.stream));
  CUDA_TRY(cudaEventRecord(g.ev1, g.stream));
  CUDA_TRY(cudaStreamSynchronize(g.stream));
  float ms = 0.f;
  CUDA_TRY(cudaEventElapsedTime(&ms, g.ev0, g.ev1));
  g.stats.gpu_ms = ms;
  if (g.profile) prof_collect(g);
  if (g.h_counters[6] != 0) {
    const int zero = 0;
    cudaMemcpyToSymbol(g_i8_timeout, &zero, sizeof(int));
    g.i8_timeouts_seen++;
    g_last_error = "gpsat: a barrier wait of the int8 tensor pipeline timed out; the results of this call are invalid "
                   "(per-expert outputs must not be used)";
    return 3;
  }
  return 0;
}

// ---- lml + gradient at fixed theta ----------------------------------------------------------
int gpsat_gpr_lml_grad(gpsat_engine* h, const gpsat_batch* batch, int mem, const double* theta, double* lml,
                       double* grad, int32_t* status) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, false)) return 2;
  REQUIRE(theta != nullptr && lml != nullptr, "theta / lml is NULL");
  Engine& g = h->e;
  begin_call(g);
  const int64_t E = batch->E;
  if (E == 0) return end_call(g);
  const int P = batch->D + 2;
  BatchDev bd;
  if (engine_load_batch(g, batch, mem, false, bd)) return 1;
  const int S = choose_slots(g, E, g.Tmax, g.Tmax, 0);
  int ld = 0;
  if (engine_alloc_slots(g, S, g.Tmax, g.Tmax, ld, grad != nullptr)) return 1;
  bd.skip_fp64 = g.lauum_planes8 ? 1 : 0;       // the gradient's K^-1 products read the sweep's planes: FP64 tiles unread
  // device-side theta / outputs
  size_t o_theta = 0, o_lml = (size_t)E * P * 8, o_grad = o_lml + (size_t)E * 8, o_status = o_grad + (size_t)E * P * 8;
  if (g.bout.ensure(o_status + (size_t)E * 4)) return 1;
  char* base = g.bout.as<char>();
  const void* dtheta = nullptr;
  if (stage_in(g, mem, theta, (size_t)E * P * 8, base + o_theta, &dtheta)) return 1;
  double* dlml = mem == GPSAT_MEM_DEVICE ? lml : reinterpret_cast<double*>(base + o_lml);
  double* dgrad = grad ? (mem == GPSAT_MEM_DEVICE ? grad : reinterpret_cast<double*>(base + o_grad)) : nullptr;
  int* dstatus = status ? (mem == GPSAT_MEM_DEVICE ? status : reinterpret_cast<int*>(base + o_status)) : nullptr;
  SlotDesc* slots = g.bslots.as<SlotDesc>();
  const SlotMem* mems = g.bmems.as<SlotMem>();
  for (int64_t b0 = 0; b0 < E; b0 += S) {
    int Tb = (int)((g.hn[g.order[b0]] + 1 + NB - 1) / NB);
    eval_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(slots, mems, bd, (const double*)dtheta, (int)b0, S, 0, ld);
    if (launch_sweep(g, S, Tb, Tb, grad != nullptr)) return 1;
    eval_finish_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(slots, bd, (const double*)dtheta, dlml, dgrad, dstatus, S);
    g.stats.kernel_launches += 2;
    g.stats.ticks++;
  }
  for (int64_t e = 0; e < E; ++e) g.stats.sweep_flops += (grad ? 1.0 : 1.0 / 3.0) * (double)g.hn[e] * g.hn[e] * g.hn[e];
  g.stats.evals = E; g.stats.slots = S;
  if (mem == GPSAT_MEM_HOST) {
    CUDA_TRY(cudaMemcpyAsync(lml, dlml, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
    if (grad) CUDA_TRY(cudaMemcpyAsync(grad, dgrad, (size_t)E * P * 8, cudaMemcpyDeviceToHost, g.stream));
    if (status) CUDA_TRY(cudaMemcpyAsync(status, dstatus, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)E * (8 + (grad ? P * 8 : 0) + (status ? 4 : 0));
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

// ---- prediction at fixed theta --------------------------------------------------------------
}  // extern "C"

static int predict_impl(gpsat_engine* h, const gpsat_batch* batch, int mem, const double* theta, double* fmean,
                        double* fvar, double* yvar, double* fcov, double* lml, int32_t* status) {
  REQUIRE(h != nullptr, "engine is NULL");
  if (check_batch(batch, true)) return 2;
  REQUIRE(theta && fmean && fvar && yvar, "theta / output is NULL");
  Engine& g = h->e;
  begin_call(g);
  const int64_t E = batch->E;
  if (E == 0) return end_call(g);
  const int P = batch->D + 2;
  const int64_t M = batch->off_m[E];
  BatchDev bd;
  if (engine_load_batch(g, batch, mem, true, bd)) return 1;
  const int S = choose_slots(g, E, g.Tmax, g.TRmax, 0);
  int ld = 0;
  if (engine_alloc_slots(g, S, g.Tmax, g.TRmax, ld, false)) return 1;
  bd.skip_fp64 = 1;                              // prediction pass: only the rows from the observation row on are read back
  size_t o_theta = 0, o_lml = (size_t)E * P * 8, o_status = o_lml + (size_t)E * 8;
  size_t o_pred = (o_status + (size_t)E * 4 + 15) & ~(size_t)15;
  if (g.bout.ensure(o_pred + (size_t)std::max<int64_t>(M, 1) * 3 * 8)) return 1;
  char* base = g.bout.as<char>();
  const void* dtheta = nullptr;
  if (stage_in(g, mem, theta, (size_t)E * P * 8, base + o_theta, &dtheta)) return 1;
  const bool dev = mem == GPSAT_MEM_DEVICE;
  double* dlml = lml ? (dev ? lml : reinterpret_cast<double*>(base + o_lml)) : nullptr;
  int* dstatus = status ? (dev ? status : reinterpret_cast<int*>(base + o_status)) : nullptr;
  double* dfm = dev ? fmean : reinterpret_cast<double*>(base + o_pred);
  double* dfv = dev ? fvar : dfm + M;
  double* dyv = dev ? yvar : dfv + M;
  // full covariance blocks: expert e owns m_e x m_e doubles at off_c[e]
  double* dcov = nullptr;
  int64_t* doffc = nullptr;
  int64_t ctotal = 0;
  if (fcov) {
    std::vector<int64_t> offc(E + 1, 0);
    for (int64_t e = 0; e < E; ++e) offc[e + 1] = offc[e] + g.hm[e] * g.hm[e];
    ctotal = offc[E];
    if (g.bmisc.ensure((E + 1) * 8)) return 1;
    doffc = g.bmisc.as<int64_t>();
    CUDA_TRY(cudaMemcpyAsync(doffc, offc.data(), (E + 1) * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));    // offc is a stack-lifetime source
    if (dev) dcov = fcov;
    else {
      if (g.bcov.ensure(std::max<int64_t>(ctotal, 1) * 8)) return 1;
      dcov = g.bcov.as<double>();
    }
  }
  SlotDesc* slots = g.bslots.as<SlotDesc>();
  const SlotMem* mems = g.bmems.as<SlotMem>();
  for (int64_t b0 = 0; b0 < E; b0 += S) {
    int Tb = 0, TRb = 0; int64_t mb = 0;
    for (int64_t k = b0; k < std::min<int64_t>(E, b0 + S); ++k) {
      int e = g.order[k];
      Tb = std::max(Tb, (int)((g.hn[e] + 1 + NB - 1) / NB));
      TRb = std::max(TRb, (int)((g.hn[e] + 1 + g.hm[e] + NB - 1) / NB));
      mb = std::max(mb, g.hm[e]);
    }
    eval_init_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(slots, mems, bd, (const double*)dtheta, (int)b0, S, 1, ld);
    if (launch_sweep(g, S, Tb, TRb, false)) return 1;
    if (mb > 0) {
      ProfScope ps(g, PROF_REDUCE);
      predict_reduce_kernel<<<dim3((unsigned)((mb * 32 + 255) / 256), S), 256, 0, g.stream>>>(slots, bd.off_m, dfm, dfv, dyv);
      g.stats.kernel_launches++;
      if (fcov) {
        const unsigned nt = (unsigned)((mb + 63) / 64);
        predict_cov_kernel<<<dim3(nt, nt, S), 256, 0, g.stream>>>(slots, doffc, dcov);
        g.stats.kernel_launches++;
      }
    }
    eval_finish_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(slots, bd, (const double*)dtheta, dlml, nullptr, dstatus, S);
    g.stats.kernel_launches += 2;
    g.stats.ticks++;
  }
  for (int64_t e = 0; e < E; ++e) {
    double n = (double)g.hn[e];
    g.stats.sweep_flops += n * n * n / 3.0 + n * n * (double)g.hm[e];
  }
  g.stats.evals = E; g.stats.slots = S;
  if (!dev) {
    if (M > 0) {
      CUDA_TRY(cudaMemcpyAsync(fmean, dfm, (size_t)M * 8, cudaMemcpyDeviceToHost, g.stream));
      CUDA_TRY(cudaMemcpyAsync(fvar, dfv, (size_t)M * 8, cudaMemcpyDeviceToHost, g.stream));
      CUDA_TRY(cudaMemcpyAsync(yvar, dyv, (size_t)M * 8, cudaMemcpyDeviceToHost, g.stream));
    }
    if (lml) CUDA_TRY(cudaMemcpyAsync(lml, dlml, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
    if (status) CUDA_TRY(cudaMemcpyAsync(status, dstatus, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    if (fcov && ctotal > 0) CUDA_TRY(cudaMemcpyAsync(fcov, dcov, (size_t)ctotal * 8, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)M * 24 + (double)E * 12 + (fcov ? (double)ctotal * 8 : 0.0);
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

extern "C" {

int gpsat_gpr_predict(gpsat_engine* h, const gpsat_batch* batch, int mem, const double* theta, double* fmean,
                      double* fvar, double* yvar, double* lml, int32_t* status) {
  return predict_impl(h, batch, mem, theta, fmean, fvar, yvar, nullptr, lml, status);
}

int gpsat_gpr_predict_cov(gpsat_engine* h, const gpsat_batch* batch, int mem, const double* theta, double* fmean,
                          double* fvar, double* yvar, double* fcov, double* lml, int32_t* status) {
  REQUIRE(fcov != nullptr, "fcov is NULL");
  return predict_impl(h, batch, mem, theta, fmean, fvar, yvar, fcov, lml, status);
}

}  // extern "C"

#include "engine_fit.inl"
