// Host-side engine: memory, launch sequences, C ABI.  Included at the end of engine.cu.
namespace gpsat {

// Default of the precision rule (slot_set_theta in engine.cu): the measured envelope behind it is
// profiles/r02_precision_envelope.md.
constexpr double I8_RHO_MAX_DEFAULT = 4096.0;
constexpr double I8_RHO_LOW_DEFAULT = 32.0;
constexpr double I8_TIER_GUARD_DEFAULT = 1e-5;

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + (bytes >> 3);
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(&p, bytes); want = bytes; }
    if (e != cudaSuccess) {
      g_last_error = std::string("cudaMalloc(") + std::to_string(bytes) + "): " + cudaGetErrorString(e);
      return 1;
    }
    cap = want;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct Engine {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int* h_counters = nullptr;          // pinned [8]: [0..4] fit counters, [6] int8-pipeline time-out flag of the last call
  int64_t i8_timeouts_seen = 0;       // calls that failed with a time-out since the engine was created
  DevBuf bX, by, bXs, boffn, boffm, border, btheta, bmisc, bslots, bmems, bA, bside, bstates, bout, bcov, bZ, bsg;
  gpsat_stats stats{};
  // host copy of the current batch shape
  std::vector<int64_t> hn, hm;
  std::vector<int> order;
  int Tmax = 0, TRmax = 0;
  // optional per-kernel timing (gpsat_engine_set_profile)
  // K^-1 tile products of the gradient on the int8 tensor path (i8mma.cuh); GPSAT_B200_LAUUM=dmma turns it off
  bool i8_lauum = true;
  bool i8_persistent = false;       // GPSAT_B200_LAUUM=i8p: first persistent variant (register epilogue; slower, kept for reference)
  bool i8_lauum_q = false;          // GPSAT_B200_LAUUM=i8q: persistent variant with the Q tile parked in an L2-resident scratch
                                    // (lauum_q_kernel.cuh): bitwise the same results, measured 11 % slower at c2
  DevBuf bQ;                        // its scratch: one FP64 tile per CTA
  int num_sms = 148;
  int cur_kind = GPSAT_MATERN32, cur_D = 3;     // kernel family / input dimension of the loaded batch
  DevBuf bUq;
  // potrf main loop of step_kernel on the int8 tensor path (step_kernel.cuh); GPSAT_B200_STEP=i8|dmma
  bool i8_step = true;
  bool i8_diag_sum = true;          // GPSAT_B200_DIAGSUM=0: per-tile SYRK + RED instead of the one-go int8 Schur complement
  bool lauum_planes8 = false;       // set by engine_alloc_slots: lauum_grad_i8_kernel reads d.Uq8
  // precision rule (slot_set_theta in engine.cu): evaluations with (var + noise) / noise above this run on FP64 mma.sync
  double i8_rho_max = I8_RHO_MAX_DEFAULT;
  double i8_rho_low = I8_RHO_LOW_DEFAULT;
  double i8_tier_guard = I8_TIER_GUARD_DEFAULT;
  DevBuf bLq, bUq8, bynorm;
  bool profile = false;
  std::vector<cudaEvent_t> evpool;
  std::vector<int> evtype;          // kernel type of event pair k (events 2k, 2k+1)
  std::vector<int> evcount;         // launches inside event pair k
};

enum ProfType : int { PROF_DIAG = 0, PROF_STEP = 1, PROF_LAUUM = 2, PROF_OPT = 3, PROF_REDUCE = 4 };

// Brackets one kernel launch with a pair of events on the engine's stream when profiling is on.
// One CUDA-event pair around a GROUP of consecutive launches of one kernel type (all step launches of a sweep share one
// pair: two event records per launch kept the host from staying ahead of the GPU and showed up as 10 - 30 ms of idle
// device per c2 bench step in the profiled, device-resident leg).  `launches` is what n_<type> counts.
struct ProfScope {
  Engine& g;
  bool on;
  size_t k = 0;
  ProfScope(Engine& g_, int type, int launches = 1) : g(g_), on(g_.profile) {
    if (!on) return;
    k = g.evtype.size();
    while (g.evpool.size() < 2 * (k + 1)) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      g.evpool.push_back(e);
    }
    g.evtype.push_back(type);
    g.evcount.push_back(launches);
    cudaEventRecord(g.evpool[2 * k], g.stream);
  }
  ~ProfScope() {
    if (on) cudaEventRecord(g.evpool[2 * k + 1], g.stream);
  }
};

static void prof_collect(Engine& g) {
  double* ms[5] = {&g.stats.ms_diag, &g.stats.ms_step, &g.stats.ms_lauum, &g.stats.ms_opt, &g.stats.ms_reduce};
  int64_t* cnt[5] = {&g.stats.n_diag, &g.stats.n_step, &g.stats.n_lauum, &g.stats.n_opt, &g.stats.n_reduce};
  for (size_t k = 0; k < g.evtype.size(); ++k) {
    float t = 0.f;
    cudaEventElapsedTime(&t, g.evpool[2 * k], g.evpool[2 * k + 1]);
    *ms[g.evtype[k]] += t;
    *cnt[g.evtype[k]] += g.evcount[k];
  }
  g.evtype.clear();
  g.evcount.clear();
}

static int align_up(int v, int a) { return (v + a - 1) / a * a; }

static int engine_setup_kernels() {
  CUDA_TRY(cudaFuncSetAttribute(diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DIAG_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, STEP_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(lauum_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LAUUM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(lauum_grad_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LAUUM_I8_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(lauum_grad_i8q_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LAUUM_I8Q_SMEM_BYTES));
#define I8P_ATTR(K) \
  CUDA_TRY(cudaFuncSetAttribute(lauum_grad_i8p_kernel<K, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, LAUUM_I8P_SMEM_BYTES)); \
  CUDA_TRY(cudaFuncSetAttribute(lauum_grad_i8p_kernel<K, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, LAUUM_I8P_SMEM_BYTES));
  I8P_ATTR(GPSAT_MATERN12) I8P_ATTR(GPSAT_MATERN32) I8P_ATTR(GPSAT_MATERN52) I8P_ATTR(GPSAT_RBF)
#undef I8P_ATTR
  CUDA_TRY(cudaFuncSetAttribute(sgpr_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SGPR_GEMM_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(vff_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, VFF_GRAD_SMEM_BYTES));
  CUDA_TRY(cudaFuncSetAttribute(vff_dim_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, VFF_DIM_SMEM_BYTES));
  return 0;
}

static int check_batch(const gpsat_batch* b, bool need_pred) {
  REQUIRE(b != nullptr, "batch is NULL");
  REQUIRE(b->kernel >= 0 && b->kernel <= 3, "unknown kernel id");
  REQUIRE(b->D >= 1 && b->D <= GPSAT_MAX_DIMS, "D out of range [1, 6]");
  REQUIRE(b->E >= 0, "E < 0");
  REQUIRE(b->off_n != nullptr, "off_n is NULL");
  REQUIRE(b->E == 0 || (b->X != nullptr && b->y != nullptr), "X / y is NULL");
  for (int64_t e = 0; e < b->E; ++e) {
    REQUIRE(b->off_n[e + 1] >= b->off_n[e], "off_n must be non-decreasing");
    REQUIRE(b->off_n[e + 1] - b->off_n[e] >= 1, "every expert needs at least one observation");
    REQUIRE(b->off_n[e + 1] - b->off_n[e] < (1 << 20), "expert too large");
  }
  if (need_pred) {
    REQUIRE(b->off_m != nullptr, "off_m is NULL");
    REQUIRE(b->off_m[b->E] == 0 || b->Xs != nullptr, "Xs is NULL");
    for (int64_t e = 0; e < b->E; ++e) REQUIRE(b->off_m[e + 1] >= b->off_m[e], "off_m must be non-decreasing");
  }
  return 0;
}

// Bring the batch to the device (or adopt device pointers) and build the size-sorted order.
static int engine_load_batch(Engine& g, const gpsat_batch* b, int mem, bool with_pred, BatchDev& out) {
  const int64_t E = b->E;
  const int D = b->D;
  g.hn.resize(E); g.hm.assign(E, 0);
  int64_t nmax = 0;
  for (int64_t e = 0; e < E; ++e) { g.hn[e] = b->off_n[e + 1] - b->off_n[e]; nmax = std::max(nmax, g.hn[e]); }
  if (with_pred) for (int64_t e = 0; e < E; ++e) g.hm[e] = b->off_m[e + 1] - b->off_m[e];
  g.order.resize(E);
  std::iota(g.order.begin(), g.order.end(), 0);
  std::stable_sort(g.order.begin(), g.order.end(), [&](int a, int c) { return g.hn[a] > g.hn[c]; });
  g.Tmax = 0; g.TRmax = 0;
  for (int64_t e = 0; e < E; ++e) {
    g.Tmax = std::max(g.Tmax, (int)((g.hn[e] + 1 + NB - 1) / NB));
    g.TRmax = std::max(g.TRmax, (int)((g.hn[e] + 1 + g.hm[e] + NB - 1) / NB));
  }
  const int64_t N = b->off_n[E], M = with_pred ? b->off_m[E] : 0;
  if (g.boffn.ensure((E + 1) * 8) || g.boffm.ensure((E + 1) * 8) || g.border.ensure(std::max<int64_t>(E, 1) * 4)) return 1;
  CUDA_TRY(cudaMemcpyAsync(g.boffn.p, b->off_n, (E + 1) * 8, cudaMemcpyHostToDevice, g.stream));
  if (with_pred) CUDA_TRY(cudaMemcpyAsync(g.boffm.p, b->off_m, (E + 1) * 8, cudaMemcpyHostToDevice, g.stream));
  CUDA_TRY(cudaMemcpyAsync(g.border.p, g.order.data(), E * 4, cudaMemcpyHostToDevice, g.stream));
  g.stats.h2d_bytes += (double)(E + 1) * 8 * (with_pred ? 2 : 1) + E * 4;
  out.kind = b->kernel; out.D = D; out.P = D + 2; out.E = E;
  g.cur_kind = b->kernel; g.cur_D = D;
  out.off_n = g.boffn.as<int64_t>();
  out.off_m = with_pred ? g.boffm.as<int64_t>() : nullptr;
  out.order = g.border.as<int>();
  if (mem == GPSAT_MEM_DEVICE) {
    out.X = b->X; out.y = b->y; out.Xs = b->Xs;
  } else {
    if (g.bX.ensure(std::max<int64_t>(N * D, 1) * 8) || g.by.ensure(std::max<int64_t>(N, 1) * 8)) return 1;
    CUDA_TRY(cudaMemcpyAsync(g.bX.p, b->X, N * D * 8, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync(g.by.p, b->y, N * 8, cudaMemcpyHostToDevice, g.stream));
    g.stats.h2d_bytes += (double)N * (D + 1) * 8;
    out.X = g.bX.as<double>(); out.y = g.by.as<double>(); out.Xs = nullptr;
    if (with_pred && M > 0) {
      if (g.bXs.ensure(M * D * 8)) return 1;
      CUDA_TRY(cudaMemcpyAsync(g.bXs.p, b->Xs, M * D * 8, cudaMemcpyHostToDevice, g.stream));
      g.stats.h2d_bytes += (double)M * D * 8;
      out.Xs = g.bXs.as<double>();
    }
  }
  out.ynorm2 = nullptr;
  out.i8_rho_max = g.i8_rho_max;
  out.i8_rho_low = g.i8_rho_low;
  out.skip_fp64 = 0;
  out.i8_tier_guard = g.i8_tier_guard;
  if (g.i8_step && E > 0) {
    if (g.bynorm.ensure(E * 8)) return 1;
    ynorm_kernel<<<(unsigned)E, 128, 0, g.stream>>>(out.off_n, out.y, g.bynorm.as<double>());
    g.stats.kernel_launches++;
    out.ynorm2 = g.bynorm.as<double>();
  }
  return 0;
}

// Carve per-slot scratch for S slots with `rows` x `ld` matrices.
static size_t i8_slot_bytes(int Tmax) {
  return (size_t)Tmax * (Tmax + 1) / 2 * I8_S * I8_PLANE_B + (size_t)Tmax * NB * sizeof(int);
}
// planes of the strict lower block triangle (all tile rows, tile columns 0 .. Tmax-2 are ever read)
static size_t i8l_slot_bytes(int TRmax) { return std::max<size_t>((size_t)TRmax * (TRmax - 1) / 2 * I8L_TILE_B, 16); }
static size_t i8u_slot_bytes(int Tmax) { return (size_t)Tmax * (Tmax + 1) / 2 * I8L_TILE_B; }
// (independent of the batch composition except for the size cap, so an expert gives the same bits alone or in a batch)
static bool i8_step_usable(const Engine& g, int Tmax, int TRmax) { return g.i8_step && TRmax <= I8L_MAX_T; }

static int engine_alloc_slots(Engine& g, int S, int Tmax, int TRmax, int& ld_out, bool with_grad = true) {
  const int ld = Tmax * NB;
  const size_t rows = (size_t)TRmax * NB;
  const size_t a_d = rows * ld;
  const size_t tile_d = (size_t)Tmax * NB * NB;
  const size_t ntile = (size_t)Tmax * (Tmax + 1) / 2;
  size_t side_d = 2 * tile_d + align_up(Tmax, 2) + 2 + ntile * PG + 2;   // Wd, Ld, logdet, quad, gpart, notpd
  side_d = (side_d + 1) & ~(size_t)1;
  if (g.bA.ensure(a_d * 8 * S) || g.bside.ensure(side_d * 8 * S) || g.bslots.ensure(sizeof(SlotDesc) * S) ||
      g.bmems.ensure(sizeof(SlotMem) * S))
    return 1;
  const bool i8l = i8_step_usable(g, Tmax, TRmax);
  const size_t lq_b = i8l_slot_bytes(TRmax);
  if (i8l && g.bLq.ensure(lq_b * S)) return 1;
  const bool i8u = i8l && with_grad;
  const size_t uq8_b = i8u_slot_bytes(Tmax);
  if (i8u && g.bUq8.ensure(uq8_b * S)) return 1;
  // the gradient's K^-1 products read the sweep's planes when there are any; otherwise they slice U themselves
  g.lauum_planes8 = i8u && g.i8_lauum && !g.i8_persistent;
  const bool i8 = g.i8_lauum && with_grad && !g.lauum_planes8;
  const size_t uq_b = i8_slot_bytes(Tmax);
  if (i8 && g.bUq.ensure(uq_b * S)) return 1;
  std::vector<SlotMem> mems(S);
  for (int s = 0; s < S; ++s) {
    double* sd = g.bside.as<double>() + side_d * s;
    SlotMem& m = mems[s];
    m.A = g.bA.as<double>() + a_d * s;
    m.Wd = sd;
    m.Ld = sd + tile_d;
    m.logdet = sd + 2 * tile_d;
    m.quad = m.logdet + align_up(Tmax, 2);
    m.gpart = m.quad + 2;
    m.notpd = reinterpret_cast<int*>(m.gpart + ntile * PG);
    m.Uq = i8 ? g.bUq.as<int8_t>() + uq_b * s : nullptr;
    m.uex = i8 ? reinterpret_cast<int*>(m.Uq + (size_t)Tmax * (Tmax + 1) / 2 * I8_S * I8_PLANE_B) : nullptr;
    m.Lq = i8l ? g.bLq.as<int8_t>() + lq_b * s : nullptr;
    m.Uq8 = i8u ? g.bUq8.as<int8_t>() + uq8_b * s : nullptr;
  }
  CUDA_TRY(cudaMemcpyAsync(g.bmems.p, mems.data(), sizeof(SlotMem) * S, cudaMemcpyHostToDevice, g.stream));
  CUDA_TRY(cudaStreamSynchronize(g.stream));   // `mems` is a stack-lifetime source
  ld_out = ld;
  return 0;
}

static int choose_slots(Engine& g, int64_t E, int Tmax, int TRmax, int max_slots) {
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  size_t per = (size_t)TRmax * NB * Tmax * NB * 8 + (size_t)2 * Tmax * NB * NB * 8 + (1 << 16);
  if (g.i8_lauum) per += i8_slot_bytes(Tmax);
  if (i8_step_usable(g, Tmax, TRmax)) per += i8l_slot_bytes(TRmax) + i8u_slot_bytes(Tmax);
  size_t avail = (size_t)((free_b + g.bA.cap + g.bside.cap + g.bUq.cap + g.bLq.cap + g.bUq8.cap) * 0.80);
  int64_t S = max_slots > 0 ? max_slots : 592;      // 4 x 148 SMs: amortises the per-step launches
  S = std::min<int64_t>(S, E);
  S = std::min<int64_t>(S, (int64_t)(avail / per));
  return (int)std::max<int64_t>(S, 1);
}

template <int KIND>
static void launch_i8p_k(Engine& g, const SlotDesc* slots, int S, int ntile) {
  const int grid = std::min(g.num_sms, ntile * S);
  if (g.cur_D == 3)
    lauum_grad_i8p_kernel<KIND, 3><<<grid, I8P_THREADS, LAUUM_I8P_SMEM_BYTES, g.stream>>>(slots, S, ntile);
  else
    lauum_grad_i8p_kernel<KIND, 0><<<grid, I8P_THREADS, LAUUM_I8P_SMEM_BYTES, g.stream>>>(slots, S, ntile);
}
static void launch_i8p(Engine& g, const SlotDesc* slots, int S, int ntile) {
  switch (g.cur_kind) {
    case GPSAT_MATERN12: launch_i8p_k<GPSAT_MATERN12>(g, slots, S, ntile); break;
    case GPSAT_MATERN32: launch_i8p_k<GPSAT_MATERN32>(g, slots, S, ntile); break;
    case GPSAT_MATERN52: launch_i8p_k<GPSAT_MATERN52>(g, slots, S, ntile); break;
    default: launch_i8p_k<GPSAT_RBF>(g, slots, S, ntile); break;
  }
}

// One blocked sweep over all occupied slots.
// dmma_slots: number of occupied slots the precision rule moved to FP64 mma.sync for this evaluation, when the caller
// knows it (the fit loop reads it back with its counters); -1 = unknown, launch the FP64 gradient kernel regardless.
static int launch_sweep(Engine& g, int S, int Tmax, int TRmax, bool do_inv, const SlotDesc* slots = nullptr,
                        bool with_lauum = true, int dmma_slots = -1) {
  if (!slots) slots = g.bslots.as<SlotDesc>();
  {
    ProfScope ps(g, PROF_DIAG);
    diag_kernel<<<S, GPSAT_THREADS, DIAG_SMEM_BYTES, g.stream>>>(slots, 0);
  }
  g.stats.kernel_launches++;
  {
    int nstep = 0;
    for (int i = 0; i < Tmax; ++i) nstep += ((TRmax - 1 - i) + (do_inv ? i : 0)) > 0 ? 1 : 0;
    ProfScope ps(g, PROF_STEP, nstep);
    for (int i = 0; i < Tmax; ++i) {
      // step i factors block column i and, in its (i+1, i) tile, the diagonal tile i+1
      int ntile = (TRmax - 1 - i) + (do_inv ? i : 0);
      if (ntile > 0) {
        step_kernel<<<S * ntile, GPSAT_THREADS, STEP_SMEM_BYTES, g.stream>>>(slots, i, do_inv ? 1 : 0, S, ntile,
                                                                                      g.i8_diag_sum ? 1 : 0);
        g.stats.kernel_launches++;
      }
    }
  }
  if (do_inv && with_lauum) {
    ProfScope ps(g, PROF_LAUUM);
    const int ntile = Tmax * (Tmax + 1) / 2;
    if (g.i8_lauum) {
      if (g.lauum_planes8) {
        if (g.i8_lauum_q) {
          const int grid = std::min(g.num_sms, ntile * S);
          if (g.bQ.ensure((size_t)g.num_sms * I8Q_SCRATCH_D * sizeof(double))) return 1;
          lauum_grad_i8q_kernel<<<grid, I8Q_THREADS, LAUUM_I8Q_SMEM_BYTES, g.stream>>>(slots, S, ntile, g.bQ.as<double>());
        } else {
          lauum_grad_i8_kernel<<<dim3(ntile, S), GPSAT_THREADS, LAUUM_I8_SMEM_BYTES, g.stream>>>(slots, 1);
        }
        g.stats.kernel_launches++;
        if (g.i8_rho_max > 0.0 && dmma_slots != 0) {      // slots the precision rule took off the planes for this evaluation
          lauum_grad_kernel<<<dim3(ntile, S), GPSAT_THREADS, LAUUM_SMEM_BYTES, g.stream>>>(slots, 1);
          g.stats.kernel_launches++;
        }
      } else {
        u_rowmax_kernel<<<dim3(Tmax * NB / 8, S), 256, 0, g.stream>>>(slots);
        u_slice_kernel<<<dim3(ntile, S), 256, 0, g.stream>>>(slots);
        if (g.i8_persistent)
          launch_i8p(g, slots, S, ntile);
        else
          lauum_grad_i8_kernel<<<dim3(ntile, S), GPSAT_THREADS, LAUUM_I8_SMEM_BYTES, g.stream>>>(slots, 0);
        g.stats.kernel_launches += 3;
      }
    } else {
      lauum_grad_kernel<<<dim3(ntile, S), GPSAT_THREADS, LAUUM_SMEM_BYTES, g.stream>>>(slots, 0);
      g.stats.kernel_launches++;
    }
  }
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace gpsat

#include "engine_api.inl"
