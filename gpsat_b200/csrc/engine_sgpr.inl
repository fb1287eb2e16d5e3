// Host orchestration of the sparse-GPR (SGPR) path: ELBO at fixed theta and prediction.
// Replaces GPflowSGPRModel.get_objective_function_value / predict
// (GPSat/models/gpflow_models.py:860-862, :229-230 inherited) for a batch of experts.
// Included at the end of engine_fit.inl (global namespace, using namespace gpsat).

struct SgprWave {
  std::vector<SgprDesc> descs;
  std::vector<SlotDesc> sk, sb;
};

// Per-slot workspace carved out of g.bsg (all sizes in doubles).
struct SgprLayout {
  size_t ct, aq, w, wb, bm, part, cts, t1, t2, ak, ab, side_k, side_b, total;
  size_t tile_d;
  int Tmax;
};

static SgprLayout sgpr_layout(int npad, int mpad, int Mp, int nsplit, bool with_pred) {
  SgprLayout L{};
  const size_t mm = (size_t)Mp * Mp;
  L.Tmax = Mp / NB;
  L.tile_d = (size_t)L.Tmax * NB * NB;
  const size_t side = 2 * L.tile_d + align_up(L.Tmax, 2) + 4;     // Wd, Ld, logdet, quad, notpd
  size_t o = 0;
  L.ct = o; o += (size_t)npad * Mp;
  L.aq = o; o += (size_t)npad * Mp;
  L.w = o; o += mm;
  L.wb = o; o += with_pred ? mm : 0;
  L.bm = o; o += mm;
  L.part = o; o += (size_t)nsplit * mm;
  L.cts = o; o += with_pred ? (size_t)mpad * Mp : 0;
  L.t1 = o; o += with_pred ? (size_t)mpad * Mp : 0;
  L.t2 = o; o += with_pred ? (size_t)mpad * Mp : 0;
  L.ak = o; o += mm;
  L.ab = o; o += mm;
  L.side_k = o; o += side;
  L.side_b = o; o += side;
  L.total = (o + 1) & ~(size_t)1;
  return L;
}

static void sgpr_fill_slot(SlotDesc& s, int expert, int M, int kind, int D, double* A, double* side,
                           const SgprLayout& L) {
  s = SlotDesc{};
  s.expert = expert;
  s.n = M; s.m = 0;
  s.T = s.TR = (M + 1 + NB - 1) / NB;
  s.ld = s.T * NB;
  s.kind = kind; s.D = D;
  s.A = A;
  s.Wd = side; s.Ld = side + L.tile_d; s.logdet = side + 2 * L.tile_d;
  s.quad = s.logdet + align_up(L.Tmax, 2);
  s.gpart = nullptr;
  s.notpd = reinterpret_cast<int*>(s.quad + 2);
  s.pre = nullptr; s.ldpre = 0;
}

static int sgpr_impl(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                     const double* theta, bool with_pred, double* fmean, double* fvar, double* yvar, double* elbo,
                     int32_t* status) {
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
  const int D = batch->D, P = D + 2;
  int Mmax = 0;
  for (int64_t e = 0; e < E; ++e) {
    const int64_t Me = off_z[e + 1] - off_z[e];
    REQUIRE(Me >= 1 && Me < (1 << 16), "every expert needs between 1 and 65535 inducing points");
    Mmax = std::max<int>(Mmax, (int)Me);
  }
  const bool dev = mem == GPSAT_MEM_DEVICE;
  BatchDev bd;
  if (engine_load_batch(g, batch, mem, with_pred, bd)) return 1;
  const int64_t NZ = off_z[E], Mtot = with_pred ? batch->off_m[E] : 0;
  const double* dZ = Z;
  if (!dev) {
    if (g.bZ.ensure((size_t)NZ * D * 8)) return 1;
    CUDA_TRY(cudaMemcpyAsync(g.bZ.p, Z, (size_t)NZ * D * 8, cudaMemcpyHostToDevice, g.stream));
    g.stats.h2d_bytes += (double)NZ * D * 8;
    dZ = g.bZ.as<double>();
  }
  // hyper-parameters are needed on the host to fill the descriptors
  std::vector<double> hth((size_t)E * P);
  if (dev) {
    CUDA_TRY(cudaMemcpyAsync(hth.data(), theta, (size_t)E * P * 8, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
  } else {
    std::memcpy(hth.data(), theta, (size_t)E * P * 8);
  }
  int64_t nmax = 0, mmax = 0;
  for (int64_t e = 0; e < E; ++e) { nmax = std::max(nmax, g.hn[e]); mmax = std::max(mmax, g.hm[e]); }
  const int npad = align_up((int)nmax, NB), mpad = align_up((int)std::max<int64_t>(mmax, 1), NB);
  const int Mp = align_up(Mmax + 1, NB), MT = Mp / NB;
  // wave size and split-K factor of the SYRK: enough CTAs to fill the machine a few times over
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  const int ntile_l = MT * (MT + 1) / 2;
  int S = (int)std::min<int64_t>(E, 148);
  int nsplit = 1;
  SgprLayout L{};
  for (;;) {
    nsplit = std::max(1, std::min({16, npad / NB, (592 + S * ntile_l - 1) / (S * ntile_l)}));
    L = sgpr_layout(npad, mpad, Mp, nsplit, with_pred);
    if (L.total * 8 * S <= (size_t)((free_b + g.bsg.cap) * 0.8) || S == 1) break;
    S = std::max(1, S / 2);
  }
  // outputs
  size_t o_elbo = 0, o_status = (size_t)E * 8;
  size_t o_pred = (o_status + (size_t)E * 4 + 15) & ~(size_t)15;
  size_t o_zero = o_pred + (size_t)std::max<int64_t>(Mtot, 1) * 3 * 8;
  if (g.bout.ensure(o_zero + (size_t)Mp * 8)) return 1;
  char* base = g.bout.as<char>();
  double* delbo = (dev && elbo) ? elbo : reinterpret_cast<double*>(base + o_elbo);
  int* dstatus = status ? (dev ? status : reinterpret_cast<int*>(base + o_status)) : nullptr;
  double* dfm = dev ? fmean : reinterpret_cast<double*>(base + o_pred);
  double* dfv = dev ? fvar : dfm + Mtot;
  double* dyv = dev ? yvar : dfv + Mtot;
  double* dzero = reinterpret_cast<double*>(base + o_zero);
  CUDA_TRY(cudaMemsetAsync(dzero, 0, (size_t)Mp * 8, g.stream));
  if (g.bsg.ensure(L.total * 8 * S) || g.bslots.ensure(sizeof(SlotDesc) * 2 * S) || g.bmisc.ensure(sizeof(SgprDesc) * S))
    return 1;
  double* ws = g.bsg.as<double>();
  SlotDesc* dsk = g.bslots.as<SlotDesc>();
  SlotDesc* dsb = dsk + S;
  SgprDesc* ddesc = g.bmisc.as<SgprDesc>();
  SgprWave w;
  w.descs.resize(S); w.sk.resize(S); w.sb.resize(S);
  for (int64_t b0 = 0; b0 < E; b0 += S) {
    int nb = 0, mb = 0, Mb = 0, cnt = 0;
    for (int s = 0; s < S; ++s) {
      SgprDesc& d = w.descs[s];
      d = SgprDesc{};
      if (b0 + s >= E) { d.expert = -1; w.sk[s] = SlotDesc{}; w.sb[s] = SlotDesc{}; w.sk[s].expert = w.sb[s].expert = -1; continue; }
      const int e = g.order[b0 + s];
      ++cnt;
      double* base_s = ws + L.total * s;
      d.expert = e;
      d.n = (int)g.hn[e]; d.m = (int)g.hm[e]; d.M = (int)(off_z[e + 1] - off_z[e]);
      d.D = D; d.kind = batch->kernel;
      d.npad = align_up(d.n, NB); d.mpad = align_up(std::max(d.m, 1), NB); d.Mp = align_up(d.M + 1, NB);
      d.X = bd.X + batch->off_n[e] * D; d.y = bd.y + batch->off_n[e];
      d.Z = dZ + off_z[e] * D;
      d.Xs = (with_pred && d.m > 0) ? bd.Xs + batch->off_m[e] * D : nullptr;
      const double* th = hth.data() + (size_t)e * P;
      for (int k = 0; k < D; ++k) d.inv_ls[k] = 1.0 / th[k];
      d.var = th[D]; d.noise = th[D + 1];
      d.Ct = base_s + L.ct; d.Aq = base_s + L.aq; d.W = base_s + L.w; d.WB = base_s + L.wb; d.Bm = base_s + L.bm;
      d.part = base_s + L.part; d.Cts = base_s + L.cts; d.T1 = base_s + L.t1; d.T2 = base_s + L.t2;
      // K slot: the M inducing points as "observations" with y = 0 and noise = jitter
      SlotDesc& k = w.sk[s];
      sgpr_fill_slot(k, e, d.M, d.kind, D, base_s + L.ak, base_s + L.side_k, L);
      k.X = d.Z; k.y = dzero; k.Xs = nullptr;
      for (int q = 0; q < D; ++q) k.inv_ls[q] = d.inv_ls[q];
      k.var = d.var; k.noise = SGPR_JITTER;
      // B slot: factor the precomputed B' (row M = A y / sigma rides along as the "y" row)
      SlotDesc& b = w.sb[s];
      sgpr_fill_slot(b, e, d.M, GPSAT_PRECOMPUTED, 0, base_s + L.ab, base_s + L.side_b, L);
      b.X = nullptr; b.y = d.Bm + (size_t)d.M * d.Mp; b.Xs = nullptr;
      b.var = 0.0; b.noise = 0.0;
      b.pre = d.Bm; b.ldpre = d.Mp;
      nb = std::max(nb, d.npad); mb = std::max(mb, d.mpad); Mb = std::max(Mb, d.Mp);
    }
    CUDA_TRY(cudaStreamSynchronize(g.stream));   // the previous wave has consumed the staging vectors
    CUDA_TRY(cudaMemcpyAsync(ddesc, w.descs.data(), sizeof(SgprDesc) * S, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync(dsk, w.sk.data(), sizeof(SlotDesc) * S, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaMemcpyAsync(dsb, w.sb.data(), sizeof(SlotDesc) * S, cudaMemcpyHostToDevice, g.stream));
    const int MTb = Mb / NB, ntl = MTb * (MTb + 1) / 2;
    sgpr_reset_kernel<<<(2 * S + 127) / 128, 128, 0, g.stream>>>(dsk, 2 * S);
    kuf_build_kernel<<<dim3(nb / NB, MTb, S), GPSAT_THREADS, 0, g.stream>>>(ddesc, 0);
    if (launch_sweep(g, S, MTb, MTb, true, dsk, false)) return 1;
    transpose_w_kernel<<<dim3(Mb / 32, Mb / 32, S), 256, 0, g.stream>>>(ddesc, dsk, 0);
    sgpr_gemm_kernel<<<dim3(MTb * (nb / NB), S), GPSAT_THREADS, SGPR_GEMM_SMEM_BYTES, g.stream>>>(ddesc, SGPR_GEMM_A, 1);
    sgpr_gemm_kernel<<<dim3(ntl * nsplit, S), GPSAT_THREADS, SGPR_GEMM_SMEM_BYTES, g.stream>>>(ddesc, SGPR_GEMM_SYRK, nsplit);
    sgpr_reduce_kernel<<<dim3(std::min(256, (Mb * Mb + 255) / 256), S), 256, 0, g.stream>>>(ddesc, nsplit);
    if (launch_sweep(g, S, MTb, MTb, with_pred, dsb, false)) return 1;
    sgpr_finish_kernel<<<(S + 127) / 128, 128, 0, g.stream>>>(ddesc, dsk, dsb, delbo, dstatus, S);
    g.stats.kernel_launches += 7;
    if (with_pred) {
      transpose_w_kernel<<<dim3(Mb / 32, Mb / 32, S), 256, 0, g.stream>>>(ddesc, dsb, 1);
      kuf_build_kernel<<<dim3(mb / NB, MTb, S), GPSAT_THREADS, 0, g.stream>>>(ddesc, 1);
      sgpr_gemm_kernel<<<dim3((mb / NB) * MTb, S), GPSAT_THREADS, SGPR_GEMM_SMEM_BYTES, g.stream>>>(ddesc, SGPR_GEMM_T1, 1);
      sgpr_gemm_kernel<<<dim3((mb / NB) * MTb, S), GPSAT_THREADS, SGPR_GEMM_SMEM_BYTES, g.stream>>>(ddesc, SGPR_GEMM_T2, 1);
      sgpr_predict_reduce_kernel<<<dim3((unsigned)((mb * 32 + 255) / 256), S), 256, 0, g.stream>>>(ddesc, dsb, bd.off_m,
                                                                                                  dfm, dfv, dyv);
      g.stats.kernel_launches += 5;
    }
    g.stats.ticks++;
    (void)cnt;
  }
  for (int64_t e = 0; e < E; ++e) {
    const double n = (double)g.hn[e], M = (double)(off_z[e + 1] - off_z[e]);
    g.stats.sweep_flops += 2.0 * M * M * M / 3.0 + 2.0 * n * M * M + 2.0 * (double)g.hm[e] * M * M;
  }
  g.stats.evals = E; g.stats.slots = S;
  if (!dev) {
    if (with_pred && Mtot > 0) {
      CUDA_TRY(cudaMemcpyAsync(fmean, dfm, (size_t)Mtot * 8, cudaMemcpyDeviceToHost, g.stream));
      CUDA_TRY(cudaMemcpyAsync(fvar, dfv, (size_t)Mtot * 8, cudaMemcpyDeviceToHost, g.stream));
      CUDA_TRY(cudaMemcpyAsync(yvar, dyv, (size_t)Mtot * 8, cudaMemcpyDeviceToHost, g.stream));
    }
    if (elbo) CUDA_TRY(cudaMemcpyAsync(elbo, delbo, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
    if (status) CUDA_TRY(cudaMemcpyAsync(status, dstatus, (size_t)E * 4, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)Mtot * 24 + (double)E * 12;
  }
  CUDA_TRY(cudaGetLastError());
  return end_call(g);
}

extern "C" {

int gpsat_sgpr_elbo(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                    const double* theta, double* elbo, int32_t* status) {
  return sgpr_impl(h, batch, off_z, Z, mem, theta, false, nullptr, nullptr, nullptr, elbo, status);
}

int gpsat_sgpr_predict(gpsat_engine* h, const gpsat_batch* batch, const int64_t* off_z, const double* Z, int mem,
                       const double* theta, double* fmean, double* fvar, double* yvar, double* elbo, int32_t* status) {
  return sgpr_impl(h, batch, off_z, Z, mem, theta, true, fmean, fvar, yvar, elbo, status);
}

}  // extern "C"
