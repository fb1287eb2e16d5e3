// C ABI of the batched local selection (select_kernels.cuh).  Included at the end of engine_post.inl.
static int select_spec_from(const gpsat_select_cond* conds, int ncond, int C, int CL, SelectSpec& spec) {
  REQUIRE(ncond >= 0 && ncond <= SEL_MAX_COND, "too many select conditions (max 8)");
  REQUIRE(ncond == 0 || conds != nullptr, "conds is NULL");
  spec.ncond = ncond;
  for (int q = 0; q < ncond; ++q) {
    const gpsat_select_cond& c = conds[q];
    REQUIRE(c.kind == 0 || c.kind == 1, "select condition kind must be 0 (compare) or 1 (radius)");
    REQUIRE(c.op >= 0 && c.op <= 4, "select condition op out of range");
    REQUIRE(c.ncol >= 1 && c.ncol <= 3 && (c.kind == 1 || c.ncol == 1), "select condition ncol out of range");
    spec.c[q].kind = c.kind; spec.c[q].op = c.op; spec.c[q].ncol = c.ncol; spec.c[q].val = c.val;
    for (int k = 0; k < 3; ++k) {
      const bool used = k < c.ncol;
      REQUIRE(!used || (c.col[k] >= 0 && c.col[k] < C && c.loc_col[k] >= 0 && c.loc_col[k] < CL),
              "select condition column index out of range");
      spec.c[q].col[k] = used ? c.col[k] : 0;
      spec.c[q].loc[k] = used ? c.loc_col[k] : 0;
    }
  }
  return 0;
}

extern "C" {

int gpsat_local_select(gpsat_engine* h, int64_t N, int32_t C, const double* data, int64_t E, int32_t CL,
                       const double* locs, int32_t ncond, const gpsat_select_cond* conds, int64_t* counts,
                       const int64_t* off, int64_t* idx) {
  REQUIRE(h != nullptr, "engine is NULL");
  REQUIRE(N >= 0 && E >= 0 && C >= 1 && CL >= 1 && E < 65536, "bad sizes");
  REQUIRE(counts != nullptr, "counts is NULL");
  REQUIRE((idx == nullptr) == (off == nullptr), "off and idx go together");
  SelectSpec spec;
  if (int rc = select_spec_from(conds, ncond, C, CL, spec)) return rc;
  Engine& g = h->e;
  begin_call(g);
  if (E == 0) return end_call(g);
  if (N == 0) { for (int64_t e = 0; e < E; ++e) counts[e] = 0; return end_call(g); }
  REQUIRE(data && locs, "data / locs is NULL");
  const int nch = (int)((N + SEL_CHUNK - 1) / SEL_CHUNK);
  const int64_t total = off ? off[E] : 0;
  // g.bX: point table, g.bXs: locations, g.bmisc: chunk counts, g.bout: [counts | off | idx]
  if (g.bX.ensure((size_t)N * C * 8) || g.bXs.ensure((size_t)E * CL * 8) || g.bmisc.ensure((size_t)E * nch * 4) ||
      g.bout.ensure((size_t)(2 * E + 1 + std::max<int64_t>(total, 1)) * 8))
    return 1;
  CUDA_TRY(cudaMemcpyAsync(g.bX.p, data, (size_t)N * C * 8, cudaMemcpyHostToDevice, g.stream));
  CUDA_TRY(cudaMemcpyAsync(g.bXs.p, locs, (size_t)E * CL * 8, cudaMemcpyHostToDevice, g.stream));
  g.stats.h2d_bytes += (double)N * C * 8 + (double)E * CL * 8;
  int64_t* dcounts = g.bout.as<int64_t>();
  int64_t* doff = dcounts + E;
  int64_t* didx = doff + E + 1;
  int* dcnt = g.bmisc.as<int>();
  select_count_kernel<<<dim3(nch, (unsigned)E), 256, 0, g.stream>>>(N, C, g.bX.as<double>(), CL, g.bXs.as<double>(), spec,
                                                                  nch, dcnt);
  select_scan_kernel<<<(unsigned)((E + 127) / 128), 128, 0, g.stream>>>((int)E, nch, dcnt, dcounts);
  g.stats.kernel_launches += 2;
  CUDA_TRY(cudaMemcpyAsync(counts, dcounts, (size_t)E * 8, cudaMemcpyDeviceToHost, g.stream));
  g.stats.d2h_bytes += (double)E * 8;
  if (idx) {
    CUDA_TRY(cudaMemcpyAsync(doff, off, (size_t)(E + 1) * 8, cudaMemcpyHostToDevice, g.stream));
    select_fill_kernel<<<dim3(nch, (unsigned)E), 256, 0, g.stream>>>(N, C, g.bX.as<double>(), CL, g.bXs.as<double>(), spec,
                                                                   nch, dcnt, doff, didx);
    g.stats.kernel_launches++;
    if (total > 0) CUDA_TRY(cudaMemcpyAsync(idx, didx, (size_t)total * 8, cudaMemcpyDeviceToHost, g.stream));
    g.stats.d2h_bytes += (double)total * 8;
  }
  CUDA_TRY(cudaGetLastError());
  if (int rc = end_call(g)) return rc;
  if (idx)
    for (int64_t e = 0; e < E; ++e) REQUIRE(counts[e] == off[e + 1] - off[e], "off does not match the selection counts");
  return 0;
}

}  // extern "C"
