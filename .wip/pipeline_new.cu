// Queue the kernels for T blocks whose audio and tables are already in the staging areas.
// A call may be issued in chunks of `chunk` periods (all kernels of chunk c, then chunk c+1, ...;
// chunk <= 0: all at once): every intermediate buffer is indexed by the period number within the
// call, so later chunks just read earlier columns.  `before_chunk(c, k0, n)` runs on the host before
// chunk c is queued (the host-buffer path uses it to make the stream wait for that chunk's upload),
// `after_chunk(c, k0, n)` after it.  Nothing here allocates unless per-kernel profiling is on.
template <class Before, class After>
void run_pipeline(bear_cuda_renderer *r, int T, int chunk, Before &&before_chunk, After &&after_chunk,
                  bool device_tables = false)
{
  Layout &p = r->lay;
  cudaStream_t st = r->stream;
  std::vector<cudaEvent_t> chain;
  std::vector<int> ids;
  auto mark = [&](int kernel_id) {
    if (!r->profiling) return;
    cudaEvent_t e;
    if (r->ev_free.empty()) CK(cudaEventCreate(&e));
    else { e = r->ev_free.back(); r->ev_free.pop_back(); }
    CK(cudaEventRecord(e, st));
    chain.push_back(e);
    if (kernel_id >= 0) ids.push_back(kernel_id);
  };
  if (chunk <= 0 || chunk > T) chunk = T;
  const bool shared_scene = p.W > 1;
  // bus rows of this call; a renderer that shares a scene alternates between its two sets, so that a
  // peer still reading the previous call's rows is never overwritten (one barrier per chunk suffices)
  float *const bus = r->bus[r->parity];
  p.cdir = bus;
  p.qdif = bus + (size_t)2 * p.L * p.cs;
  for (int w = 0; w < p.W; w++) {
    const float *pb = shared_scene ? r->peer_bus[w][r->parity] : bus;
    p.peer_cdir[w] = pb;
    p.peer_qdif[w] = pb + (size_t)2 * p.L * p.cs;
  }
  p.S = object_splits_for(r, chunk);     // one value per call: chunk c+1 reads the last period of chunk c
  uint64_t launches = 0;
  if (device_tables) { launch_validate_views(p, T, st); launches++; }
  int nchunks = 0;
  for (int k0 = 0; k0 < T; nchunks++) {
    const int n = std::min(chunk, T - k0);
    before_chunk(nchunks, k0, n);
    p.k0 = k0;
    if (p.N + p.M > 0) {
      mark(BEAR_CUDA_KERNEL_MIX_AUX); launch_mix_prep(p, n, st);
      mark(BEAR_CUDA_KERNEL_MIX); launch_mix(p, n, st);
      launches += 2;
      if (p.S > 1) { mark(BEAR_CUDA_KERNEL_MIX_AUX); launch_bus_sum(p, n, st); launches++; }
    }
    if (p.H > 0) { mark(BEAR_CUDA_KERNEL_HOA_MIX); launch_hoa_mix(p, n, st); launches++; }
    if (shared_scene) {
      // every rank's bus rows of this chunk are complete before anyone reads them
      mark(BEAR_CUDA_KERNEL_EXCHANGE);
      NCK(nccl().AllReduce(r->d_barrier, r->d_barrier, 1, kNcclInt32, kNcclSum, r->comm, st));
    }
    if (p.dif && p.l1 > p.l0) { mark(BEAR_CUDA_KERNEL_DECORRELATE); launch_decorrelate(p, n, st); launches++; }
    if (p.l1 > p.l0 || p.H > 0) { mark(BEAR_CUDA_KERNEL_BUS_FDL); launch_bus_fdl(p, n, st); launches++; }
    p.KS = cmac_splits_for(r, n);
    p.ks_dev = cmac_groups_by_view(p, n);
    mark(BEAR_CUDA_KERNEL_CMAC); launches += launch_cmac(p, n, st);
    mark(BEAR_CUDA_KERNEL_IFFT_OUT); launch_ifft_out(p, n, st); launches++;
    p.KS = r->ks_cfg;
    p.ks_dev = 0;
    if (shared_scene) {
      // the stereo partial mixes of the ranks are summed (north_star: "only per-block stereo partial
      // mixes are summed with NCCL"), chunk by chunk so that the sum of chunk c runs under chunk c+1
      mark(BEAR_CUDA_KERNEL_EXCHANGE);
      NCK(nccl().GroupStart());
      for (int e = 0; e < 2; e++) {
        float *row = p.out + (size_t)e * p.os + (size_t)k0 * p.B;
        if (r->comm_root >= 0)
          NCK(nccl().Reduce(row, row, (size_t)n * p.B, kNcclFloat32, kNcclSum, r->comm_root, r->comm, st));
        else
          NCK(nccl().AllReduce(row, row, (size_t)n * p.B, kNcclFloat32, kNcclSum, r->comm, st));
      }
      NCK(nccl().GroupEnd());
    }
    after_chunk(nchunks, k0, n);
    k0 += n;
  }
  p.k0 = 0;
  mark(BEAR_CUDA_KERNEL_CARRY);

  CarryPlan plan{};
  int n = 0, rows = 0;
  auto add = [&](float *base, int nrows, long long stride, int count, long long shift) {
    if (nrows <= 0 || count <= 0) return;
    plan.e[n] = CarryRow{base, nrows, stride, count, shift};
    plan.row_start[n] = rows;
    rows += nrows;
    n++;
  };
  const long long TB = (long long)T * p.B;
  add(p.xin, p.N + p.M + p.H, (long long)p.xs, p.hist, TB);
  {
    // the last period of the bus rows becomes the prefix of the rows the NEXT call writes
    float *next = shared_scene ? r->bus[r->parity ^ 1] : bus;
    const long long from = (long long)(bus - next) + TB;
    if (p.N + p.M > 0) add(next, 2 * p.L, (long long)p.cs, p.B, from);
    if (p.N > 0) add(next + (size_t)2 * p.L * p.cs, p.L, (long long)p.cs, p.B, from);
  }
  if (p.dif) add(p.w, p.L, (long long)p.ws, p.B + p.hist, TB);
  if (p.H > 0) {
    add(p.amb, p.A, (long long)p.ws, p.B + p.hist, TB);
    add(p.g_hoa, 1, 0, p.A * p.H, (long long)T * p.A * p.H);
  }
  if (p.N > 0) {
    add(p.g_dir, 1, 0, p.L * p.N, (long long)T * p.L * p.N);
    add(p.g_dif, 1, 0, p.L * p.N, (long long)T * p.L * p.N);
    add(p.d_dir, 1, 0, 2 * p.N, (long long)T * 2 * p.N);
  }
  if (p.M > 0) {
    add(p.g_ds, 1, 0, p.L * p.M, (long long)T * p.L * p.M);
    add(p.d_ds, 1, 0, 2 * p.M, (long long)T * 2 * p.M);
  }
  add(p.d_stat, 1, 0, 2 * 2 * p.L, (long long)T * 2 * p.L);
  add(reinterpret_cast<float *>(p.view), 1, 0, 1, T);
  plan.n = n;
  plan.row_start[n] = rows;
  plan.fdl = p.fdl;
  plan.fdl_stride = (long long)p.R * p.B;
  plan.fdl_channels = p.CH;
  plan.fdl_B = p.B;
  plan.fdl_P = p.P;
  plan.fdl_T = T;
  launch_carry(plan, st);
  launches++;
  mark(-1);
  if (r->profiling) {
    r->ev_pending.push_back(std::move(chain));
    r->ev_kernel.push_back(std::move(ids));
  }
  CK(cudaGetLastError());
  r->launches += launches;
  r->calls++;
  if (shared_scene) r->parity ^= 1;
}

void run_pipeline(bear_cuda_renderer *r, int T, bool device_tables = false)
{
  run_pipeline(r, T, r->chunk, [](int, int, int) {}, [](int, int, int) {}, device_tables);
}

// Error bits raised by the kernels of the calls queued so far (a table delay out of range, a view
// index out of range on the device-resident path); synchronises the render stream.
void fetch_kernel_errors(bear_cuda_renderer *r)
{
  CK(cudaMemcpyAsync(r->h_err, r->lay.err_flag, sizeof(int), cudaMemcpyDeviceToHost, r->stream));
  CK(cudaStreamSynchronize(r->stream));
  const int bits = r->h_err[0];
  if (!bits) return;
  CK(cudaMemsetAsync(r->lay.err_flag, 0, sizeof(int), r->stream));
  r->h_err[0] = 0;
  if (bits & kErrViewRange) throw InvalidArgument{"brir_index out of range (device tables); view 0 was rendered instead"};
  throw InvalidArgument{"a table delay is negative or longer than history_samples - 4 samples (" +
                        std::to_string(r->lay.hist - 4) + "); it was clamped for this call -- create the renderer "
                        "with a larger max_delay_samples"};
}

