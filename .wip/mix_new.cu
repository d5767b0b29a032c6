// ----------------------------------------------------------------------------------------------
// mix_prep_kernel: one record per (period, channel) for mix_kernel -- the per-ear delay ramp in
// samples (validated against the history length, then clamped) and the channel's NON-ZERO gains as
// short lists of (loudspeaker, start value, per-period delta).  Panning gains are sparse: a point
// source feeds 2-4 of the 24 loudspeakers, so mix_kernel walks these lists instead of the dense
// [L x N] tables of rcl::GainMatrix (per_ear_delay.cpp:31-36, dsp.cpp:127-129); terms it skips would
// have added exact zeros.  A channel with more than kMixK non-zero gains (extent, "dense") is marked
// and takes the table-driven path of mix_kernel.
// Record (24 words): [d0L, ddL, d0R, ddR | nd + 256 nq, 4 direct speakers (bytes), 4 diffuse
// speakers, - | 4 x (g0, dg) direct | 4 x (q0, dq) diffuse]; nd / nq = 255: dense.
// grid: (ceil((N+M)/128), periods); block 128; one thread per channel (coalesced table reads).
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) mix_prep_kernel(Layout p, int T)
{
  const int c = blockIdx.x * 128 + threadIdx.x, k = p.k0 + blockIdx.y;
  const int C = p.N + p.M, L = p.L;
  if (c >= C) return;
  const bool obj = c < p.N;
  const int ci = obj ? c : c - p.N, ld = obj ? p.N : p.M;
  float *rec = p.mix_recs + ((size_t)k * C + c) * kMixRecWords;
  // delays of periods k-1 (slot k) and k (slot k+1), seconds -> samples
  const float *ds = (obj ? p.d_dir : p.d_ds) + (size_t)k * 2 * ld + 2 * ci;
  const float dmax = (float)(p.hist - 4);
  float4 dl;
  {
    float a0 = ds[0] * p.fs, a1 = ds[2 * ld] * p.fs, b0 = ds[1] * p.fs, b1 = ds[2 * ld + 1] * p.fs;
    // a table delay outside [0, history - 4] cannot be rendered (the reference allows up to 1 s,
    // per_ear_delay.cpp:25): flagged, reported by the call, and clamped so that no read leaves the tile
    if (!(a1 >= 0.f && a1 <= dmax) || !(b1 >= 0.f && b1 <= dmax)) atomicOr(p.err_flag, kErrDelayRange);
    a0 = fminf(fmaxf(a0, 0.f), dmax); a1 = fminf(fmaxf(a1, 0.f), dmax);
    b0 = fminf(fmaxf(b0, 0.f), dmax); b1 = fminf(fmaxf(b1, 0.f), dmax);
    dl = make_float4(a0, a1 - a0, b0, b1 - b0);
  }
  unsigned cnt[2] = {0u, 0u}, spk[2] = {0u, 0u};
  const int nlists = obj ? 2 : 1;
  for (int w = 0; w < nlists; w++) {
    const float *g = (w ? p.g_dif : (obj ? p.g_dir : p.g_ds)) + (size_t)k * L * ld + ci;
    float2 *dst = reinterpret_cast<float2 *>(rec + 8 + 8 * w);
    for (int l = 0; l < L; l++) {
      const float g0 = g[(size_t)l * ld], g1 = g[(size_t)(L + l) * ld];
      if (g0 != 0.f || g1 != 0.f) {
        if (cnt[w] < (unsigned)kMixK) {
          dst[cnt[w]] = make_float2(g0, g1 - g0);
          spk[w] |= (unsigned)l << (8 * cnt[w]);
        }
        cnt[w]++;
      }
    }
    if (cnt[w] > (unsigned)kMixK) cnt[w] = 255u;
  }
  reinterpret_cast<float4 *>(rec)[0] = dl;
  reinterpret_cast<uint4 *>(rec)[1] = make_uint4(cnt[0] | cnt[1] << 8, spk[0], spk[1], 0u);
}

// ----------------------------------------------------------------------------------------------
// mix_kernel: objects + DirectSpeakers onto the 2L direct bus and the L diffuse bus.
//   cdir[2l+e][t] = sum_c g_c,l(t) * D[x_c, tau_c,e(t)](t)      (per_ear_delay.cpp:22-56)
//   qdif[l][t]    = sum_c q_c,l(t) * x_c(t)                     (dsp.cpp:127-129)
// grid: (periods * B / kMixTile, S); block: kMixTile.  One output sample per thread; the 24 (left,
// right) direct-bus accumulators are packed pairs, so one FFMA2 adds a channel to both ears of a
// loudspeaker.  The CTA walks its channels in chunks of kMixChunk: input rows (history + tile) and
// the chunk's records arrive by bulk copies (TMA, 1-D) on an mbarrier, double buffered, so the copies
// of chunk i+1 run under the arithmetic of chunk i.  Per channel: two Lagrange-3 reads at the ramped
// per-ear delays, then one ramped-gain FMA + one FFMA2 per non-zero loudspeaker.  Which accumulator a
// list entry feeds is the same for all threads (the lists are per channel and period), so the
// dispatch is a CTA-uniform switch.
// ----------------------------------------------------------------------------------------------
#define BEAR_CASES24(M) M(0) M(1) M(2) M(3) M(4) M(5) M(6) M(7) M(8) M(9) M(10) M(11) M(12) M(13) M(14) M(15) M(16) \
  M(17) M(18) M(19) M(20) M(21) M(22) M(23)

__device__ __forceinline__ void mix_add_direct(f32x2 (&acc)[kMaxSpeakers], unsigned l, f32x2 g2, f32x2 x2)
{
#define BEAR_CASE(i) case i: acc[i] = fma2(g2, x2, acc[i]); break;
  switch (l) { BEAR_CASES24(BEAR_CASE) default: break; }
#undef BEAR_CASE
}
__device__ __forceinline__ void mix_add_diffuse(float (&acc)[kMaxSpeakers], unsigned l, float g, float x)
{
#define BEAR_CASE(i) case i: acc[i] = fmaf(g, x, acc[i]); break;
  switch (l) { BEAR_CASES24(BEAR_CASE) default: break; }
#undef BEAR_CASE
}

__device__ __forceinline__ void mbar_wait(unsigned bar_addr, unsigned parity)
{
  unsigned done;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done) : "r"(bar_addr), "r"(parity) : "memory");
  } while (!done);
}

__global__ void __launch_bounds__(kMixTile, 4) mix_kernel(Layout p, int T)
{
  extern __shared__ __align__(16) float smx[];
  const int XW = p.hist + kMixTile;
  float *const xt0 = smx;                                    // [2][kMixChunk][XW] input rows
  float *const rc0 = smx + 2 * kMixChunk * XW;               // [2][kMixChunk][kMixRecWords] records
  __shared__ __align__(8) unsigned long long bars[2];
  const int tid = threadIdx.x;
  const int t = p.k0 * p.B + blockIdx.x * kMixTile + tid;
  const int k = p.k0 + (blockIdx.x * kMixTile) / p.B;
  const int n = t - k * p.B;
  const int s = blockIdx.y;
  const int C = p.N + p.M, L = p.L;
  const int per = (C + p.S - 1) / p.S;
  const int c0 = s * per, c1 = min(C, c0 + per);
  const float invB = 1.0f / (float)p.B;
  const float rg = (float)(n + p.goff) * invB;
  const float rd = (float)(n + p.doff) * invB;

  f32x2 acc[kMaxSpeakers];     // (left, right) of loudspeaker l
  float accQ[kMaxSpeakers];    // diffuse bus
#pragma unroll
  for (int l = 0; l < kMaxSpeakers; l++) { acc[l] = 0ull; accQ[l] = 0.f; }

  const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = (c1 - c0 + kMixChunk - 1) / kMixChunk;
  // warp 0 requests chunk ci: one bulk copy per input row (lane = row) and one for the records
  auto request = [&](int ci) {
    const int cb = c0 + ci * kMixChunk, cnt = min(kMixChunk, c1 - cb), st = ci & 1;
    const unsigned bar = bar0 + 8 * st;
    if (tid == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                   ::"r"(bar), "r"(cnt * (XW + kMixRecWords) * 4) : "memory");
    __syncwarp();
    if (tid < cnt) {
      const float *src = p.xin + (size_t)(cb + tid) * p.xs + (t - tid);   // column (hist + t0) - hist
      const unsigned dst = (unsigned)__cvta_generic_to_shared(xt0 + (st * kMixChunk + tid) * XW);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "l"(src), "r"(XW * 4), "r"(bar) : "memory");
    }
    if (tid == 0) {
      const float *src = p.mix_recs + ((size_t)k * C + cb) * kMixRecWords;
      const unsigned dst = (unsigned)__cvta_generic_to_shared(rc0 + st * kMixChunk * kMixRecWords);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "l"(src), "r"(cnt * kMixRecWords * 4), "r"(bar) : "memory");
    }
  };
  if (tid < 32 && nchunks > 0) request(0);
  for (int ci = 0; ci < nchunks; ci++) {
    const int st = ci & 1;
    if (tid < 32 && ci + 1 < nchunks) request(ci + 1);   // its buffer was released by the barrier below
    mbar_wait(bar0 + 8 * st, (ci >> 1) & 1);
    const int cb = c0 + ci * kMixChunk, cnt = min(kMixChunk, c1 - cb);
    const float *xrow = xt0 + st * kMixChunk * XW + p.hist + tid;
    const float *recs = rc0 + st * kMixChunk * kMixRecWords;
    // channels two at a time: both channels' delayed reads are issued before either's accumulation
    constexpr int UC = 2;
    for (int j0 = 0; j0 < cnt; j0 += UC) {
      f32x2 x2[UC];
#pragma unroll
      for (int u = 0; u < UC; u++) {
        const int j = min(j0 + u, cnt - 1);
        const float4 dl = *reinterpret_cast<const float4 *>(recs + j * kMixRecWords);
        const float *xr = xrow + j * XW;
        x2[u] = delayed_read2(xr, ramped_delay(dl.x, dl.y, rd), xr, ramped_delay(dl.z, dl.w, rd));
      }
#pragma unroll
      for (int u = 0; u < UC; u++) {
        const int j = j0 + u;
        if (j >= cnt) break;
        const float *rec = recs + j * kMixRecWords;
        const uint4 meta = *reinterpret_cast<const uint4 *>(rec + 4);
        const unsigned nd = meta.x & 0xffu, nq = meta.x >> 8;
        if (nd <= (unsigned)kMixK) {
          unsigned spk = meta.y;
#pragma unroll 1
          for (unsigned i = 0; i < nd; i++, spk >>= 8) {
            const float2 gd = *reinterpret_cast<const float2 *>(rec + 8 + 2 * i);
            mix_add_direct(acc, spk & 0xffu, splat2(fmaf(rg, gd.y, gd.x)), x2[u]);
          }
        } else {   // dense gains: the table itself, every thread the same address
          const int c = cb + j;
          const bool obj = c < p.N;
          const int ld = obj ? p.N : p.M;
          const float *g = (obj ? p.g_dir + (size_t)k * L * ld + c : p.g_ds + (size_t)k * L * ld + (c - p.N));
#pragma unroll
          for (int l = 0; l < kMaxSpeakers; l++)
            if (l < L) {
              const float g0 = __ldg(g + (size_t)l * ld), g1 = __ldg(g + (size_t)(L + l) * ld);
              acc[l] = fma2(splat2(fmaf(rg, g1 - g0, g0)), x2[u], acc[l]);
            }
        }
        if (nq) {
          const float x0 = xrow[j * XW];
          if (nq <= (unsigned)kMixK) {
            unsigned spk = meta.z;
#pragma unroll 1
            for (unsigned i = 0; i < nq; i++, spk >>= 8) {
              const float2 qd = *reinterpret_cast<const float2 *>(rec + 16 + 2 * i);
              mix_add_diffuse(accQ, spk & 0xffu, fmaf(rg, qd.y, qd.x), x0);
            }
          } else {
            const int c = cb + j;
            const float *g = p.g_dif + (size_t)k * L * p.N + c;
#pragma unroll
            for (int l = 0; l < kMaxSpeakers; l++)
              if (l < L) {
                const float g0 = __ldg(g + (size_t)l * p.N), g1 = __ldg(g + (size_t)(L + l) * p.N);
                accQ[l] = fmaf(fmaf(rg, g1 - g0, g0), x0, accQ[l]);
              }
          }
        }
      }
    }
    __syncthreads();   // every thread is done with buffer st before warp 0 requests chunk ci + 2 into it
  }
  // split 0 writes the bus rows themselves, the other splits their partial sums (bus_sum_kernel adds them)
  float *od, *oq;
  size_t os;
  if (s == 0) {
    od = p.cdir + (size_t)p.B + t; oq = p.qdif + (size_t)p.B + t; os = p.cs;
  } else {
    od = p.bus_part + (size_t)(s - 1) * 3 * L * p.bps + (t - p.k0 * p.B); oq = od + (size_t)2 * L * p.bps; os = p.bps;
  }
#pragma unroll
  for (int l = 0; l < kMaxSpeakers; l++) {
    if (l < L) {
      float vL, vR;
      unpack2(acc[l], vL, vR);
      od[(size_t)(2 * l) * os] = vL;
      od[(size_t)(2 * l + 1) * os] = vR;
      oq[(size_t)l * os] = accQ[l];
    }
  }
}

// bus_sum_kernel: bus row += the partial sums of object splits 1..S-1, in split order (launches that
// are too short to fill the GPU with one CTA per tile split their channels over several CTAs).
// grid: (ceil(periods * B / 1024), 3L); block 256; four samples per thread.
__global__ void __launch_bounds__(256) bus_sum_kernel(Layout p, int T)
{
  const int row = blockIdx.y, L = p.L;
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i >= (size_t)T * p.B) return;
  float *dst = (row < 2 * L ? p.cdir + (size_t)row * p.cs : p.qdif + (size_t)(row - 2 * L) * p.cs) + (size_t)(p.k0 + 1) * p.B + i;
  float4 v = *reinterpret_cast<float4 *>(dst);
  for (int s = 1; s < p.S; s++) {
    const float4 a = *reinterpret_cast<const float4 *>(p.bus_part + ((size_t)(s - 1) * 3 * L + row) * p.bps + i);
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
  }
  *reinterpret_cast<float4 *>(dst) = v;
}

size_t mix_smem_bytes(int hist);
void launch_mix_prep(const Layout &p, int T, cudaStream_t st)
{
  mix_prep_kernel<<<dim3((p.N + p.M + 127) / 128, T), 128, 0, st>>>(p, T);
}
void launch_mix(const Layout &p, int T, cudaStream_t st)
{
  dim3 grid(T * p.B / kMixTile, p.S);
  mix_kernel<<<grid, kMixTile, mix_smem_bytes(p.hist), st>>>(p, T);
}
void launch_bus_sum(const Layout &p, int T, cudaStream_t st)   // only launches with S > 1
{
  bus_sum_kernel<<<dim3((T * p.B + 1023) / 1024, 3 * p.L), 256, 0, st>>>(p, T);
}

// validate_views_kernel: tables written by a device-side producer are not seen by the host, so the
// view indices of a device-resident call are checked here: out-of-range ones raise the error flag and
// are replaced by view 0, so that no later kernel indexes the filter table out of bounds.
__global__ void __launch_bounds__(256) validate_views_kernel(Layout p, int T)
{
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= T) return;
  const int v = p.view[1 + i];
  if (v < 0 || v >= p.V) {
    atomicOr(p.err_flag, kErrViewRange);
    p.view[1 + i] = 0;
  }
}
void launch_validate_views(const Layout &p, int T, cudaStream_t st)
{
  validate_views_kernel<<<(T + 255) / 256, 256, 0, st>>>(p, T);
}

