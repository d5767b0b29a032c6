// ----------------------------------------------------------------------------------------------
// The gain mix of objects and DirectSpeakers onto the bus (a2, a3, a4, a7 of SURVEY.md section 8).
//
// Panning gains are sparse -- a point source feeds 2-4 of the 24 loudspeakers -- and the reference's
// dense [L x N] rcl::GainMatrix (per_ear_delay.cpp:31-36, dsp.cpp:127-129) spends its time adding
// zeros.  Here every (period, channel) gets a RECORD that lists its non-zero loudspeakers in up to
// kMixSlots "slots" (mix_prep_kernel), the channels of a period are SORTED by that loudspeaker set
// (mix_order_kernel), and mix_kernel walks them in sorted order with kMixSlots packed accumulators:
// all channels of a run share the slot -> loudspeaker map, so the inner loop has no table look-up, no
// indexed accumulator and no branch that depends on the gains; when the loudspeaker set changes the
// slot sums are added to the bus rows in global memory (a few dozen times per tile of 1024 channels).
// Channels with more than kMixSlots non-zero loudspeakers (extent, "dense") are mixed by
// mix_dense_kernel from the tables themselves.  Skipped terms would have added exact zeros; the order
// of summation over channels differs from the reference's, which is within the FP32 bar.
// ----------------------------------------------------------------------------------------------

// mix_prep_kernel: the record and the sort key of every (period, channel).
// Record (kMixRecWords = 40 words):
//   [0..3]   per-ear delay ramp in samples: d0L, ddL, d0R, ddR (validated, then clamped to the history)
//   [4]      number of slots | has_diffuse << 8        [5], [6] loudspeaker of slots 0..3, 4..7 (bytes)
//   [8..23]  direct gains, slot pairs:  (g0[2i], g0[2i+1], dg[2i], dg[2i+1]),  i = 0..3
//   [24..39] diffuse gains, same layout (objects only)
// Key: bit 31 dense | loudspeaker mask << 1 | has_diffuse.
// grid: (ceil((N+M)/128), periods); block 128; one thread per channel (coalesced table reads).
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
  const float *gd = (obj ? p.g_dir : p.g_ds) + (size_t)k * L * ld + ci;   // [slot k: L rows][slot k+1: L rows]
  const float *gq = p.g_dif + (size_t)k * L * ld + ci;
  unsigned mask = 0u, any_q = 0u;
  for (int l = 0; l < L; l++) {
    const bool nzd = gd[(size_t)l * ld] != 0.f || gd[(size_t)(L + l) * ld] != 0.f;
    const bool nzq = obj && (gq[(size_t)l * ld] != 0.f || gq[(size_t)(L + l) * ld] != 0.f);
    if (nzd || nzq) mask |= 1u << l;
    if (nzq) any_q = 1u;
  }
  const int ns = __popc(mask);
  unsigned *keys = p.mix_keys + (size_t)k * C;
  if (ns > kMixSlots) {
    keys[c] = 0x80000000u | any_q;
    reinterpret_cast<float4 *>(rec)[0] = dl;
    reinterpret_cast<uint4 *>(rec)[1] = make_uint4(255u | any_q << 8, 0u, 0u, 0u);
    return;
  }
  keys[c] = mask << 1 | any_q;
  float g0[kMixSlots], dg[kMixSlots], q0[kMixSlots], dq[kMixSlots];
  unsigned spk[2] = {0u, 0u};
  unsigned m = mask;
#pragma unroll
  for (int s = 0; s < kMixSlots; s++) {
    g0[s] = dg[s] = q0[s] = dq[s] = 0.f;
    if (m) {
      const int l = __ffs(m) - 1;
      m &= m - 1;
      spk[s >> 2] |= (unsigned)l << (8 * (s & 3));
      const float a = gd[(size_t)l * ld], b = gd[(size_t)(L + l) * ld];
      g0[s] = a; dg[s] = b - a;
      if (obj) {
        const float qa = gq[(size_t)l * ld], qb = gq[(size_t)(L + l) * ld];
        q0[s] = qa; dq[s] = qb - qa;
      }
    }
  }
  float4 *r4 = reinterpret_cast<float4 *>(rec);
  r4[0] = dl;
  reinterpret_cast<uint4 *>(rec)[1] = make_uint4((unsigned)ns | any_q << 8, spk[0], spk[1], 0u);
#pragma unroll
  for (int i = 0; i < kMixSlots / 2; i++) {
    r4[2 + i] = make_float4(g0[2 * i], g0[2 * i + 1], dg[2 * i], dg[2 * i + 1]);
    r4[2 + kMixSlots / 2 + i] = make_float4(q0[2 * i], q0[2 * i + 1], dq[2 * i], dq[2 * i + 1]);
  }
}

// mix_order_kernel: the channels of one (period, object split) sorted by key (bitonic sort in shared
// memory), written as order[k][c0 + i] = channel | has_diffuse << 30 | run_start << 31, where run_start
// marks the first channel of a run with a new loudspeaker set; dense channels sort to the end.
// counts[k][split] = (listed channels, dense channels).
// grid: (periods, S); block 256; dynamic shared memory: 8 bytes x the next power of two >= channels per split.
__global__ void __launch_bounds__(256) mix_order_kernel(Layout p, int T)
{
  extern __shared__ unsigned long long skey[];
  const int k = p.k0 + blockIdx.x, s = blockIdx.y;
  const int C = p.N + p.M;
  const int per = (C + p.S - 1) / p.S;
  const int c0 = min(C, s * per), c1 = min(C, c0 + per), len = c1 - c0;
  int n2 = 1;
  while (n2 < len) n2 <<= 1;
  const unsigned *keys = p.mix_keys + (size_t)k * C;
  for (int i = threadIdx.x; i < n2; i += 256)
    skey[i] = i < len ? ((unsigned long long)keys[c0 + i] << 32 | (unsigned)(c0 + i)) : ~0ull;
  __syncthreads();
  for (int size = 2; size <= n2; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < n2 / 2; i += 256) {
        const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
        const bool up = (lo & size) == 0;
        const unsigned long long a = skey[lo], b = skey[hi];
        if ((a > b) == up) { skey[lo] = b; skey[hi] = a; }
      }
      __syncthreads();
    }
  unsigned *order = p.mix_order + (size_t)k * C + c0;
  int ndense = 0;
  for (int i = threadIdx.x; i < len; i += 256) {
    const unsigned key = (unsigned)(skey[i] >> 32), c = (unsigned)skey[i];
    const unsigned prev = i ? (unsigned)(skey[i - 1] >> 32) : ~0u;
    const unsigned start = (i == 0 || (key >> 1) != (prev >> 1)) ? 1u : 0u;
    order[i] = c | (key & 1u) << 30 | start << 31;
    ndense += key >> 31;
  }
  __shared__ int sdense;
  if (threadIdx.x == 0) sdense = 0;
  __syncthreads();
  if (ndense) atomicAdd(&sdense, ndense);
  __syncthreads();
  if (threadIdx.x == 0) {
    p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2] = len - sdense;
    p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2 + 1] = sdense;
  }
}

__device__ __forceinline__ void mbar_wait(unsigned bar_addr, unsigned parity)
{
  unsigned done;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done) : "r"(bar_addr), "r"(parity) : "memory");
  } while (!done);
}

// ----------------------------------------------------------------------------------------------
// mix_kernel: the listed channels of one object split onto the 2L direct and L diffuse bus rows.
//   cdir[2l+e][t] += g_c,l(t) * D[x_c, tau_c,e(t)](t)      (per_ear_delay.cpp:22-56)
//   qdif[l][t]    += q_c,l(t) * x_c(t)                     (dsp.cpp:127-129)
// grid: (periods * B / kMixTile, S); block: kMixTile; one output sample per thread.  The CTA zeroes
// its tile of the 3L rows, then walks its channels in sorted order in chunks of kMixChunk: warp 0
// reads the chunk's order entries and requests, per channel, the input row (history + tile) and the
// record by bulk copies (TMA, 1-D) on an mbarrier, double buffered, so the copies of chunk i+1 run
// under the arithmetic of chunk i.  Per channel: two Lagrange-3 reads at the ramped per-ear delays,
// then per slot pair one packed gain ramp and two FFMA2 into the (left, right) slot accumulators.
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kMixTile, 5) mix_kernel(Layout p, int T)
{
  extern __shared__ __align__(16) float smx[];
  const int XW = p.hist + kMixTile;
  float *const xt0 = smx;                                    // [2][kMixChunk][XW] input rows
  float *const rc0 = smx + 2 * kMixChunk * XW;               // [2][kMixChunk][kMixRecWords] records
  __shared__ __align__(8) unsigned long long bars[2];
  __shared__ unsigned sord[2][kMixChunk];                    // order entries of the staged chunks
  const int tid = threadIdx.x;
  const int t = p.k0 * p.B + blockIdx.x * kMixTile + tid;
  const int k = p.k0 + (blockIdx.x * kMixTile) / p.B;
  const int n = t - k * p.B;
  const int s = blockIdx.y;
  const int C = p.N + p.M, L = p.L;
  const int per = (C + p.S - 1) / p.S;
  const int c0 = min(C, s * per);
  const int nlisted = p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2];
  const float invB = 1.0f / (float)p.B;
  const float rg = (float)(n + p.goff) * invB;
  const float rd = (float)(n + p.doff) * invB;
  const f32x2 rg2 = splat2(rg);

  // split 0 writes the bus rows themselves, the other splits their partial sums (bus_sum_kernel adds them)
  float *od, *oq;
  size_t os;
  if (s == 0) {
    od = p.cdir + (size_t)p.B + t; oq = p.qdif + (size_t)p.B + t; os = p.cs;
  } else {
    od = p.bus_part + (size_t)(s - 1) * 3 * L * p.bps + (t - p.k0 * p.B); oq = od + (size_t)2 * L * p.bps; os = p.bps;
  }
  for (int l = 0; l < L; l++) { od[(size_t)(2 * l) * os] = 0.f; od[(size_t)(2 * l + 1) * os] = 0.f; oq[(size_t)l * os] = 0.f; }

  f32x2 acc[kMixSlots];          // (left, right) sum of slot i over the current run
  f32x2 accQ[kMixSlots / 2];     // diffuse sums of slots (2i, 2i+1)
#pragma unroll
  for (int i = 0; i < kMixSlots; i++) acc[i] = 0ull;
#pragma unroll
  for (int i = 0; i < kMixSlots / 2; i++) accQ[i] = 0ull;
  unsigned nact = 0u, spk_lo = 0u, spk_hi = 0u;   // slots in use and their loudspeakers (same in all threads)
  bool qdirty = false;
  // end of a run: the slot sums go to the bus rows of their loudspeakers
  auto flush = [&]() {
#pragma unroll
    for (int i = 0; i < kMixSlots; i++) {
      if (i < (int)nact) {
        const unsigned l = ((i < 4 ? spk_lo : spk_hi) >> (8 * (i & 3))) & 0xffu;
        float vL, vR;
        unpack2(acc[i], vL, vR);
        float *pd = od + (size_t)(2 * l) * os;
        pd[0] += vL;
        pd[os] += vR;
        if (qdirty) {
          float qa, qb;
          unpack2(accQ[i >> 1], qa, qb);
          oq[(size_t)l * os] += (i & 1) ? qb : qa;
        }
      }
      acc[i] = 0ull;
    }
#pragma unroll
    for (int i = 0; i < kMixSlots / 2; i++) accQ[i] = 0ull;
    qdirty = false;
  };

  const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = (nlisted + kMixChunk - 1) / kMixChunk;
  const unsigned *order = p.mix_order + (size_t)k * C + c0;
  // warp 0 requests chunk ci: lane i reads the i-th order entry and requests that channel's input row
  // and record (the diffuse half of a record only when the channel has diffuse gains)
  auto request = [&](int ci) {
    const int i0 = ci * kMixChunk, cnt = min(kMixChunk, nlisted - i0), st = ci & 1;
    const unsigned bar = bar0 + 8 * st;
    unsigned e = 0u, bytes = 0u;
    if (tid < cnt) {
      e = order[i0 + tid];
      sord[st][tid] = e;
      bytes = (unsigned)XW * 4u + ((e >> 30 & 1u) ? kMixRecWords * 4u : (kMixRecWords - kMixSlots * 2) * 4u);
    }
    unsigned total = bytes;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if (tid == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(total) : "memory");
    __syncwarp();
    if (tid < cnt) {
      const unsigned c = e & 0xffffffu;
      const float *src = p.xin + (size_t)c * p.xs + (t - tid);   // column (hist + t0) - hist
      const unsigned dst = (unsigned)__cvta_generic_to_shared(xt0 + (st * kMixChunk + tid) * XW);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "l"(src), "r"(XW * 4), "r"(bar) : "memory");
      const float *rsrc = p.mix_recs + ((size_t)k * C + c) * kMixRecWords;
      const unsigned rdst = (unsigned)__cvta_generic_to_shared(rc0 + (st * kMixChunk + tid) * kMixRecWords);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(rdst), "l"(rsrc), "r"(bytes - (unsigned)XW * 4u), "r"(bar) : "memory");
    }
  };
  if (tid < 32 && nchunks > 0) request(0);
  for (int ci = 0; ci < nchunks; ci++) {
    const int st = ci & 1;
    if (tid < 32 && ci + 1 < nchunks) request(ci + 1);   // its buffer was released by the barrier below
    mbar_wait(bar0 + 8 * st, (ci >> 1) & 1);
    const int cnt = min(kMixChunk, nlisted - ci * kMixChunk);
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
        const float4 *rec = reinterpret_cast<const float4 *>(recs + j * kMixRecWords);
        const unsigned e = sord[st][j];
        if (e >> 31) {   // a new loudspeaker set: close the run, take the slot map of this one
          flush();
          const uint4 meta = *reinterpret_cast<const uint4 *>(rec + 1);
          nact = meta.x & 0xffu; spk_lo = meta.y; spk_hi = meta.z;
        }
#pragma unroll
        for (int i = 0; i < kMixSlots / 2; i++) {
          if (i == kMixSlots / 4 && nact <= (unsigned)kMixSlots / 2) break;   // upper half of the slots unused
          const float4 g = rec[2 + i];
          float ga, gb;
          unpack2(fma2(rg2, pack2(g.z, g.w), pack2(g.x, g.y)), ga, gb);
          acc[2 * i] = fma2(splat2(ga), x2[u], acc[2 * i]);
          acc[2 * i + 1] = fma2(splat2(gb), x2[u], acc[2 * i + 1]);
        }
        if (e >> 30 & 1u) {
          const f32x2 x0 = splat2(xrow[j * XW]);
          qdirty = true;
#pragma unroll
          for (int i = 0; i < kMixSlots / 2; i++) {
            if (i == kMixSlots / 4 && nact <= (unsigned)kMixSlots / 2) break;
            const float4 q = rec[2 + kMixSlots / 2 + i];
            accQ[i] = fma2(fma2(rg2, pack2(q.z, q.w), pack2(q.x, q.y)), x0, accQ[i]);
          }
        }
      }
    }
    __syncthreads();   // every thread is done with buffer st before warp 0 requests chunk ci + 2 into it
  }
  flush();
}

// mix_dense_kernel: the channels whose gains are dense (more than kMixSlots non-zero loudspeakers:
// objects with extent), straight from the gain tables with all 3L accumulators in registers, added to
// the bus rows.  Every thread reads the same gain address (a broadcast from L1), the delayed reads
// come from global memory.  Periods without dense channels return at once.
// grid: periods * B / kMixTile; block kMixTile.
__global__ void __launch_bounds__(kMixTile) mix_dense_kernel(Layout p, int T)
{
  const int tid = threadIdx.x;
  const int t = p.k0 * p.B + blockIdx.x * kMixTile + tid;
  const int k = p.k0 + (blockIdx.x * kMixTile) / p.B;
  const int n = t - k * p.B;
  const int C = p.N + p.M, L = p.L;
  const int per = (C + p.S - 1) / p.S;
  int total = 0;
  for (int s = 0; s < p.S; s++) total += p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2 + 1];
  if (total == 0) return;
  const float invB = 1.0f / (float)p.B;
  const float rg = (float)(n + p.goff) * invB, rd = (float)(n + p.doff) * invB;
  f32x2 acc[kMaxSpeakers];
  float accQ[kMaxSpeakers];
#pragma unroll
  for (int l = 0; l < kMaxSpeakers; l++) { acc[l] = 0ull; accQ[l] = 0.f; }
  for (int s = 0; s < p.S; s++) {
    const int c0 = min(C, s * per);
    const int nlisted = p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2];
    const int ndense = p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2 + 1];
    for (int i = 0; i < ndense; i++) {
      const unsigned e = p.mix_order[(size_t)k * C + c0 + nlisted + i];
      const int c = (int)(e & 0xffffffu);
      const float4 dl = *reinterpret_cast<const float4 *>(p.mix_recs + ((size_t)k * C + c) * kMixRecWords);
      const float *xr = p.xin + (size_t)c * p.xs + p.hist + t;
      const f32x2 x2 = delayed_read2(xr, ramped_delay(dl.x, dl.y, rd), xr, ramped_delay(dl.z, dl.w, rd));
      const bool obj = c < p.N;
      const int ld = obj ? p.N : p.M;
      const float *g = (obj ? p.g_dir + (size_t)k * L * ld + c : p.g_ds + (size_t)k * L * ld + (c - p.N));
#pragma unroll
      for (int l = 0; l < kMaxSpeakers; l++)
        if (l < L) {
          const float g0 = __ldg(g + (size_t)l * ld), g1 = __ldg(g + (size_t)(L + l) * ld);
          acc[l] = fma2(splat2(fmaf(rg, g1 - g0, g0)), x2, acc[l]);
        }
      if (e >> 30 & 1u) {
        const float x0 = xr[0];
        const float *q = p.g_dif + (size_t)k * L * p.N + c;
#pragma unroll
        for (int l = 0; l < kMaxSpeakers; l++)
          if (l < L) {
            const float q0 = __ldg(q + (size_t)l * p.N), q1 = __ldg(q + (size_t)(L + l) * p.N);
            accQ[l] = fmaf(fmaf(rg, q1 - q0, q0), x0, accQ[l]);
          }
      }
    }
  }
  float *od = p.cdir + (size_t)p.B + t, *oq = p.qdif + (size_t)p.B + t;
#pragma unroll
  for (int l = 0; l < kMaxSpeakers; l++)
    if (l < L) {
      float vL, vR;
      unpack2(acc[l], vL, vR);
      od[(size_t)(2 * l) * p.cs] += vL;
      od[(size_t)(2 * l + 1) * p.cs] += vR;
      oq[(size_t)l * p.cs] += accQ[l];
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
