// sm_100a kernels of the BEAR render core.  FP32 throughout (parity bar: max-abs 1e-5 against
// the CPU oracle), no tensor cores: the path is HBM/L2-bound complex MACs plus shared-memory
// FFTs (DESIGN.md "Kernels").
//
// Signal flow per call of T periods (reference: visr_bear/src/dsp.cpp:117-205):
//   mix_prep_kernel   per (period, channel): delay ramps in samples, the set of loudspeakers with a non-zero
//                     gain, a 128-byte record of <= 6 (start, delta) gain slots; sort key = the set
//   mix_order_kernel  channels of a period sorted by set -> runs of channels that share their slots
//   mix_kernel        x -> per-ear Lagrange-3 delay -> ramped gain mix -> direct bus (48 rows)
//                     x -> ramped gain mix -> diffuse bus (24 rows); slot sums added per run   [a2,a3,a4,a7]
//   mix_dense_kernel  the same for channels with more than 6 non-zero loudspeakers (extent)
//   bus_sum_kernel    short launches split their channels over several CTAs per tile: partial rows summed
//   hoa_mix_kernel    hoa in -> ramped matrix -> ambisonic bus (16 ch)                 [a10]
//   decorrelate_kernel  diffuse bus -> overlap-save FFT conv with the decorrelators    [a5]
//   bus_fdl_kernel    direct bus + static-delayed decorrelator outputs -> 48 ch bus -> rFFT(2B)
//                     -> frequency-domain delay line; HOA bus (fixed delay commuted in front of
//                     the decode filters) -> rFFT(2B) -> FDL                           [a6,a8,a12]
//   view_group_kernel periods of the launch sorted by BRIR view -> work items of <= 5 periods
//   cmac_kernel       Y[e] = sum_{l,p} FDL[2l+e][k-p] * H[view][2l+e][p] (+ HOA decode terms),
//                     for the new view and, in a view-change block, the old one; one filter load
//                     serves all periods of a work item                               [a9,a11]
//   ifft_out_kernel   sum partial spectra -> irFFT -> overlap-save discard -> view crossfade -> out [a9,a13]
//   carry_kernel      moves histories / table state / the last P delay-line spectra to the front
//   unpack_kernel     single-period host call: scatters the packed upload (audio + tables)
//   frames_kernel     interleaved file frames (float / 16-, 24-, 32-bit PCM) -> planar input rows
//   control_kernel    optional: the Control graph's table producers on the device      [a15]
//   validate_views_kernel  device-resident tables: view indices checked before they index the filters
// One scene on several GPUs (bear_cuda_comm_connect): peer_sync_kernel (counters in the peers' memory),
//   bus_gather_kernel / bus_gather_thin_kernel (this rank's loudspeakers' rows summed over all ranks,
//   peers read over NVLink with ordinary loads), out_sum_kernel (stereo partial mixes on the receiver).
// Every bulk copy (cp.async.bulk) in this file reads this GPU's own memory; there is no multicast.
#include "kernels.cuh"

#include <type_traits>

namespace bearcuda {

// ----------------------------------------------------------------------------------------------
// small complex helpers
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
  return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_conj(float2 a, float2 w)  // a * conj(w)
{
  return make_float2(fmaf(a.x, w.x, a.y * w.y), fmaf(a.y, w.x, -a.x * w.y));
}
// table of exp(-i pi q / M), q in [0, M); q in [M, 2M) by symmetry.  (The table itself has 2M
// entries: [M, 2M) holds the per-pass twiddles of the radix-8 passes, see fft_shared.)
__device__ __forceinline__ float2 ldtw(const float2 *__restrict__ tw, int q, int M)
{
  if (q < M) return __ldg(tw + q);
  float2 w = __ldg(tw + q - M);
  return make_float2(-w.x, -w.y);
}

// ----------------------------------------------------------------------------------------------
// Complex FFT of length M = 2^logM in shared memory: Stockham autosort with radix-8 passes held in
// registers (then one radix-4 or radix-2 pass if log2 M is not a multiple of 3).  Natural order in
// and out, unnormalised.  Both buffers hold M + M/8 float2: element i lives at padi(i) so that the
// stride-8 scatter of a radix-8 pass is bank-conflict free.  `a` holds the input, `b` is scratch;
// returns the buffer that holds the result.  All threads of the CTA must call it.
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ int padi(int i) { return i + (i >> 3); }
__host__ __device__ constexpr int fft_smem_floats2(int M) { return 2 * (M + (M >> 3)); }

template <bool INV>
__device__ __forceinline__ void dft4(float2 a0, float2 a1, float2 a2, float2 a3, float2 &x0, float2 &x1, float2 &x2,
                                     float2 &x3)
{
  const float2 s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), t = csub(a1, a3);
  const float2 r = INV ? make_float2(-t.y, t.x) : make_float2(t.y, -t.x);   // -+i (a1 - a3)
  x0 = cadd(s02, s13); x2 = csub(s02, s13);
  x1 = cadd(d02, r);   x3 = csub(d02, r);
}
// v[r] <- sum_n v[n] exp(-+2 pi i n r / 8)
template <bool INV>
__device__ __forceinline__ void dft8(float2 (&v)[8])
{
  const float s = 0.70710678118654752440f;
  const float2 a0 = cadd(v[0], v[4]), a1 = cadd(v[1], v[5]), a2 = cadd(v[2], v[6]), a3 = cadd(v[3], v[7]);
  const float2 b0 = csub(v[0], v[4]);
  float2 t = csub(v[1], v[5]);
  const float2 b1 = INV ? make_float2((t.x - t.y) * s, (t.x + t.y) * s) : make_float2((t.x + t.y) * s, (t.y - t.x) * s);
  t = csub(v[2], v[6]);
  const float2 b2 = INV ? make_float2(-t.y, t.x) : make_float2(t.y, -t.x);
  t = csub(v[3], v[7]);
  const float2 b3 = INV ? make_float2((-t.x - t.y) * s, (t.x - t.y) * s) : make_float2((t.y - t.x) * s, (-t.x - t.y) * s);
  dft4<INV>(a0, a1, a2, a3, v[0], v[2], v[4], v[6]);
  dft4<INV>(b0, b1, b2, b3, v[1], v[3], v[5], v[7]);
}

template <bool INV>
__device__ __forceinline__ float2 *fft_shared(float2 *a, float2 *b, int M, int logM, const float2 *__restrict__ tw);

template <bool INV>
__device__ __forceinline__ float2 twid(float2 v, float2 w) { return INV ? cmul_conj(v, w) : cmul(v, w); }

// (tid, nthr): the calling thread's index within, and the size of, the thread group that works on this
// transform; several groups of one CTA may transform different buffers in lockstep.
template <bool INV>
__device__ __forceinline__ float2 *fft_shared(float2 *a, float2 *b, int M, int logM, const float2 *__restrict__ tw,
                                              int tid, int nthr)
{
  int ls = 0;
  for (; ls + 3 <= logM; ls += 3) {
    const int Ns = 1 << ls, Q = M >> 3;
    // exp(-2 pi i r k / (8 Ns)), r = 1..7, from the per-pass table behind the first M entries:
    // consecutive threads read consecutive entries (k is the low bits of j).  The twiddles of the
    // thread's first butterfly are fetched ahead of the barrier, they do not depend on the data.
    float2 w0[7];
    if (ls > 0 && tid < Q) {
      const float2 *w = tw + M + (Ns - 8) + (tid & (Ns - 1));
#pragma unroll
      for (int r = 1; r < 8; r++) w0[r - 1] = __ldg(w + (r - 1) * Ns);
    }
    __syncthreads();
    for (int j = tid; j < Q; j += nthr) {
      const int k = j & (Ns - 1);
      float2 v[8];
#pragma unroll
      for (int r = 0; r < 8; r++) v[r] = a[padi(j + r * Q)];
      if (ls > 0) {
        if (j == tid) {
#pragma unroll
          for (int r = 1; r < 8; r++) v[r] = twid<INV>(v[r], w0[r - 1]);
        } else {
          const float2 *w = tw + M + (Ns - 8) + k;
#pragma unroll
          for (int r = 1; r < 8; r++) v[r] = twid<INV>(v[r], __ldg(w + (r - 1) * Ns));
        }
      }
      dft8<INV>(v);
      const int j0 = ((j - k) << 3) + k;
#pragma unroll
      for (int r = 0; r < 8; r++) b[padi(j0 + r * Ns)] = v[r];
    }
    float2 *t = a; a = b; b = t;
  }
  if (logM - ls == 2) {
    const int Ns = 1 << ls, Q = M >> 2;
    __syncthreads();
    for (int j = tid; j < Q; j += nthr) {
      const int k = j & (Ns - 1);
      float2 v0 = a[padi(j)], v1 = a[padi(j + Q)], v2 = a[padi(j + 2 * Q)], v3 = a[padi(j + 3 * Q)];
      if (ls > 0) {
        const int q = k << (logM - 1 - ls);
        v1 = twid<INV>(v1, ldtw(tw, q, M)); v2 = twid<INV>(v2, ldtw(tw, 2 * q, M)); v3 = twid<INV>(v3, ldtw(tw, 3 * q, M));
      }
      float2 x0, x1, x2, x3;
      dft4<INV>(v0, v1, v2, v3, x0, x1, x2, x3);
      const int j0 = ((j - k) << 2) + k;
      b[padi(j0)] = x0; b[padi(j0 + Ns)] = x1; b[padi(j0 + 2 * Ns)] = x2; b[padi(j0 + 3 * Ns)] = x3;
    }
    float2 *t = a; a = b; b = t;
  } else if (logM - ls == 1) {
    const int Ns = 1 << ls, Hh = M >> 1;
    __syncthreads();
    for (int j = tid; j < Hh; j += nthr) {
      const int k = j & (Ns - 1);
      const float2 v0 = a[padi(j)];
      const float2 v1 = twid<INV>(a[padi(j + Hh)], ldtw(tw, k << (logM - ls), M));
      const int j0 = ((j - k) << 1) + k;
      b[padi(j0)] = cadd(v0, v1);
      b[padi(j0 + Ns)] = csub(v0, v1);
    }
    float2 *t = a; a = b; b = t;
  }
  __syncthreads();
  return a;
}

template <bool INV>
__device__ __forceinline__ float2 *fft_shared(float2 *a, float2 *b, int M, int logM, const float2 *__restrict__ tw)
{
  return fft_shared<INV>(a, b, M, logM, tw, threadIdx.x, blockDim.x);
}

// First radix-8 pass of fft_shared_c with the input in registers: thread tid < M/8 holds
// v[r] = x[tid + r * M/8].  The butterfly has no twiddles (Ns = 1); its output goes to `out` in the
// layout the second pass reads, so the input never makes a round trip through shared memory.
template <bool INV, int LOGM>
__device__ __forceinline__ void fft_first_pass_reg(float2 (&v)[8], float2 *out, int tid)
{
  if (tid < (1 << LOGM) / 8) {
    dft8<INV>(v);
    float2 *bp = out + 9 * tid;   // padi(8 tid + r) = 9 tid + r
#pragma unroll
    for (int r = 0; r < 8; r++) bp[r] = v[r];
  }
}

// The same transform with length and group size known at compile time (the per-period kernels are
// instantiated per period size): the passes unroll, and because M/8 and Ns >= 8 are multiples of 8
// the padded index of element j + r*stride is padi(j) + r*(stride + stride/8), a constant offset
// from one base address per butterfly instead of eight index computations.
// FIRST = 1: the first radix-8 pass has been done by fft_first_pass_reg (its output is in `a`).
// UPPER_OUT: the last pass stores elements M/2 .. M-1 of the result to upper_out[0 .. M/2) (global
// memory, consecutive threads consecutive elements) instead of shared memory -- the overlap-save
// output of an inverse transform, whose lower half is discarded.
template <bool INV, int LOGM, int NTHR, int FIRST = 0, bool UPPER_OUT = false>
__device__ __forceinline__ float2 *fft_shared_c(float2 *a, float2 *b, const float2 *__restrict__ tw, int tid,
                                                float2 *__restrict__ upper_out = nullptr)
{
  constexpr int M = 1 << LOGM, Q = M >> 3, QP = Q + (Q >> 3);
  constexpr int NP = LOGM / 3, REM = LOGM % 3;
  static_assert(LOGM >= 6, "constant-offset addressing needs M/8 to be a multiple of 8");
#pragma unroll
  for (int pass = FIRST; pass < NP; pass++) {
    const int Ns = 1 << (3 * pass), NsP = Ns + (Ns >> 3);
    float2 w0[7];
    if (pass > 0 && tid < Q) {
      const float2 *w = tw + M + (Ns - 8) + (tid & (Ns - 1));
#pragma unroll
      for (int r = 1; r < 8; r++) w0[r - 1] = __ldg(w + (r - 1) * Ns);
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < (Q + NTHR - 1) / NTHR; it++) {
      const int j = tid + it * NTHR;
      if (Q % NTHR == 0 || j < Q) {
        const int k = j & (Ns - 1);
        float2 v[8];
        const float2 *ap = a + padi(j);
#pragma unroll
        for (int r = 0; r < 8; r++) v[r] = ap[r * QP];
        if (pass > 0) {
          if (it == 0) {
#pragma unroll
            for (int r = 1; r < 8; r++) v[r] = twid<INV>(v[r], w0[r - 1]);
          } else {
            const float2 *w = tw + M + (Ns - 8) + k;
#pragma unroll
            for (int r = 1; r < 8; r++) v[r] = twid<INV>(v[r], __ldg(w + (r - 1) * Ns));
          }
        }
        dft8<INV>(v);
        float2 *bp = b + padi(((j - k) << 3) + k);
#pragma unroll
        for (int r = 0; r < 8; r++) {
          if (UPPER_OUT && REM == 0 && pass == NP - 1) {   // last pass: Ns = M/8, element j + r M/8
            if (r >= 4) upper_out[j + (r - 4) * Q] = v[r];
          } else {
            bp[pass == 0 ? r : r * NsP] = v[r];            // pass 0: padi(8j + r) = 9j + r
          }
        }
      }
    }
    float2 *t = a; a = b; b = t;
  }
  constexpr int ls = 3 * NP, Ns = 1 << ls;
  if constexpr (REM == 2) {
    constexpr int Q4 = M >> 2, Q4P = Q4 + (Q4 >> 3), NsP = Ns + (Ns >> 3);
    __syncthreads();
#pragma unroll
    for (int it = 0; it < (Q4 + NTHR - 1) / NTHR; it++) {
      const int j = tid + it * NTHR;
      if (Q4 % NTHR == 0 || j < Q4) {
        const int k = j & (Ns - 1);
        const float2 *ap = a + padi(j);
        float2 v0 = ap[0], v1 = ap[Q4P], v2 = ap[2 * Q4P], v3 = ap[3 * Q4P];
        const int q = k << (LOGM - 1 - ls);
        v1 = twid<INV>(v1, ldtw(tw, q, M)); v2 = twid<INV>(v2, ldtw(tw, 2 * q, M)); v3 = twid<INV>(v3, ldtw(tw, 3 * q, M));
        float2 x0, x1, x2, x3;
        dft4<INV>(v0, v1, v2, v3, x0, x1, x2, x3);
        float2 *bp = b + padi(((j - k) << 2) + k);
        if (UPPER_OUT) {                                   // last pass: Ns = M/4, element j + r M/4
          upper_out[j] = x2; upper_out[j + Q4] = x3;
        } else {
          bp[0] = x0; bp[NsP] = x1; bp[2 * NsP] = x2; bp[3 * NsP] = x3;
        }
      }
    }
    float2 *t = a; a = b; b = t;
  } else if constexpr (REM == 1) {
    constexpr int Hh = M >> 1, HhP = Hh + (Hh >> 3), NsP = Ns + (Ns >> 3);
    __syncthreads();
#pragma unroll
    for (int it = 0; it < (Hh + NTHR - 1) / NTHR; it++) {
      const int j = tid + it * NTHR;
      if (Hh % NTHR == 0 || j < Hh) {
        const int k = j & (Ns - 1);
        const float2 *ap = a + padi(j);
        const float2 v0 = ap[0];
        const float2 v1 = twid<INV>(ap[HhP], ldtw(tw, k << (LOGM - ls), M));
        float2 *bp = b + padi(((j - k) << 1) + k);
        if (UPPER_OUT) {                                   // last pass: Ns = M/2, element j + r M/2
          upper_out[j] = csub(v0, v1);
        } else {
          bp[0] = cadd(v0, v1);
          bp[NsP] = csub(v0, v1);
        }
      }
    }
    float2 *t = a; a = b; b = t;
  }
  __syncthreads();
  return a;
}

// real-FFT split: Z = FFT_M(x[2n] + i x[2n+1]);  X[k], X[M-k] of the length-2M real transform
__device__ __forceinline__ void rfft_post(float2 zk, float2 zm, float2 w, float2 &xk, float2 &xm)
{
  float2 bc = make_float2(zm.x, -zm.y);
  float2 e = make_float2(0.5f * (zk.x + bc.x), 0.5f * (zk.y + bc.y));
  float2 d = csub(zk, bc);
  float2 o = make_float2(0.5f * d.y, -0.5f * d.x);
  float2 wo = cmul(w, o);
  xk = cadd(e, wo);
  float2 t = csub(e, wo);
  xm = make_float2(t.x, -t.y);
}
// inverse of the above up to a factor 2 (the unnormalised inverse FFT_M then yields 2M x[n])
__device__ __forceinline__ void irfft_pre(float2 xk, float2 xm, float2 w, float2 &zk, float2 &zm)
{
  float2 b = make_float2(xm.x, -xm.y);
  float2 ev = cadd(xk, b);
  float2 od = cmul_conj(csub(xk, b), w);
  zk = make_float2(ev.x - od.y, ev.y + od.x);
  zm = make_float2(ev.x + od.y, od.x - ev.y);
}

// Two FP32 lanes in one 64-bit register: fma.rn.f32x2 (SASS FFMA2) does two IEEE fused multiply-adds
// per issued instruction -- bit-identical to two fmaf() -- which halves the issue slots of the
// accumulation (a scalar second operand is broadcast by the hardware).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
  f32x2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi)
{
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// acc += a * b with the accumulator tied to one register (no copy for the compiler to place)
__device__ __forceinline__ void fma2_acc(f32x2 &acc, f32x2 a, f32x2 b)
{
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
  f32x2 d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 splat2(float v) { return pack2(v, v); }

// Lagrange order-3 coefficients for position s in [1,2) relative to the first of the nodes 0..3
__device__ __forceinline__ void lagrange3(float s, float c[4])
{
  float s1 = s - 1.0f, s2 = s - 2.0f, s3 = s - 3.0f;
  c[0] = -(s1 * s2 * s3) * (1.0f / 6.0f);
  c[1] = (s * s2 * s3) * 0.5f;
  c[2] = -(s * s1 * s3) * 0.5f;
  c[3] = (s * s1 * s2) * (1.0f / 6.0f);
}
// Delay at ramp position r, rounded product then rounded sum (no FMA): a delay near 100 samples
// has an ulp of 7.6e-6 samples, and on white noise one ulp of delay is ~1e-6 of output, so the
// evaluation order is kept identical to the CPU restatement.
__device__ __forceinline__ float ramped_delay(float d0, float dd, float r) { return __fadd_rn(d0, __fmul_rn(dd, r)); }
// y = sum_j c_j(1.5 + mu) x[-i - j] with i = floor(d + .5), mu = d - i: realised delay d + 1.5
__device__ __forceinline__ float delayed_read(const float *__restrict__ x, float d)
{
  float fi = floorf(__fadd_rn(d, 0.5f));
  float c[4];
  lagrange3(__fadd_rn(1.5f, __fsub_rn(d, fi)), c);
  const float *q = x - (int)fi;
  float acc = c[0] * q[0];
  acc = fmaf(c[1], q[-1], acc);
  acc = fmaf(c[2], q[-2], acc);
  acc = fmaf(c[3], q[-3], acc);
  return acc;
}
// Two delayed reads at once, (delayed_read(xa, da), delayed_read(xb, db)), with the coefficient
// arithmetic in packed form: the same IEEE operations in the same order as the scalar version, two
// lanes per instruction.  (ptxas contracts a packed multiply feeding a packed add into FFMA2 even
// for mul.rn/add.rn, so there is deliberately no such pair here; ramped_delay stays scalar.)
__device__ __forceinline__ f32x2 delayed_read2(const float *__restrict__ xa, float da, const float *__restrict__ xb, float db)
{
  const f32x2 d = pack2(da, db);
  float ha, hb;
  unpack2(add2(d, splat2(0.5f)), ha, hb);
  // floor(h) for 0 <= h < 2^23 without the conversion unit: h + 2^23 rounded towards minus infinity
  // is exactly 2^23 + floor(h); its low mantissa bits are the integer, subtracting 2^23 back the float
  const float ua = __fadd_rd(ha, 8388608.0f), ub = __fadd_rd(hb, 8388608.0f);
  const float fa = ua - 8388608.0f, fb = ub - 8388608.0f;
  const int ia = __float_as_int(ua) & 0x7fffff, ib = __float_as_int(ub) & 0x7fffff;
  // position among the four nodes: s = 1.5 + mu with mu = d - floor(d + .5) in [-.5, .5).  With m2 = mu^2,
  // s (s-3) = m2 - 2.25 and (s-1)(s-2) = m2 - 0.25, so
  //   c0 = -(s-1)(s-2)(s-3)/6 = (m2 - .25) (.25 - mu/6)      c1 =  s(s-2)(s-3)/2 = (m2 - 2.25) (mu/2 - .25)
  //   c3 =  s(s-1)(s-2)/6     = (m2 - .25) (.25 + mu/6)      c2 = -s(s-1)(s-3)/2 = (m2 - 2.25) (-mu/2 - .25)
  // : 12 packed operations for the eight coefficients of both ears
  const f32x2 mu = sub2(d, pack2(fa, fb));
  const f32x2 m2 = mul2(mu, mu);
  const f32x2 a = add2(m2, splat2(-2.25f)), b = add2(m2, splat2(-0.25f));
  const f32x2 c0 = mul2(b, fma2(mu, splat2(-1.0f / 6.0f), splat2(0.25f)));
  const f32x2 c1 = mul2(a, fma2(mu, splat2(0.5f), splat2(-0.25f)));
  const f32x2 c2 = mul2(a, fma2(mu, splat2(-0.5f), splat2(-0.25f)));
  const f32x2 c3 = mul2(b, fma2(mu, splat2(1.0f / 6.0f), splat2(0.25f)));
  const float *qa = xa - ia, *qb = xb - ib;
  f32x2 acc = mul2(c0, pack2(qa[0], qb[0]));
  acc = fma2(c1, pack2(qa[-1], qb[-1]), acc);
  acc = fma2(c2, pack2(qa[-2], qb[-2]), acc);
  acc = fma2(c3, pack2(qa[-3], qb[-3]), acc);
  return acc;
}

// ----------------------------------------------------------------------------------------------
// Filter table -> partitioned spectra (one-off, at renderer creation).
// grid: n_filters * P; block: B/4.  out[(f*P + p)*B + bin], scaled.
// ----------------------------------------------------------------------------------------------
__global__ void filter_fft_kernel(const float *__restrict__ taps, int taps_len, int B, int logB, int P,
                                  const float2 *__restrict__ tw, float2 *__restrict__ out, float scale)
{
  extern __shared__ float2 sm[];
  float2 *a = sm, *b = sm + B + (B >> 3);
  const int f = blockIdx.x / P, p = blockIdx.x % P;
  const float *src = taps + (size_t)f * taps_len;
  for (int n = threadIdx.x; n < B; n += blockDim.x) {
    int t0 = p * B + 2 * n;
    float v0 = (2 * n < B && t0 < taps_len) ? src[t0] : 0.0f;
    float v1 = (2 * n + 1 < B && t0 + 1 < taps_len) ? src[t0 + 1] : 0.0f;
    a[padi(n)] = make_float2(v0 * scale, v1 * scale);
  }
  float2 *z = fft_shared<false>(a, b, B, logB, tw);
  float2 *dst = out + (size_t)blockIdx.x * B;
  for (int k = threadIdx.x; k <= B / 2; k += blockDim.x) {
    if (k == 0) {
      dst[0] = make_float2(z[0].x + z[0].y, z[0].x - z[0].y);
    } else {
      float2 xk, xm;
      rfft_post(z[padi(k)], z[padi(B - k)], __ldg(tw + k), xk, xm);
      dst[k] = xk;
      dst[B - k] = xm;
    }
  }
}

void launch_filter_fft(const float *taps, int n_filters, int taps_len, int B, int logB, int P, const float2 *tw,
                       float2 *out, float scale, cudaStream_t st)
{
  int threads = B / 8 < 32 ? 32 : B / 8;
  filter_fft_kernel<<<n_filters * P, threads, fft_smem_floats2(B) * sizeof(float2), st>>>(taps, taps_len, B, logB, P, tw, out, scale);
}

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
// Record (kMixRecWords = 8 + 4 kMixSlots words):
//   [0..3]   per-ear delay ramp in samples: d0L, d0R, ddL, ddR (validated, then clamped to the history)
//   [4]      number of slots | has_diffuse << 8        [5], [6] loudspeaker of slots 0..3, 4.. (bytes; 255 = unused)
//   [8..]    direct gains, slot pairs:  (g0[2i], g0[2i+1], dg[2i], dg[2i+1]),  i = 0 .. kMixSlots/2 - 1
//   then     diffuse gains, same layout (objects only)
// Key: bit 31 dense | loudspeaker mask << 1 | has_diffuse.
// grid: (ceil(N/128) + ceil(M/128), periods); block 128; one thread per channel (coalesced table reads),
// a CTA holds channels of one kind.  The kernel is bound by memory transactions, not by bytes, so the
// scattered accesses go through shared memory: while a thread walks its 4 L table entries (high
// loudspeakers first) it pushes the (start, delta) values of the non-zero ones onto its own stack in
// shared memory -- gathering them from global memory afterwards cost up to 24 lines per load -- and the
// 128-byte records are exchanged through shared memory so that the CTA writes them as contiguous lines.
constexpr int kPrepThreads = 128;
__global__ void __launch_bounds__(kPrepThreads) mix_prep_kernel(Layout p, int T)
{
  __shared__ float stack[4 * kMixSlots][kPrepThreads];            // [4 s + {g0, g1, q0, q1}][thread]
  __shared__ float4 recs[kPrepThreads * (kMixRecWords / 4)];
  const int tid = threadIdx.x, k = p.k0 + blockIdx.y;
  const int C = p.N + p.M, L = p.L;
  const int nbo = (p.N + kPrepThreads - 1) / kPrepThreads;
  const bool obj = (int)blockIdx.x < nbo;
  const int cb = (obj ? blockIdx.x : blockIdx.x - nbo) * kPrepThreads, ld = obj ? p.N : p.M;
  const int cnt = min(kPrepThreads, ld - cb);
  const bool on = tid < cnt;
  const int ci = cb + tid, c = obj ? ci : p.N + ci;
  constexpr int Q = kMixRecWords / 4;
  float4 rec[Q];
#pragma unroll
  for (int i = 0; i < Q; i++) rec[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (on) {
    const float *gd = (obj ? p.g_dir : p.g_ds) + (size_t)k * L * ld + ci;   // [slot k: L rows][slot k+1: L rows]
    const float *gq = p.g_dif + (size_t)k * L * ld + ci;
    unsigned mask = 0u, any_q = 0u;
    int ns = 0;
    unsigned spk[2] = {0u, 0u};
    for (int l0 = ((L - 1) & ~7); l0 >= 0; l0 -= 8) {   // eight loudspeakers' table entries are loaded before any is tested
      float va[8], vb[8], qa[8], qb[8];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int l = min(l0 + i, L - 1);
        va[i] = gd[(size_t)l * ld]; vb[i] = gd[(size_t)(L + l) * ld];
        qa[i] = obj ? gq[(size_t)l * ld] : 0.f; qb[i] = obj ? gq[(size_t)(L + l) * ld] : 0.f;
      }
#pragma unroll
      for (int i = 7; i >= 0; i--) {
        if (l0 + i < L) {
          const bool nzq = qa[i] != 0.f || qb[i] != 0.f;
          if (va[i] != 0.f || vb[i] != 0.f || nzq) {
            mask |= 1u << (l0 + i);
            if (ns < kMixSlots) {   // slots in DESCENDING loudspeaker order (see below)
              stack[4 * ns + 0][tid] = va[i]; stack[4 * ns + 1][tid] = vb[i];
              stack[4 * ns + 2][tid] = qa[i]; stack[4 * ns + 3][tid] = qb[i];
              spk[ns >> 2] |= (unsigned)(l0 + i) << (8 * (ns & 3));
            }
            ns++;
          }
          if (nzq) any_q = 1u;
        }
      }
    }
    // delays of periods k-1 (slot k) and k (slot k+1), seconds -> samples
    const float *ds = (obj ? p.d_dir : p.d_ds) + (size_t)k * 2 * ld + 2 * ci;
    const float dmax = (float)(p.hist - 4);
    {
      float a0 = ds[0] * p.fs, a1 = ds[2 * ld] * p.fs, b0 = ds[1] * p.fs, b1 = ds[2 * ld + 1] * p.fs;
      // a table delay outside [0, history - 4] cannot be rendered (the reference allows up to 1 s,
      // per_ear_delay.cpp:25): flagged, reported by the call, and clamped so that no read leaves the tile
      if (!(a1 >= 0.f && a1 <= dmax) || !(b1 >= 0.f && b1 <= dmax)) atomicOr(p.err_flag, kErrDelayRange);
      a0 = fminf(fmaxf(a0, 0.f), dmax); a1 = fminf(fmaxf(a1, 0.f), dmax);
      b0 = fminf(fmaxf(b0, 0.f), dmax); b1 = fminf(fmaxf(b1, 0.f), dmax);
      rec[0] = make_float4(a0, b0, a1 - a0, b1 - b0);
    }
    unsigned *keys = p.mix_keys + (size_t)k * C;
    if (ns > kMixSlots) {
      keys[c] = 0x80000000u | any_q;
      rec[1] = make_float4(__uint_as_float(255u | any_q << 8), 0.f, 0.f, 0.f);
    } else {
      keys[c] = mask << 1 | any_q;
      // Slots in DESCENDING loudspeaker order: the channels are sorted by the mask as a number, so
      // neighbours in the sorted order share their highest loudspeakers, i.e. their first slots, which
      // mix_kernel can keep accumulating across a run boundary.
      float g0[kMixSlots], dg[kMixSlots], q0[kMixSlots], dq[kMixSlots];
#pragma unroll
      for (int s = 0; s < kMixSlots; s++) {
        g0[s] = dg[s] = q0[s] = dq[s] = 0.f;
        if (s < ns) {
          const float a = stack[4 * s + 0][tid], b = stack[4 * s + 1][tid];
          g0[s] = a; dg[s] = b - a;
          const float qa = stack[4 * s + 2][tid], qb = stack[4 * s + 3][tid];
          q0[s] = qa; dq[s] = qb - qa;
        } else {
          spk[s >> 2] |= 255u << (8 * (s & 3));   // 255: slot not needed
        }
      }
      rec[1] = make_float4(__uint_as_float((unsigned)ns | any_q << 8), __uint_as_float(spk[0]), __uint_as_float(spk[1]), 0.f);
#pragma unroll
      for (int i = 0; i < kMixSlots / 2; i++) {
        rec[2 + i] = make_float4(g0[2 * i], g0[2 * i + 1], dg[2 * i], dg[2 * i + 1]);
        rec[2 + kMixSlots / 2 + i] = make_float4(q0[2 * i], q0[2 * i + 1], dq[2 * i], dq[2 * i + 1]);
      }
    }
  }
  // 16-byte piece i of thread t's record sits at piece (i + t) mod 8 of its 128 bytes: a quarter warp's
  // stores then hit eight different bank groups
  static_assert(Q == 8, "record swizzle assumes 8 pieces of 16 bytes");
#pragma unroll
  for (int i = 0; i < Q; i++) recs[tid * Q + ((i + tid) & (Q - 1))] = rec[i];
  __syncthreads();
  float4 *dst = reinterpret_cast<float4 *>(p.mix_recs + ((size_t)k * C + (obj ? cb : p.N + cb)) * kMixRecWords);
#pragma unroll
  for (int it = 0; it < Q; it++) {
    const int idx = it * kPrepThreads + tid, t = idx / Q, ph = idx % Q;
    if (t < cnt) dst[t * Q + ((ph - t) & (Q - 1))] = recs[idx];
  }
}

// mix_order_kernel: the channels of one (period, object split) sorted by key (bitonic sort in shared
// memory), written as order[k][c0 + i] = channel | (run length - 1) << 17 | has_diffuse << 30 |
// run_start << 31, where run_start marks the first channel of a run with a new loudspeaker set (only
// that entry carries the length); dense channels sort to the end.
// counts[k][split] = (listed channels, dense channels).
// grid: (periods, S); block 256; dynamic shared memory: 8 bytes x the next power of two >= channels per split.
__global__ void __launch_bounds__(256) mix_order_kernel(Layout p, int T)
{
  extern __shared__ __align__(16) unsigned sk[];   // [n2] sort words, then [n2] the keys by local channel
  const int k = p.k0 + blockIdx.x, s = blockIdx.y;
  const int C = p.N + p.M;
  const int per = (C + p.S - 1) / p.S;
  const int c0 = min(C, s * per), c1 = min(C, c0 + per), len = c1 - c0;
  int n2 = 1, bits = 0;
  while (n2 < len) { n2 <<= 1; bits++; }
  unsigned *kraw = sk + n2;
  const unsigned *keys = p.mix_keys + (size_t)k * C;
  // One 32-bit word per channel: (sort value of its loudspeaker set) << bits | local channel.  Up to 256
  // channels the 24-bit set itself fits (dense channels count as the full set and sort last), so the order
  // is the numeric order of the sets, which the slot-keeping form of mix_kernel relies on; beyond that
  // a (32 - bits)-bit hash of the set stands in -- all that matters there is that equal sets end up
  // next to each other (two sets with one hash merely split each other's runs; runs are detected on
  // the real keys below).  The diffuse flag is not part of the order: it does not separate runs.
  const bool exact = 24 + bits <= 32;
  const unsigned top = 0xffffffffu >> bits;        // sort value of the dense channels
  for (int i = threadIdx.x; i < n2; i += 256) {
    unsigned w = 0xffffffffu, key = 0xffffffffu;
    if (i < len) {
      key = keys[c0 + i];
      const unsigned set = key >> 1 & 0xffffffu;
      unsigned h = exact ? set : min((set * 2654435761u) >> bits, top - 1u);
      if (key >> 31) h = exact ? 0xffffffu : top;
      w = h << bits | (unsigned)i;
    }
    sk[i] = w;
    kraw[i] = key;
  }
  __syncthreads();
  // Bitonic network, two consecutive strides (2h, h) per pass: a thread holds the four elements
  // b + {0, h, 2h, 3h} in registers for both compare-exchange steps, which halves the trips through
  // shared memory and the barriers (30 instead of 55 passes for 1024 channels).
  auto cswap = [](unsigned &a, unsigned &b, bool up) {
    const unsigned lo = min(a, b), hi = max(a, b);
    a = up ? lo : hi; b = up ? hi : lo;
  };
  for (int size = 2; size <= n2; size <<= 1) {
    int stride = size >> 1;
    for (; stride >= 2; stride >>= 2) {
      const int h = stride >> 1, lh = 31 - __clz(h);
      for (int g = threadIdx.x; g < n2 / 4; g += 256) {
        const int b = ((g >> lh) << (lh + 2)) + (g & (h - 1));
        const bool up = (b & size) == 0;
        unsigned e0 = sk[b], e1 = sk[b + h], e2 = sk[b + 2 * h], e3 = sk[b + 3 * h];
        cswap(e0, e2, up); cswap(e1, e3, up);
        cswap(e0, e1, up); cswap(e2, e3, up);
        sk[b] = e0; sk[b + h] = e1; sk[b + 2 * h] = e2; sk[b + 3 * h] = e3;
      }
      __syncthreads();
    }
    if (stride == 1) {   // an odd number of strides: the last one on its own
      for (int i = threadIdx.x; i < n2 / 2; i += 256) {
        const bool up = (2 * i & size) == 0;
        uint2 v = reinterpret_cast<uint2 *>(sk)[i];
        cswap(v.x, v.y, up);
        reinterpret_cast<uint2 *>(sk)[i] = v;
      }
      __syncthreads();
    }
  }
  const unsigned lmask = (unsigned)n2 - 1u;
  unsigned *order = p.mix_order + (size_t)k * C + c0;
  int ndense = 0;
  for (int i0 = 0; i0 < len; i0 += 256) {   // (uniform trip count: the ballot below needs whole warps)
    const int i = i0 + threadIdx.x;
    const bool valid = i < len;
    unsigned key = 0u, c = 0u, start = 0u;
    if (valid) {
      const unsigned li = sk[i] & lmask;
      key = kraw[li]; c = (unsigned)c0 + li;
      const unsigned prev = i ? kraw[sk[i - 1] & lmask] : ~0u;
      start = (i == 0 || (key >> 1) != (prev >> 1)) ? 1u : 0u;
    }
    // the first channel of a run carries its length: up to the next run start among the warp's 32
    // consecutive positions, else to the end of those and on for as long as the set stays the same
    const unsigned starts = __ballot_sync(0xffffffffu, start != 0u);
    unsigned run = 0u;
    if (start) {
      const unsigned lane = threadIdx.x & 31u, rest = lane == 31u ? 0u : starts >> (lane + 1u);
      int j;
      if (rest) {
        j = i + __ffs(rest);
      } else {
        j = (i | 31) + 1;
        while (j < len && kraw[sk[j] & lmask] >> 1 == key >> 1) j++;
      }
      run = (unsigned)(min(j, len) - i - 1);
    }
    if (valid) order[i] = c | run << 17 | (key & 1u) << 30 | start << 31;
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
// its tile of the 3L rows, then walks its channels in sorted order in chunks of kMixChunk through a
// ring of kMixStages shared-memory stages: warp 0 reads a chunk's order entries and requests, per
// channel, the input row (history + tile) and the record by bulk copies (TMA, 1-D) that complete on
// the stage's "full" mbarrier, kMixStages - 1 chunks ahead of the arithmetic; a warp that is done with a stage
// arrives on its "empty" mbarrier, which is all warp 0 waits for before it refills the stage -- there
// is no CTA-wide barrier in the loop, the warps drift apart by up to a chunk.  Per channel: two
// Lagrange-3 reads at the ramped per-ear delays (the next channel's are computed before this channel's
// sums, for instruction-level parallelism), then per slot pair one packed gain ramp and two FFMA2 into
// the (left, right) slot accumulators.
// ----------------------------------------------------------------------------------------------
#ifndef BEAR_MIX_STAGES
#define BEAR_MIX_STAGES 2
#endif
constexpr int kMixStages = BEAR_MIX_STAGES;      // chunks in flight: the one being summed + kMixStages - 1 requested ahead

#ifndef BEAR_MIX_MINBLOCKS
#define BEAR_MIX_MINBLOCKS 7
#endif
// KEEP: at a run boundary only the slots whose loudspeaker changes are closed (worth it when runs are
// short: few channels per CTA); otherwise all slots are closed and start from zero (cheaper per run
// boundary when runs are long -- measured 1.94 against 2.05 ms at 1024 channels per CTA).
template <bool KEEP>
__global__ void __launch_bounds__(kMixTile, BEAR_MIX_MINBLOCKS) mix_kernel(Layout p, int T)
{
  extern __shared__ __align__(16) float smx[];
  const int XW = p.hist + kMixTile;
  float *const xt0 = smx;                                             // [stages][kMixChunk][XW] input rows
  float *const rc0 = smx + kMixStages * kMixChunk * XW;               // [stages][kMixChunk][kMixRecWords] records
  __shared__ __align__(8) unsigned long long bars[2 * kMixStages];    // full[stage], empty[stage]
  __shared__ unsigned sord[kMixStages][kMixChunk];                    // order entries of the staged chunks
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
  // Slot state, the same in all threads.  KEEP: the loudspeaker of every slot (255 = none); otherwise the
  // number of slots in use and their loudspeakers as the record packs them.
  unsigned slot_spk[kMixSlots];
#pragma unroll
  for (int i = 0; i < kMixSlots; i++) slot_spk[i] = 255u;
  unsigned nact = 0u, cur_lo = 0u, cur_hi = 0u;
  int run_left = 0;                               // channels of the current run still to come
  unsigned qdirty = 0u;                           // slots that have seen a channel with diffuse gains
  // Slot sums are added to the bus rows of their loudspeaker with reductions (red.global.add: nothing
  // comes back, so the thread does not wait for L2; every address belongs to one thread, whose stores and
  // reductions to it are performed in program order -- the result does not depend on timing).
  auto close_slot = [&](int i, unsigned l) {
    float vL, vR;
    unpack2(acc[i], vL, vR);
    float *pd = od + (size_t)(2 * l) * os;
    atomicAdd(pd, vL);
    atomicAdd(pd + os, vR);
    if (qdirty >> i & 1u) {
      float qa, qb;
      unpack2(accQ[i >> 1], qa, qb);
      atomicAdd(oq + (size_t)l * os, (i & 1) ? qb : qa);
      if (KEEP) {
        accQ[i >> 1] = (i & 1) ? pack2(qa, 0.f) : pack2(0.f, qb);
        qdirty &= ~(1u << i);
      }
    }
    if (KEEP) acc[i] = 0ull;
  };
  // all slots in use are closed and start from zero
  auto close_all = [&]() {
#pragma unroll
    for (int i = 0; i < kMixSlots; i++)
      if (KEEP ? slot_spk[i] != 255u : i < (int)nact)
        close_slot(i, KEEP ? slot_spk[i] : ((i < 4 ? cur_lo : cur_hi) >> (8 * (i & 3))) & 0xffu);
    if (!KEEP) {
#pragma unroll
      for (int i = 0; i < kMixSlots; i++) acc[i] = 0ull;
#pragma unroll
      for (int i = 0; i < kMixSlots / 2; i++) accQ[i] = 0ull;
      qdirty = 0u;
    }
  };
  // a run with a new loudspeaker set begins.  KEEP: a slot is closed only if the run needs it for another
  // loudspeaker; slots whose loudspeaker it shares, or which it does not use, simply go on
  auto retarget = [&](const uint4 meta) {
    if (KEEP) {
#pragma unroll
      for (int i = 0; i < kMixSlots; i++) {
        const unsigned want = ((i < 4 ? meta.y : meta.z) >> (8 * (i & 3))) & 0xffu;
        if (want != 255u && want != slot_spk[i]) {
          if (slot_spk[i] != 255u) close_slot(i, slot_spk[i]);
          slot_spk[i] = want;
        }
      }
    } else {
      close_all();
      nact = meta.x & 0xffu; cur_lo = meta.y; cur_hi = meta.z;
    }
  };
  auto flush = [&]() { close_all(); };

  const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < kMixStages; i++) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * i) : "memory");                       // full
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * (kMixStages + i)), "r"(kMixTile / 32) : "memory");  // empty
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = (nlisted + kMixChunk - 1) / kMixChunk;
  const unsigned *order = p.mix_order + (size_t)k * C + c0;
  unsigned e_pref = (tid < min(kMixChunk, nlisted)) ? order[tid] : 0u;   // warp 0: order entry of the chunk to request next
  // warp 0 requests chunk ci into stage ci % kMixStages: lane i takes the i-th order entry and requests
  // that channel's input row and record (the diffuse half of a record only when the channel has diffuse gains)
  auto request = [&](int ci) {
    const int i0 = ci * kMixChunk, cnt = min(kMixChunk, nlisted - i0), st = ci % kMixStages;
    const unsigned bar = bar0 + 8 * st;
    if (ci >= kMixStages)   // every warp is done with the chunk that used this stage before
      mbar_wait(bar0 + 8 * (kMixStages + st), (unsigned)(ci / kMixStages - 1) & 1u);
    const unsigned e = e_pref;
    unsigned bytes = 0u;
    if (tid < cnt) {
      sord[st][tid] = e;
      bytes = (unsigned)XW * 4u + ((e >> 30 & 1u) ? kMixRecWords * 4u : (kMixRecWords - kMixSlots * 2) * 4u);
    }
    const int inext = i0 + kMixChunk + tid;
    e_pref = (tid < kMixChunk && inext < nlisted) ? order[inext] : 0u;
    unsigned total = bytes;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    __syncwarp();   // the lanes' sord entries are ordered before lane 0's (releasing) arrive
    if (tid == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(total) : "memory");
    __syncwarp();
    if (tid < cnt) {
      const unsigned c = e & 0x1ffffu;
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
  if (tid < 32)
    for (int ci = 0; ci < min(nchunks, kMixStages - 1); ci++) request(ci);
  auto lagrange = [&](const float *rec, const float *xr) {
    const float4 dl = *reinterpret_cast<const float4 *>(rec);
    // ramped_delay for both ears: the two products in one packed multiply (the same IEEE products), the
    // sums stay scalar -- a packed add behind a packed multiply would be contracted into FFMA2
    float pa, pb;
    unpack2(mul2(pack2(dl.z, dl.w), splat2(rd)), pa, pb);
    return delayed_read2(xr, __fadd_rn(dl.x, pa), xr, __fadd_rn(dl.y, pb));
  };
  for (int ci = 0; ci < nchunks; ci++) {
    const int st = ci % kMixStages;
    if (tid < 32 && ci + kMixStages - 1 < nchunks) request(ci + kMixStages - 1);
    mbar_wait(bar0 + 8 * st, (unsigned)(ci / kMixStages) & 1u);
    const int cnt = min(kMixChunk, nlisted - ci * kMixChunk);
    const float *xr = xt0 + st * kMixChunk * XW + p.hist + tid;
    const float *rec = rc0 + st * kMixChunk * kMixRecWords;
    f32x2 x2n = lagrange(rec, xr);
    for (int j = 0; j < cnt;) {
      if (run_left == 0) {   // a new loudspeaker set: re-target the slots that change, take the run's length
        retarget(*reinterpret_cast<const uint4 *>(rec + 4));
        run_left = (int)(sord[st][j] >> 17 & 0xfffu) + 1;
      }
      const int jend = min(cnt, j + run_left);
      run_left -= jend - j;
      // the channels of the run that are in this chunk: nothing in here depends on the gains.  Two per
      // trip, so that the delayed samples of the channel after the current one (whose reads are issued
      // ahead of the current one's sums) alternate between two names instead of being moved
      auto channel = [&](int jj, const float *rc, const float *xp, const f32x2 x2) {
        const float4 *r4 = reinterpret_cast<const float4 *>(rc);
#pragma unroll
        for (int i = 0; i < kMixSlots / 2; i++) {   // all slots, used or not (unused ones hold zero gains): no branch
          const float4 g = r4[2 + i];
          float ga, gb;
          unpack2(fma2(rg2, pack2(g.z, g.w), pack2(g.x, g.y)), ga, gb);
          fma2_acc(acc[2 * i], splat2(ga), x2);
          fma2_acc(acc[2 * i + 1], splat2(gb), x2);
        }
        if (sord[st][jj] >> 30 & 1u) {
          const f32x2 x0 = splat2(xp[0]);
          qdirty = (1u << kMixSlots) - 1u;
#pragma unroll
          for (int i = 0; i < kMixSlots / 2; i++) {
            const float4 q = r4[2 + kMixSlots / 2 + i];
            fma2_acc(accQ[i], fma2(rg2, pack2(q.z, q.w), pack2(q.x, q.y)), x0);
          }
        }
      };
      for (; j + 1 < jend; j += 2, rec += 2 * kMixRecWords, xr += 2 * XW) {
        const f32x2 x2b = lagrange(rec + kMixRecWords, xr + XW);
        channel(j, rec, xr, x2n);
        if (j + 2 < cnt) x2n = lagrange(rec + 2 * kMixRecWords, xr + 2 * XW);
        channel(j + 1, rec + kMixRecWords, xr + XW, x2b);
      }
      if (j < jend) {
        const f32x2 x2 = x2n;
        if (j + 1 < cnt) x2n = lagrange(rec + kMixRecWords, xr + XW);
        channel(j, rec, xr, x2);
        j++; rec += kMixRecWords; xr += XW;
      }
    }
    __syncwarp();
    if ((tid & 31) == 0)   // this warp is done with stage st
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8 * (kMixStages + st)) : "memory");
  }
  flush();
}

// mix_dense_kernel: the channels whose gains are dense (more than kMixSlots non-zero loudspeakers:
// objects with extent), from the gain tables themselves with all 3L accumulators in registers, added
// to the bus rows.  Channels are taken in chunks of kMixChunk: their input rows (history + tile) arrive
// by bulk copies (TMA, 1-D) on an mbarrier while all threads stage the chunk's 4 x L table values per
// channel in shared memory as pair-packed (start, start, delta, delta); per channel the inner loop is
// then one broadcast load, one packed ramp and two FFMA2 per loudspeaker pair.  Periods without dense
// channels return at once.
// grid: periods * B / kMixTile; block kMixTile; dynamic shared memory: kMixChunk rows + kMixChunk x 4 L + 4 floats.
__global__ void __launch_bounds__(kMixTile) mix_dense_kernel(Layout p, int T)
{
  extern __shared__ __align__(16) float smd[];
  const int XW = p.hist + kMixTile;
  float *const xt = smd;                                       // [kMixChunk][XW] input rows
  float *const sg = smd + kMixChunk * XW;                      // [kMixChunk][2][kMaxSpeakers * 2] direct / diffuse pairs
  float *const sdl = sg + kMixChunk * 4 * kMaxSpeakers;        // [kMixChunk][4] delay ramps
  __shared__ __align__(8) unsigned long long bar;
  __shared__ unsigned sent[kMixChunk];
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
  const f32x2 rg2 = splat2(rg);
  f32x2 acc[kMaxSpeakers];
  f32x2 accQ[kMaxSpeakers / 2];
#pragma unroll
  for (int l = 0; l < kMaxSpeakers; l++) acc[l] = 0ull;
#pragma unroll
  for (int l = 0; l < kMaxSpeakers / 2; l++) accQ[l] = 0ull;
  const unsigned bar_addr = (unsigned)__cvta_generic_to_shared(&bar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  unsigned phase = 0;
  for (int s = 0; s < p.S; s++) {
    const int c0 = min(C, s * per);
    const int nlisted = p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2];
    const int ndense = p.mix_counts[((size_t)k * kMaxObjectSplits + s) * 2 + 1];
    const unsigned *order = p.mix_order + (size_t)k * C + c0 + nlisted;
    for (int i0 = 0; i0 < ndense; i0 += kMixChunk) {
      const int cnt = min(kMixChunk, ndense - i0);
      __syncthreads();   // the previous chunk has been used (and the barrier is initialised)
      if (tid < 32) {
        unsigned e = 0u;
        if (tid < cnt) { e = order[i0 + tid]; sent[tid] = e; }
        if (tid == 0)
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(cnt * XW * 4) : "memory");
        __syncwarp();
        if (tid < cnt) {
          const float *src = p.xin + (size_t)(e & 0x1ffffu) * p.xs + (t - tid);
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"((unsigned)__cvta_generic_to_shared(xt + tid * XW)), "l"(src), "r"(XW * 4), "r"(bar_addr) : "memory");
        }
      }
      // gains of the chunk: entry (channel j, loudspeaker l) -> pair layout (g0[2i], g0[2i+1], dg[2i], dg[2i+1])
      for (int idx = tid; idx < cnt * kMaxSpeakers; idx += kMixTile) {
        const int j = idx / kMaxSpeakers, l = idx - j * kMaxSpeakers;
        const unsigned e = order[i0 + j];
        const int c = (int)(e & 0x1ffffu);
        const bool obj = c < p.N;
        const int ld = obj ? p.N : p.M;
        float g0 = 0.f, g1 = 0.f, q0 = 0.f, q1 = 0.f;
        if (l < L) {
          const float *g = (obj ? p.g_dir + (size_t)k * L * ld + c : p.g_ds + (size_t)k * L * ld + (c - p.N));
          g0 = g[(size_t)l * ld]; g1 = g[(size_t)(L + l) * ld];
          if (e >> 30 & 1u) {
            const float *q = p.g_dif + (size_t)k * L * p.N + c;
            q0 = q[(size_t)l * p.N]; q1 = q[(size_t)(L + l) * p.N];
          }
        }
        float *o = sg + j * 4 * kMaxSpeakers + 4 * (l >> 1) + (l & 1);
        o[0] = g0; o[2] = g1 - g0;
        o[2 * kMaxSpeakers] = q0; o[2 * kMaxSpeakers + 2] = q1 - q0;
      }
      if (tid < cnt * 4) {
        const int j = tid >> 2;
        sdl[tid] = p.mix_recs[((size_t)k * C + (order[i0 + j] & 0x1ffffu)) * kMixRecWords + (tid & 3)];
      }
      __syncthreads();
      mbar_wait(bar_addr, phase);
      phase ^= 1u;
      for (int j = 0; j < cnt; j++) {
        const float *xr = xt + j * XW + p.hist + tid;
        const float4 dl = *reinterpret_cast<const float4 *>(sdl + 4 * j);
        const f32x2 x2 = delayed_read2(xr, ramped_delay(dl.x, dl.z, rd), xr, ramped_delay(dl.y, dl.w, rd));
        const float4 *g4 = reinterpret_cast<const float4 *>(sg + j * 4 * kMaxSpeakers);
#pragma unroll
        for (int i = 0; i < kMaxSpeakers / 2; i++) {
          const float4 g = g4[i];
          float ga, gb;
          unpack2(fma2(rg2, pack2(g.z, g.w), pack2(g.x, g.y)), ga, gb);
          fma2_acc(acc[2 * i], splat2(ga), x2);
          fma2_acc(acc[2 * i + 1], splat2(gb), x2);
        }
        if (sent[j] >> 30 & 1u) {
          const f32x2 x0 = splat2(xr[0]);
#pragma unroll
          for (int i = 0; i < kMaxSpeakers / 2; i++) {
            const float4 q = g4[kMaxSpeakers / 2 + i];
            fma2_acc(accQ[i], fma2(rg2, pack2(q.z, q.w), pack2(q.x, q.y)), x0);
          }
        }
      }
    }
  }
  float *od = p.cdir + (size_t)p.B + t, *oq = p.qdif + (size_t)p.B + t;
#pragma unroll
  for (int l = 0; l < kMaxSpeakers; l++)
    if (l < L) {
      float vL, vR, qa, qb;
      unpack2(acc[l], vL, vR);
      unpack2(accQ[l >> 1], qa, qb);
      od[(size_t)(2 * l) * p.cs] += vL;
      od[(size_t)(2 * l + 1) * p.cs] += vR;
      oq[(size_t)l * p.cs] += (l & 1) ? qb : qa;
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

// bus_gather_kernel (one scene on W GPUs): the bus rows of this rank's loudspeakers [l0, l1), summed
// over all ranks' object groups in rank order, for the periods of this launch.  The peers' rows are
// read from their memory over NVLink with ordinary 16-byte loads (cudaIpc mappings), each sample once;
// the sum lands in this GPU's memory, which is what decorrelate_kernel and bus_fdl_kernel then read.
// grid: (ceil(periods * B / 1024), 3 (l1 - l0)); block 256; four samples per thread.
__global__ void __launch_bounds__(256) bus_gather_kernel(Layout p, int T)
{
  const int nl = p.l1 - p.l0, r = blockIdx.y;
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i >= (size_t)T * p.B) return;
  const bool dir = r < 2 * nl;
  const size_t off = (size_t)(dir ? 2 * p.l0 + r : p.l0 + (r - 2 * nl)) * p.cs + (size_t)(p.k0 + 1) * p.B + i;
  float4 v[kMaxRanks];
#pragma unroll
  for (int w = 0; w < kMaxRanks; w++)
    if (w < p.W) v[w] = *reinterpret_cast<const float4 *>((dir ? p.peer_cdir[w] : p.peer_qdif[w]) + off);
  float4 a = v[0];
#pragma unroll
  for (int w = 1; w < kMaxRanks; w++)
    if (w < p.W) { a.x += v[w].x; a.y += v[w].y; a.z += v[w].z; a.w += v[w].w; }
  *reinterpret_cast<float4 *>(const_cast<float *>(dir ? p.cin : p.qin) + off) = a;
}
// The same sums by one small CTA per SM that stays resident for the whole launch (pipelined calls): the
// gather is bound by NVLink, not by the SMs, and in this form it leaves six of an SM's seven mix_kernel
// slots to the next call's mix, which runs beside it.  About 16 float4 in flight per thread whatever
// the number of ranks (W ranks x 16/W elements).
template <int W>
__global__ void __launch_bounds__(128, 6) bus_gather_thin_kernel(Layout p, int T)
{
  constexpr int U = 16 / W;
  const int nl = p.l1 - p.l0;
  const unsigned per_row = (unsigned)((size_t)T * p.B / 4), total = 3u * nl * per_row;
  const unsigned stride = gridDim.x * 128;
  for (unsigned j0 = blockIdx.x * 128 + threadIdx.x; j0 < total; j0 += U * stride) {
    float4 v[U][W];
    size_t off[U];
    bool dir[U], on[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const unsigned j = j0 + u * stride;
      on[u] = j < total;
      const unsigned r = on[u] ? j / per_row : 0, i = on[u] ? j - r * per_row : 0;
      dir[u] = r < 2u * nl;
      off[u] = (size_t)(dir[u] ? 2 * p.l0 + r : p.l0 + (r - 2 * nl)) * p.cs + (size_t)(p.k0 + 1) * p.B + 4 * (size_t)i;
#pragma unroll
      for (int w = 0; w < W; w++)
        if (on[u]) v[u][w] = *reinterpret_cast<const float4 *>((dir[u] ? p.peer_cdir[w] : p.peer_qdif[w]) + off[u]);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (!on[u]) continue;
      float4 a = v[u][0];
#pragma unroll
      for (int w = 1; w < W; w++) { a.x += v[u][w].x; a.y += v[u][w].y; a.z += v[u][w].z; a.w += v[u][w].w; }
      *reinterpret_cast<float4 *>(const_cast<float *>(dir[u] ? p.cin : p.qin) + off[u]) = a;
    }
  }
}
template <int W>
static void launch_bus_gather_thin(const Layout &p, int T, cudaStream_t st)
{
  if (p.W == W) bus_gather_thin_kernel<W><<<148, 128, 0, st>>>(p, T);
  else if constexpr (W < kMaxRanks) launch_bus_gather_thin<W + 1>(p, T, st);
}
void launch_bus_gather(const Layout &p, int T, cudaStream_t st, bool thin)
{
  if (p.l1 <= p.l0) return;
  if (thin) launch_bus_gather_thin<2>(p, T, st);
  else bus_gather_kernel<<<dim3((T * p.B + 1023) / 1024, 3 * (p.l1 - p.l0)), 256, 0, st>>>(p, T);
}

// peer_sync_kernel: one warp; lane w looks after rank w.  Everything this GPU wrote before the launch
// (kernel boundary) is made visible system-wide before the counters are stored; a lane that waits
// reads its counter with acquire semantics, so kernels launched after this one see the peers' data.
// A peer that never arrives raises the error flag after ~20 s instead of hanging the GPU.
__global__ void __launch_bounds__(32) peer_sync_kernel(PeerSync s)
{
  const int w = threadIdx.x;
  if (w >= s.W) return;
  if (s.signal >> w & 1) {
    __threadfence_system();
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(s.flags[w] + s.rank), "r"(s.value) : "memory");
  }
  if (s.wait >> w & 1) {
    const long long t0 = clock64();
    for (;;) {
      int v;
      asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(s.flags[s.rank] + w) : "memory");
      if (v - s.value >= 0) break;
      if (clock64() - t0 > kPeerTimeoutCycles) { atomicOr(s.err, kErrPeerTimeout); break; }
    }
  }
}
void launch_peer_sync(const PeerSync &s, cudaStream_t st) { peer_sync_kernel<<<1, 32, 0, st>>>(s); }

// out_sum_kernel: the stereo result of a shared scene on the rank that receives it = the sum of the
// partial mixes every rank's ifft_out_kernel stored into this GPU's inbox.  grid (ceil(n/1024), 2) x 256.
__global__ void __launch_bounds__(256) out_sum_kernel(float *out, const float *inbox, size_t os, int W, size_t col0, size_t n)
{
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (i >= n) return;
  const size_t off = (size_t)blockIdx.y * os + col0 + i;
  float4 v[kMaxRanks];
#pragma unroll
  for (int w = 0; w < kMaxRanks; w++)
    if (w < W) v[w] = *reinterpret_cast<const float4 *>(inbox + (size_t)w * 2 * os + off);
  float4 a = v[0];
#pragma unroll
  for (int w = 1; w < kMaxRanks; w++)
    if (w < W) { a.x += v[w].x; a.y += v[w].y; a.z += v[w].z; a.w += v[w].w; }
  *reinterpret_cast<float4 *>(out + off) = a;
}
void launch_out_sum(float *out, const float *inbox, size_t os, int W, size_t col0, size_t n, cudaStream_t st)
{
  out_sum_kernel<<<dim3((unsigned)((n + 1023) / 1024), 2), 256, 0, st>>>(out, inbox, os, W, col0, n);
}

size_t mix_smem_bytes(int hist);
static size_t mix_dense_smem_bytes(int hist)
{
  return (size_t)kMixChunk * ((hist + kMixTile) + 4 * kMaxSpeakers + 4) * sizeof(float);
}
void launch_mix_prep(const Layout &p, int T, cudaStream_t st)
{
  const int C = p.N + p.M, per = (C + p.S - 1) / p.S;
  const dim3 grid((p.N + kPrepThreads - 1) / kPrepThreads + (p.M + kPrepThreads - 1) / kPrepThreads, T);
  mix_prep_kernel<<<grid, kPrepThreads, 0, st>>>(p, T);
  int n2 = 1;
  while (n2 < per) n2 <<= 1;
  mix_order_kernel<<<dim3(T, p.S), 256, (size_t)n2 * sizeof(unsigned long long), st>>>(p, T);
}
void launch_mix(const Layout &p, int T, cudaStream_t st)
{
  dim3 grid(T * p.B / kMixTile, p.S);
  const int per = (p.N + p.M + p.S - 1) / p.S;   // channels per CTA
  if (per <= 256) mix_kernel<true><<<grid, kMixTile, mix_smem_bytes(p.hist), st>>>(p, T);
  else mix_kernel<false><<<grid, kMixTile, mix_smem_bytes(p.hist), st>>>(p, T);
  mix_dense_kernel<<<T * p.B / kMixTile, kMixTile, mix_dense_smem_bytes(p.hist), st>>>(p, T);
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

// ----------------------------------------------------------------------------------------------
// hoa_mix_kernel: ambisonic bus amb[m][t] = sum_c G[m,c](ramp) hoa_in[c][t]   (dsp.cpp:178-180)
// grid: T*B/kMixTile; block kMixTile.
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kMixTile) hoa_mix_kernel(Layout p, int T)
{
  const int t = p.k0 * p.B + blockIdx.x * kMixTile + threadIdx.x;
  const int k = p.k0 + (blockIdx.x * kMixTile) / p.B;
  const int n = t - k * p.B;
  const float rg = (float)(n + p.goff) / (float)p.B;
  const float *x0 = p.xin + (size_t)(p.N + p.M) * p.xs + p.hist + t;
  for (int m0 = 0; m0 < p.A; m0 += 16) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = 0.0f;
    for (int c = 0; c < p.H; c++) {
      const float x = __ldg(x0 + (size_t)c * p.xs);
#pragma unroll
      for (int i = 0; i < 16; i++) {
        const int m = m0 + i;
        if (m < p.A) {
          const float g0 = __ldg(p.g_hoa + ((size_t)k * p.A + m) * p.H + c);
          const float g1 = __ldg(p.g_hoa + ((size_t)(k + 1) * p.A + m) * p.H + c);
          acc[i] = fmaf(fmaf(rg, g1 - g0, g0), x, acc[i]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 16; i++)
      if (m0 + i < p.A) p.amb[(size_t)(m0 + i) * p.ws + p.B + p.hist + t] = acc[i];
  }
}

void launch_hoa_mix(const Layout &p, int T, cudaStream_t st)
{
  hoa_mix_kernel<<<T * p.B / kMixTile, kMixTile, 0, st>>>(p, T);
}

// The per-period FFT kernels are compiled once per supported period size (128 .. 4096).
template <class F>
static void dispatch_logb(int logB, F &&f)
{
  switch (logB) {
    case 7: f(std::integral_constant<int, 7>{}); break;
    case 8: f(std::integral_constant<int, 8>{}); break;
    case 9: f(std::integral_constant<int, 9>{}); break;
    case 10: f(std::integral_constant<int, 10>{}); break;
    case 11: f(std::integral_constant<int, 11>{}); break;
    case 12: f(std::integral_constant<int, 12>{}); break;
    default: break;   // bear_cuda_create rejects other period sizes
  }
}

// ----------------------------------------------------------------------------------------------
// decorrelate_kernel: w[l] = dec[l] (*) q[l], one partition overlap-save (dsp.cpp:61-72,132-135).
// grid: T*L (l fastest); block: B/4; smem 2*B float2.
// ----------------------------------------------------------------------------------------------
template <int LOGB>
__global__ void decorrelate_kernel(Layout p, int T)
{
  extern __shared__ float2 sm[];
  constexpr int M = 1 << LOGB, NTHR = (M >> 3) < 32 ? 32 : (M >> 3);
  float2 *a = sm, *b = sm + M + (M >> 3);
  const int nl = p.l1 - p.l0;                                  // loudspeakers this rank convolves
  const int l = p.l0 + blockIdx.x % nl, k = p.k0 + blockIdx.x / nl;
  const int tid = threadIdx.x;
  constexpr int nthr = NTHR;
  // window [block k-1 | block k] of the diffuse bus (sum of the partials) straight into the registers
  // of the first FFT pass: thread tid takes samples tid + r M/8, eight loads in flight
  constexpr int Q = M >> 3;
  float2 v0[8];
#pragma unroll
  for (int r = 0; r < 8; r++) v0[r] = make_float2(0.f, 0.f);
  if (tid < Q) {
    const float2 *src = reinterpret_cast<const float2 *>(p.qin + (size_t)l * p.cs + (size_t)k * p.B) + tid;
#pragma unroll
    for (int r = 0; r < 8; r++) v0[r] = src[r * Q];
  }
  fft_first_pass_reg<false, LOGB>(v0, b, tid);
  float2 *z = fft_shared_c<false, LOGB, NTHR, 1>(b, a, p.tw, tid);
  float2 *o = (z == a) ? b : a;
  const float2 *h = p.h_dec + (size_t)l * M;
  for (int kk = tid; kk <= M / 2; kk += nthr) {
    if (kk == 0) {
      const float2 h0 = __ldg(h);
      const float x0 = (z[0].x + z[0].y) * h0.x, xM = (z[0].x - z[0].y) * h0.y;
      o[0] = make_float2(x0 + xM, x0 - xM);
    } else {
      const float2 w = __ldg(p.tw + kk);
      float2 xk, xm, zk, zm;
      rfft_post(z[padi(kk)], z[padi(M - kk)], w, xk, xm);
      xk = cmul(xk, __ldg(h + kk));
      xm = cmul(xm, __ldg(h + M - kk));
      irfft_pre(xk, xm, w, zk, zm);
      o[padi(kk)] = zk;
      o[padi(M - kk)] = zm;
    }
  }
  // inverse transform; its last pass writes the kept half (the last B samples) straight to the row
  float2 *dst = reinterpret_cast<float2 *>(p.w + (size_t)l * p.ws + p.B + p.hist + (size_t)k * p.B);
  fft_shared_c<true, LOGB, NTHR, 0, true>(o, z, p.tw, tid, dst);
}

void launch_decorrelate(const Layout &p, int T, cudaStream_t st)
{
  const int threads = p.B / 8 < 32 ? 32 : p.B / 8;
  const size_t smem = fft_smem_floats2(p.B) * sizeof(float2);
  dispatch_logb(p.logB, [&](auto lb) { decorrelate_kernel<decltype(lb)::value><<<T * (p.l1 - p.l0), threads, smem, st>>>(p, T); });
}

// ----------------------------------------------------------------------------------------------
// bus_fdl_kernel: assemble the 2B-sample window [block k-1, block k] of the convolver inputs and
// write their packed spectra into the frequency-domain delay line.
//   bus channel 2l+e : sum of the direct partials + decorrelator output l read through the
//                      static delay line 2l+e (dsp.cpp:137-158), ramped when the view changed
//   HOA channel a    : ambisonic bus a read through the fixed HOA delay.  The reference delays
//                      the decoder OUTPUT (dsp.cpp:190-197); delay and decode filter are both
//                      LTI, so the delay is commuted in front and the HOA terms join the same
//                      frequency-domain accumulation as the BRIR terms.
// A CTA transforms two channels with two thread groups in lockstep: both ears of one loudspeaker
// (they read the same decorrelator row through different delays) or two HOA channels.  The rows the
// fractional-delay gathers read are first staged in shared memory with 16-byte cp.async copies.
// grid: T * (L + ceil(A/2)) (unit fastest); block: 2 * B/8.
// ----------------------------------------------------------------------------------------------
template <int LOGB>
__global__ void bus_fdl_kernel(Layout p, int T, int nunits, int stage_direct)
{
  extern __shared__ __align__(16) float2 sm[];
  constexpr int M = 1 << LOGB, NTHR = (M >> 3) < 32 ? 32 : (M >> 3);
  const int unit = blockIdx.x % nunits, k = p.k0 + blockIdx.x / nunits;
  constexpr int nthr = NTHR;                                          // threads per group (blockDim.x / 2)
  const int g = threadIdx.x >= nthr, tid = threadIdx.x - g * nthr;   // group 0 / 1
  const int FB = M + (M >> 3);                                       // padded FFT buffer, float2
  float2 *a = sm + (size_t)g * 2 * FB, *b = a + FB;
  const int RW = p.hist + 2 * M;                                     // staged row: history + window
  float *rows = reinterpret_cast<float *>(sm + 4 * FB);              // [2][RW]
  float *cwin = rows + 2 * RW;                                       // [2][2M] direct-bus windows (S == 1)
  const float invB = 1.0f / (float)p.B;
  const int L2 = 2 * p.L;
  const int nl = p.l1 - p.l0;                                        // loudspeakers this rank convolves
  const bool bus = unit < nl;
  const int spk = p.l0 + unit, hu = unit - nl;                       // loudspeaker / HOA channel pair
  const int ch = bus ? 2 * spk + g : L2 + 2 * hu + g;
  const bool active = bus || ch < L2 + p.A;                          // odd A: last group idles

  // static delays at the ends of blocks k-2, k-1, k (table slot = block + 2); loaded before the
  // copies below so that their latency overlaps
  float ta = 0.f, tb = 0.f, tc = 0.f;
  if (bus) {
    ta = p.d_stat[(size_t)k * L2 + ch] * p.fs;
    tb = p.d_stat[(size_t)(k + 1) * L2 + ch] * p.fs;
    tc = p.d_stat[(size_t)(k + 2) * L2 + ch] * p.fs;
    // same rule as the per-channel delays (mix_prep_kernel): flag, then clamp to what the staged row holds
    const float dmax = (float)(p.hist - 4);
    if (tid == 0 && !(tc >= 0.f && tc <= dmax)) atomicOr(p.err_flag, kErrDelayRange);
    ta = fminf(fmaxf(ta, 0.f), dmax); tb = fminf(fmaxf(tb, 0.f), dmax); tc = fminf(fmaxf(tc, 0.f), dmax);
  }
  // stage the source rows -- columns [k*B, k*B + hist + 2B) of the (B + hist)-prefixed row -- and, for
  // bus units, both ears' direct-bus windows [block k-1 | block k]: one bulk copy (TMA, 1-D) each,
  // issued by thread 0, landing on one mbarrier that every thread then waits on
  __shared__ __align__(8) unsigned long long sbar;
  const unsigned sbar_addr = (unsigned)__cvta_generic_to_shared(&sbar);
  const bool cstaged = bus && stage_direct;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sbar_addr) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const unsigned row_bytes = (unsigned)RW * 4u, win_bytes = 2u * M * 4u;
    const int nrows = bus ? (p.dif ? 1 : 0) : min(2, p.A - 2 * hu);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 ::"r"(sbar_addr), "r"(nrows * row_bytes + (cstaged ? 2u * win_bytes : 0u)) : "memory");
    for (int r = 0; r < nrows; r++) {
      const float *src = (bus ? p.w + (size_t)spk * p.ws : p.amb + (size_t)(2 * hu + r) * p.ws) + (size_t)k * p.B;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"((unsigned)__cvta_generic_to_shared(rows + r * RW)), "l"(src), "r"(row_bytes), "r"(sbar_addr) : "memory");
    }
    if (cstaged)
      for (int r = 0; r < 2; r++) {
        const float *src = p.cin + (size_t)(2 * spk + r) * p.cs + (size_t)k * p.B;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"((unsigned)__cvta_generic_to_shared(cwin + r * 2 * M)), "l"(src), "r"(win_bytes), "r"(sbar_addr) : "memory");
      }
  }
  __syncthreads();   // the barrier is initialised and armed before anyone polls it
  {
    unsigned done;
    do {
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}"
                   : "=r"(done) : "r"(sbar_addr) : "memory");
    } while (!done);
  }

  constexpr int Q = M >> 3;
  float2 v0[8];
#pragma unroll
  for (int r = 0; r < 8; r++) v0[r] = make_float2(0.f, 0.f);
  if (active) {
    if (bus) {
      const float *wrow = rows + p.hist;    // window sample 0
      const bool dif = p.dif != 0;
      const float dda = tb - ta, ddb = tc - tb;
      const float2 *cw = reinterpret_cast<const float2 *>(cwin + g * 2 * M);
      if (tid < Q) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
          const int n = tid + r * Q;
          float2 v = make_float2(0.f, 0.f);
          if (cstaged) {
            v = cw[n];
          } else {
            v = reinterpret_cast<const float2 *>(p.cin + (size_t)ch * p.cs + (size_t)k * p.B)[n];
          }
          if (dif) {
            const int t = 2 * n;            // window sample, both of the pair lie in the same block
            const bool h = t >= p.B;        // false: block k-1, true: block k
            const int nn = h ? t - p.B : t;
            const float d0 = h ? tb : ta, dd = h ? ddb : dda;
            float ra, rb;
            unpack2(delayed_read2(wrow + t, ramped_delay(d0, dd, (float)(nn + p.doff) * invB),
                                  wrow + t + 1, ramped_delay(d0, dd, (float)(nn + 1 + p.doff) * invB)), ra, rb);
            v.x += ra; v.y += rb;
          }
          v0[r] = v;
        }
      }
    } else {
      const float *arow = rows + g * RW + p.hist;
      if (tid < Q) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
          const int n = tid + r * Q;
          v0[r] = make_float2(delayed_read(arow + 2 * n, p.hoa_delay), delayed_read(arow + 2 * n + 1, p.hoa_delay));
        }
      }
    }
  }
  // the assembled window goes into the first FFT pass from registers (thread tid holds samples tid + r M/8)
  fft_first_pass_reg<false, LOGB>(v0, b, tid);
  float2 *z = fft_shared_c<false, LOGB, NTHR, 1>(b, a, p.tw, tid);
  if (!active) return;
  float2 *dst = p.fdl + ((size_t)ch * p.R + p.P + k) * M;   // period k lives in slot P + k
  for (int kk = tid; kk <= M / 2; kk += nthr) {
    if (kk == 0) {
      dst[0] = make_float2(z[0].x + z[0].y, z[0].x - z[0].y);
    } else {
      float2 xk, xm;
      rfft_post(z[padi(kk)], z[padi(M - kk)], __ldg(p.tw + kk), xk, xm);
      dst[kk] = xk;
      dst[M - kk] = xm;
    }
  }
}

// shared memory of bus_fdl_kernel; the direct-bus windows are staged only when there is a single
// partial (S == 1) and everything still fits (it does up to B = 2048)
// (the bulk copies read cin, which is always this GPU's own memory: with peers it is the gathered sum)
static bool bus_fdl_stage_direct(int B, int hist)
{
  const size_t base = (size_t)2 * fft_smem_floats2(B) * sizeof(float2) + (size_t)2 * (hist + 2 * B) * sizeof(float);
  return base + (size_t)2 * 2 * B * sizeof(float) <= 200 * 1024;
}
static size_t bus_fdl_smem_bytes(int B, int hist)
{
  return (size_t)2 * fft_smem_floats2(B) * sizeof(float2) + (size_t)2 * (hist + 2 * B) * sizeof(float) +
         (bus_fdl_stage_direct(B, hist) ? (size_t)2 * 2 * B * sizeof(float) : 0);
}

void launch_bus_fdl(const Layout &p, int T, cudaStream_t st)
{
  const int threads = 2 * (p.B / 8 < 32 ? 32 : p.B / 8);
  const int nunits = (p.l1 - p.l0) + (p.H > 0 ? (p.A + 1) / 2 : 0);
  if (nunits == 0) return;
  const size_t smem = bus_fdl_smem_bytes(p.B, p.hist);
  const int stage = bus_fdl_stage_direct(p.B, p.hist) ? 1 : 0;
  dispatch_logb(p.logB, [&](auto lb) {
    bus_fdl_kernel<decltype(lb)::value><<<T * nunits, threads, smem, st>>>(p, T, nunits, stage);
  });
}

// ----------------------------------------------------------------------------------------------
// cmac_kernel: the partitioned complex multiply-accumulate (the roofline kernel).
//   Y_w[k][e][bin] = sum_{l<L, part<P} FDL[2l+e][k-part][bin] * H[view_w][2l+e][part][bin]
//                  + sum_{a<A} FDL[2L+a][k][bin] * Hhoa[a][e][bin]            w in {new, old}
// View-major: the CTA for period j loads the filter partitions of view(j) ONCE and uses them for
// both spectra that need them -- the new-view spectrum of period j and, when period j+1 changes
// view, the old-view spectrum of period j+1.  Walking the partitions in order, the FDL operand of
// the second accumulation is the first one's operand of the previous step, so each term costs one
// filter load and one FDL load.  (j = -1 is the view the previous call ended on.)
// Work is split KS ways by loudspeaker (the HOA decode terms form one more unit); partial spectra
// are summed by ifft_out_kernel.  Each thread owns two adjacent bins (one 16-byte load per operand
// per term), consecutive threads consecutive bin pairs: every warp request is 512 contiguous bytes.
// grid: (B/(2*threads), KS, 2(T+1)); block: min(B/2, 128).
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void cmac2(float4 &acc, const float4 x, const float4 h, const bool packed0)
{
  // bin 0 of the packed spectrum holds (DC, Nyquist), both real: two real products instead of a complex one
  // order of the two products of each component as in cmac2_packed: x.re * h first, then x.im * (i h)
  acc.x = fmaf(x.x, h.x, acc.x);
  if (packed0) {
    acc.y = fmaf(x.y, h.y, acc.y);
  } else {
    acc.y = fmaf(x.x, h.y, acc.y);
    acc.x = fmaf(-x.y, h.y, acc.x);
    acc.y = fmaf(x.y, h.x, acc.y);
  }
  acc.z = fmaf(x.z, h.z, acc.z); acc.z = fmaf(-x.w, h.w, acc.z);
  acc.w = fmaf(x.z, h.w, acc.w); acc.w = fmaf(x.w, h.z, acc.w);
}
__device__ __forceinline__ float4 ldg_stream(const float4 *p)
{
  float4 v;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// Partial spectra of up to CNT periods that use the same view: the new-view spectrum of each period
// and (HAS_B) the old-view spectrum of the period after it.  One filter load serves all of them.
// Flags are per CTA, not per period: a period of the item that does not need one of its two spectra
// still accumulates it (its result is not stored), so the inner loop has no branches.
template <int CNT, bool HAS_B>
__device__ __forceinline__ void cmac_item(const Layout &p, const int (&j)[kCmacGroup], int e, int ks, int pair, int v,
                                          int T, unsigned maskA, unsigned maskB)
{
  const int B = p.B, P = p.P, L = p.L, R = p.R;
  const int nl = p.l1 - p.l0;                              // units: this rank's loudspeakers, then the HOA decode terms
  const int nunits = nl + (p.H > 0 ? 1 : 0);
  const int per = (nunits + p.KS - 1) / p.KS;
  const int u0 = ks * per, u1 = min(nunits, u0 + per);
  const bool packed0 = pair == 0;
  const int hB = B >> 1;                                   // float4 per spectrum
  float4 accA[CNT], accB[HAS_B ? CNT : 1];
  int slotA[CNT];
#pragma unroll
  for (int g = 0; g < CNT; g++) {
    accA[g] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (HAS_B) accB[g] = accA[g];
    slotA[g] = P + j[g];                                   // slot of period j (j >= -1)
  }
  const float4 *fdl4 = reinterpret_cast<const float4 *>(p.fdl) + pair;
  for (int u = u0; u < u1; u++) {
    if (u < nl) {
      const int ch = 2 * (p.l0 + u) + e;
      const float4 *x = fdl4 + (size_t)ch * R * hB;
      const float4 *h = reinterpret_cast<const float4 *>(p.h_brir) + ((size_t)(v * 2 * L + ch) * P) * hB + pair;
      float4 xb[HAS_B ? CNT : 1];
      int slot[CNT];
#pragma unroll
      for (int g = 0; g < CNT; g++) {
        slot[g] = slotA[g];
        if (HAS_B) xb[g] = __ldg(x + (size_t)(slotA[g] + 1) * hB);   // period j+1
      }
#pragma unroll(CNT == 1 ? 6 : (CNT <= 3 ? 3 : 2))
      for (int part = 0; part < P; part++) {
        const float4 hv = ldg_stream(h + (size_t)part * hB);
#pragma unroll
        for (int g = 0; g < CNT; g++) {
          const float4 xa = __ldg(x + (size_t)slot[g] * hB);
          cmac2(accA[g], xa, hv, packed0);
          if (HAS_B) { cmac2(accB[g], xb[g], hv, packed0); xb[g] = xa; }
          slot[g]--;
        }
      }
    } else {
      for (int am = 0; am < p.A; am++) {
        const float4 *x = fdl4 + (size_t)(2 * L + am) * R * hB;
        const float4 hv = __ldg(reinterpret_cast<const float4 *>(p.h_hoa) + ((size_t)am * 2 + e) * hB + pair);
#pragma unroll
        for (int g = 0; g < CNT; g++) {
          cmac2(accA[g], __ldg(x + (size_t)slotA[g] * hB), hv, packed0);
          if (HAS_B) cmac2(accB[g], __ldg(x + (size_t)(slotA[g] + 1) * hB), hv, packed0);
        }
      }
    }
  }
  // partial spectra are scratch of this launch: indexed by the period within the launch
  float4 *y = reinterpret_cast<float4 *>(p.ypart) + pair;
#pragma unroll
  for (int g = 0; g < CNT; g++) {
    const int jl = j[g] - p.k0;
    if (maskA >> g & 1) y[((((size_t)jl * 2 + e) * 2 + 0) * p.KS + ks) * hB] = accA[g];
    if (HAS_B && (maskB >> g & 1)) y[((((size_t)(jl + 1) * 2 + e) * 2 + 1) * p.KS + ks) * hB] = accB[g];
  }
}

// The same for the grouped kernel, with the operands prefetched through shared memory: every thread
// copies its own 16-byte pieces (one filter piece + CNT delay-line pieces per term) with cp.async
// into a private ring of kCmacSlots pieces and reads them back itself, so no barrier is needed and
// loads in flight cost no registers.  Terms of one loudspeaker: part = -1 (delay-line slot of period
// j+1, the first operand of the old-view accumulation) then part = 0 .. P-1.
// The accumulation uses packed FMAs: acc(re,im) += x.re * (h.re,h.im) + x.im * (-h.im,h.re), two
// FFMA2 per bin.  Only the warp that owns packed bin 0 (DC, Nyquist) takes the scalar path.
constexpr int kCmacSlots = 21;     // float4 pieces per thread
constexpr int kCmacRingStride = 128;   // float4 between the pieces of one thread (= max threads per CTA)
struct CmacAcc { f32x2 lo, hi; };  // bins 2*pair and 2*pair+1

__device__ __forceinline__ void cmac2_packed(CmacAcc &acc, const float4 x, const f32x2 h01, const f32x2 r01,
                                             const f32x2 h23, const f32x2 r23)
{
  acc.lo = fma2(splat2(x.x), h01, acc.lo); acc.lo = fma2(splat2(x.y), r01, acc.lo);
  acc.hi = fma2(splat2(x.z), h23, acc.hi); acc.hi = fma2(splat2(x.w), r23, acc.hi);
}
__device__ __forceinline__ void cmac2_scalar(CmacAcc &acc, const float4 x, const float4 h, const bool packed0)
{
  float4 a;
  unpack2(acc.lo, a.x, a.y); unpack2(acc.hi, a.z, a.w);
  cmac2(a, x, h, packed0);
  acc.lo = pack2(a.x, a.y); acc.hi = pack2(a.z, a.w);
}

// delay-line pieces g = G .. CNT-1 of one term (the shared-memory offset is an immediate)
template <int G, int CNT>
__device__ __forceinline__ void cmac_request_fdl(unsigned pdst, const float4 *px, int (&poff)[CNT], int hB)
{
  if constexpr (G < CNT) {
    asm volatile("cp.async.cg.shared.global [%0+%2], [%1], 16;" ::"r"(pdst), "l"(px + poff[G]),
                 "n"((1 + G) * kCmacRingStride * (int)sizeof(float4)) : "memory");
    poff[G] -= hB;
    cmac_request_fdl<G + 1, CNT>(pdst, px, poff, hB);
  }
}

template <int CNT, bool HAS_B>
__device__ __forceinline__ void cmac_item_pipe(const Layout &p, const int (&j)[kCmacGroup], int e, int ks, int KS,
                                               int pair, int v, unsigned maskA, unsigned maskB, float4 *ring)
{
  constexpr int W = 1 + CNT;                 // pieces per term
  constexpr int S = kCmacSlots / W;          // stages
  constexpr int RS = kCmacRingStride;
  constexpr unsigned kStageBytes = W * RS * sizeof(float4);
  const int B = p.B, P = p.P, L = p.L, R = p.R;
  const int nl = p.l1 - p.l0;
  const int nunits = nl + (p.H > 0 ? 1 : 0);
  const int per = (nunits + KS - 1) / KS;
  const int u0 = ks * per, u1 = min(nunits, u0 + per), ub = min(u1, nl);  // [u0, ub): BRIR units
  const bool packed0 = pair == 0;
  const bool scalar_warp = blockIdx.x == 0 && threadIdx.x < 32;           // the warp with packed bin 0
  const int hB = B >> 1;
  CmacAcc accA[CNT], accB[HAS_B ? CNT : 1];
  float4 xb[HAS_B ? CNT : 1];
  int slotA[CNT];
#pragma unroll
  for (int g = 0; g < CNT; g++) {
    accA[g].lo = accA[g].hi = 0ull;
    if (HAS_B) { accB[g] = accA[g]; xb[g] = make_float4(0.f, 0.f, 0.f, 0.f); }
    slotA[g] = P + j[g];
  }
  const float4 *fdl4 = reinterpret_cast<const float4 *>(p.fdl) + pair;
  constexpr int first_part = HAS_B ? -1 : 0;
  const int tpu = P - first_part;                              // terms per unit
  const int nterms = (ub - u0) * tpu;
  // ---- producer: walks (unit, part) one term ahead of the consumer by S-1 stages
  const int ringlen = R * hB;                                  // float4 per delay-line channel
  int poff[CNT];                                               // float4 offset of the next delay-line piece
#pragma unroll
  for (int g = 0; g < CNT; g++) {
    poff[g] = (HAS_B ? slotA[g] + 1 : slotA[g]) * hB;
  }
  const float4 *px = fdl4 + (size_t)(2 * (p.l0 + u0) + e) * ringlen;
  const float4 *ph = reinterpret_cast<const float4 *>(p.h_brir) + ((size_t)(v * 2 * L + 2 * (p.l0 + u0) + e) * P) * hB + pair;
  const unsigned ring0 = (unsigned)__cvta_generic_to_shared(ring + threadIdx.x);
  unsigned pdst = ring0;
  int pstage = 0, ppart = first_part, pt = 0;
  auto request = [&]() {
    if (pt < nterms) {
      if (ppart >= 0) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(pdst), "l"(ph) : "memory");
        ph += hB;
      }
      cmac_request_fdl<0, CNT>(pdst, px, poff, hB);
      pt++;
      pdst += kStageBytes;
      if (++pstage == S) { pstage = 0; pdst = ring0; }
      if (++ppart == P) {                                      // next loudspeaker, same ear
        ppart = first_part;
        px += 2 * (size_t)ringlen;
        ph += (size_t)P * hB;
#pragma unroll
        for (int g = 0; g < CNT; g++) poff[g] += tpu * hB;
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll 1
  for (int i = 0; i < S - 1; i++) request();
  // ---- consumer
  const float4 *csrc = ring + threadIdx.x;
  int cstage = 0, part = first_part;
#pragma unroll 1
  for (int t = 0; t < nterms; t++) {
    request();
    // "memory": the reads below must not be moved above the wait (there is no barrier to stop them)
    asm volatile("cp.async.wait_group %0;" ::"n"(S - 1) : "memory");
    if (HAS_B && part < 0) {
#pragma unroll
      for (int g = 0; g < CNT; g++) xb[g] = csrc[(1 + g) * RS];
    } else {
      const float4 hv = csrc[0];
      if (scalar_warp) {
#pragma unroll
        for (int g = 0; g < CNT; g++) {
          const float4 xa = csrc[(1 + g) * RS];
          cmac2_scalar(accA[g], xa, hv, packed0);
          if (HAS_B) { cmac2_scalar(accB[g], xb[g], hv, packed0); xb[g] = xa; }
        }
      } else {
        const f32x2 h01 = pack2(hv.x, hv.y), r01 = pack2(-hv.y, hv.x), h23 = pack2(hv.z, hv.w), r23 = pack2(-hv.w, hv.z);
#pragma unroll
        for (int g = 0; g < CNT; g++) {
          const float4 xa = csrc[(1 + g) * RS];
          cmac2_packed(accA[g], xa, h01, r01, h23, r23);
          if (HAS_B) { cmac2_packed(accB[g], xb[g], h01, r01, h23, r23); xb[g] = xa; }
        }
      }
    }
    if (++part == P) part = first_part;
    csrc += W * RS;
    if (++cstage == S) { cstage = 0; csrc = ring + threadIdx.x; }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (p.H > 0 && u0 <= nl && nl < u1) {   // the HOA decode unit (index nl) is in this split
    for (int am = 0; am < p.A; am++) {
      const float4 *x = fdl4 + (size_t)(2 * L + am) * ringlen;
      const float4 hv = __ldg(reinterpret_cast<const float4 *>(p.h_hoa) + ((size_t)am * 2 + e) * hB + pair);
#pragma unroll
      for (int g = 0; g < CNT; g++) {
        cmac2_scalar(accA[g], __ldg(x + (size_t)slotA[g] * hB), hv, packed0);
        if (HAS_B) cmac2_scalar(accB[g], __ldg(x + (size_t)(slotA[g] + 1) * hB), hv, packed0);
      }
    }
  }
  float4 *y = reinterpret_cast<float4 *>(p.ypart) + pair;
#pragma unroll
  for (int g = 0; g < CNT; g++) {
    const int jl = j[g] - p.k0;
    float4 a;
    if (maskA >> g & 1) {
      unpack2(accA[g].lo, a.x, a.y); unpack2(accA[g].hi, a.z, a.w);
      y[((((size_t)jl * 2 + e) * 2 + 0) * KS + ks) * hB] = a;
    }
    if (HAS_B && (maskB >> g & 1)) {
      unpack2(accB[g].lo, a.x, a.y); unpack2(accB[g].hi, a.z, a.w);
      y[((((size_t)(jl + 1) * 2 + e) * 2 + 1) * KS + ks) * hB] = a;
    }
  }
}

template <int CNT>
__device__ __forceinline__ void cmac_item_pipe_flags(const Layout &p, const int (&j)[kCmacGroup], int e, int ks, int KS,
                                                     int pair, int v, unsigned maskA, unsigned maskB, float4 *ring)
{
  if (maskB) cmac_item_pipe<CNT, true>(p, j, e, ks, KS, pair, v, maskA, maskB, ring);
  else cmac_item_pipe<CNT, false>(p, j, e, ks, KS, pair, v, maskA, maskB, ring);
}

template <int CNT>
__device__ __forceinline__ void cmac_item_flags(const Layout &p, const int (&j)[kCmacGroup], int e, int ks, int pair,
                                                int v, int T, unsigned maskA, unsigned maskB)
{
  if (maskB) cmac_item<CNT, true>(p, j, e, ks, pair, v, T, maskA, maskB);
  else cmac_item<CNT, false>(p, j, e, ks, pair, v, T, maskA, maskB);
}

// GROUPED = false: item z is period k0 - 1 + z.  GROUPED = true: item z is entry z of the table
// view_group_kernel made for this launch (up to kCmacGroup periods of one view).
template <bool GROUPED>
__global__ void __launch_bounds__(128, 4) cmac_kernel(Layout p, int T)
{
  const int pair = blockIdx.x * blockDim.x + threadIdx.x;  // bins 2*pair, 2*pair+1
  const int ks = blockIdx.y;
  const int item = blockIdx.z >> 1, e = blockIdx.z & 1;
  int j[kCmacGroup];
  int cnt = 1;
  int KS = p.KS;
  if (GROUPED) {
    // the builder decided how many of the launch's p.KS splits to use once it knew the item count
    KS = p.grp_count[1];
    if (item >= p.grp_count[0] || ks >= KS) return;
    const int2 it = p.grp_items[item];
    cnt = it.y;
#pragma unroll
    for (int g = 0; g < kCmacGroup; g++) j[g] = p.k0 - 1 + p.grp_sorted[it.x + min(g, cnt - 1)];
  } else {
#pragma unroll
    for (int g = 0; g < kCmacGroup; g++) j[g] = p.k0 - 1 + item;
  }
  const int v = p.view[j[0] + 1];                           // view of period j (slot 0: previous call)
  unsigned maskA = 0, maskB = 0;
#pragma unroll
  for (int g = 0; g < kCmacGroup; g++) {
    if (g < cnt) {
      if (j[g] >= p.k0) maskA |= 1u << g;                   // earlier launches already did their periods
      if (j[g] + 1 < p.k0 + T && p.view[j[g] + 2] != v) maskB |= 1u << g;   // period j+1 must be in this launch
    }
  }
  if (!maskA && !maskB) return;
  if (!GROUPED) { cmac_item_flags<1>(p, j, e, ks, pair, v, T, maskA, maskB); return; }
  extern __shared__ __align__(16) float4 cmac_ring[];
  switch (cnt) {
    case 1: cmac_item_pipe_flags<1>(p, j, e, ks, KS, pair, v, maskA, maskB, cmac_ring); break;
    case 2: cmac_item_pipe_flags<2>(p, j, e, ks, KS, pair, v, maskA, maskB, cmac_ring); break;
    case 3: cmac_item_pipe_flags<3>(p, j, e, ks, KS, pair, v, maskA, maskB, cmac_ring); break;
    case 4: cmac_item_pipe_flags<4>(p, j, e, ks, KS, pair, v, maskA, maskB, cmac_ring); break;
#if BEAR_CMAC_GROUP == 6
    case 5: cmac_item_pipe_flags<5>(p, j, e, ks, KS, pair, v, maskA, maskB, cmac_ring); break;
#endif
    default: cmac_item_pipe_flags<kCmacGroup>(p, j, e, ks, KS, pair, v, maskA, maskB, cmac_ring); break;
  }
}
static_assert(kCmacGroup == 5 || kCmacGroup == 6, "cmac_kernel dispatches on the item size");

// view_group_kernel: the periods k0-1 .. k0+n-1 of a launch sorted by view and cut into items of up
// to kCmacGroup periods (counting sort in one CTA).  grp_sorted holds period indices relative to
// k0-1, grp_items (first, count) pairs into it, grp_count the number of items.
// grid: 1; block: 1024; dynamic shared memory: 3*V + 2048 ints.
__global__ void __launch_bounds__(1024) view_group_kernel(Layout p, int n)
{
  extern __shared__ int gsh[];
  const int V = p.V, tid = threadIdx.x;
  int *cnt = gsh, *off = gsh + V, *ioff = gsh + 2 * V, *scanA = gsh + 3 * V, *scanB = scanA + 1024;
  const int32_t *view = p.view + p.k0;                       // view[i] = view of period k0 - 1 + i
  for (int v = tid; v < V; v += 1024) cnt[v] = 0;
  __syncthreads();
  for (int i = tid; i <= n; i += 1024) atomicAdd(&cnt[view[i]], 1);
  __syncthreads();
  // exclusive scans over the views of the counts and of the items per view
  const int cpt = (V + 1023) / 1024, v0 = tid * cpt, v1 = min(V, v0 + cpt);
  int sumc = 0, sumi = 0;
  for (int v = v0; v < v1; v++) { sumc += cnt[v]; sumi += (cnt[v] + kCmacGroup - 1) / kCmacGroup; }
  scanA[tid] = sumc; scanB[tid] = sumi;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    const int a = tid >= d ? scanA[tid - d] : 0, b = tid >= d ? scanB[tid - d] : 0;
    __syncthreads();
    scanA[tid] += a; scanB[tid] += b;
    __syncthreads();
  }
  int oc = scanA[tid] - sumc, oi = scanB[tid] - sumi;
  for (int v = v0; v < v1; v++) {
    off[v] = oc; ioff[v] = oi;
    const int c = cnt[v], ni = (c + kCmacGroup - 1) / kCmacGroup;
    for (int s = 0; s < ni; s++) p.grp_items[oi + s] = make_int2(oc + s * kCmacGroup, min(kCmacGroup, c - s * kCmacGroup));
    oc += c; oi += ni;
  }
  if (tid == 1023) {
    // splits: about 12 CTAs per SM over the items there turned out to be (at most the launch's p.KS)
    const int nitems = scanB[1023];
    const int threads = p.B / 2 < 128 ? p.B / 2 : 128;
    const int base = (p.B / 2 / threads) * 2 * max(nitems, 1);
    const int nunits = (p.l1 - p.l0) + (p.H > 0 ? 1 : 0);
    p.grp_count[0] = nitems;
    p.grp_count[1] = max(1, min(min((12 * 148 + base - 1) / base, nunits), p.KS));
  }
  __syncthreads();
  for (int i = tid; i <= n; i += 1024) p.grp_sorted[atomicAdd(&off[view[i]], 1)] = i;
}

// Grouping pays when views repeat within the launch; it needs the view histogram in shared memory.
bool cmac_groups_by_view(const Layout &p, int T) { return T >= 8 && p.V <= 2560 && p.grp_items != nullptr; }

int launch_cmac(const Layout &p, int T, cudaStream_t st)
{
  int threads = p.B / 2 < 128 ? p.B / 2 : 128;
  dim3 grid(p.B / 2 / threads, p.KS, 2 * (T + 1));
  if (cmac_groups_by_view(p, T)) {
    view_group_kernel<<<1, 1024, (3 * p.V + 2048) * sizeof(int), st>>>(p, T);
    cmac_kernel<true><<<grid, threads, (size_t)kCmacSlots * kCmacRingStride * sizeof(float4), st>>>(p, T);
    return 2;
  }
  cmac_kernel<false><<<grid, threads, 0, st>>>(p, T);
  return 1;
}

// ----------------------------------------------------------------------------------------------
// ifft_out_kernel: out[e] block k = irFFT(sum of partial spectra), last B samples; in a
// view-change block the old-view and new-view results are crossfaded linearly
// (rcl::InterpolatingFirFilterMatrix with transitionSamples = period, dsp.cpp:73-87).
// grid: 2T; block: B/4.
// ----------------------------------------------------------------------------------------------
template <int LOGB>
__global__ void __launch_bounds__(512) ifft_out_kernel(Layout p, int T)
{
  extern __shared__ float2 sm[];
  constexpr int M = 1 << LOGB;
  constexpr int NTHR = (M >> 2) < 32 ? 32 : ((M >> 2) > 256 ? 256 : (M >> 2));   // threads per group
  const int FB = M + (M >> 3);
  const int k = p.k0 + (blockIdx.x >> 1), e = blockIdx.x & 1;
  // two thread groups in lockstep: group 0 transforms the new-view spectrum, group 1 the old-view
  // one of a view-change period (idle otherwise, it only keeps the barriers company)
  constexpr int nthr = NTHR;
  const int g = threadIdx.x >= nthr, tid = threadIdx.x - g * nthr;
  float2 *a = sm + (size_t)g * 2 * FB, *b = a + FB;
  const bool xf = p.view[k + 1] != p.view[k];
  const bool work = g == 0 || xf;
  const float invB = 1.0f / (float)p.B;
  float2 *dst = reinterpret_cast<float2 *>(p.out + (size_t)e * p.os + (size_t)k * p.B);
  const int KS = p.ks_dev ? p.grp_count[1] : p.KS;
  if (work) {
    const float2 *y = p.ypart + ((((size_t)(k - p.k0) * 2 + e) * 2 + g) * KS) * M;   // g = 1: old view
    for (int kk = tid; kk <= M / 2; kk += nthr) {
      // partial spectra summed in order; the loads of four partials are issued together
      float2 xk = make_float2(0.f, 0.f), xm = xk;
      const int km = kk ? M - kk : 0;
      int s = 0;
      for (; s + 4 <= KS; s += 4) {
        float2 u[4], v[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { u[i] = y[(size_t)(s + i) * M + kk]; v[i] = y[(size_t)(s + i) * M + km]; }
#pragma unroll
        for (int i = 0; i < 4; i++) { xk.x += u[i].x; xk.y += u[i].y; xm.x += v[i].x; xm.y += v[i].y; }
      }
      for (; s < KS; s++) {
        const float2 u = y[(size_t)s * M + kk], v = y[(size_t)s * M + km];
        xk.x += u.x; xk.y += u.y; xm.x += v.x; xm.y += v.y;
      }
      if (kk == 0) {
        a[0] = make_float2(xk.x + xk.y, xk.x - xk.y);
      } else {
        float2 zk, zm;
        irfft_pre(xk, xm, __ldg(p.tw + kk), zk, zm);
        a[padi(kk)] = zk;
        a[padi(M - kk)] = zm;
      }
    }
  }
  float2 *z = fft_shared_c<true, LOGB, NTHR>(a, b, p.tw, tid);
  // old view faded out (in place), then the new view faded in on top of it
  if (g == 1 && xf) {
    for (int n = tid; n < M / 2; n += nthr) {
      float2 v = z[padi(M / 2 + n)];
      const float r0 = (float)(2 * n + p.xoff) * invB, r1 = (float)(2 * n + 1 + p.xoff) * invB;
      v.x *= (1.0f - r0); v.y *= (1.0f - r1);
      z[padi(M / 2 + n)] = v;
    }
  }
  __syncthreads();
  if (g == 0) {
    // group 1's result sits in the same one of its two buffers as ours (same pass count)
    const float2 *zo = z + 2 * FB;
    for (int n = tid; n < M / 2; n += nthr) {
      float2 v = z[padi(M / 2 + n)];
      if (xf) {
        const float r0 = (float)(2 * n + p.xoff) * invB, r1 = (float)(2 * n + 1 + p.xoff) * invB;
        const float2 o = zo[padi(M / 2 + n)];
        v.x = fmaf(r0, v.x, o.x); v.y = fmaf(r1, v.y, o.y);
      }
      dst[n] = v;
    }
  }
}

void launch_ifft_out(const Layout &p, int T, cudaStream_t st)
{
  // B/4 threads per group (the FFT passes use the first B/8 of them, the partial-sum and output
  // loops all): a single period is two CTAs, so the sum over the partial spectra is spread wide
  const int threads = 2 * (p.B / 4 < 32 ? 32 : (p.B / 4 > 256 ? 256 : p.B / 4));
  const size_t smem = 2 * fft_smem_floats2(p.B) * sizeof(float2);
  dispatch_logb(p.logB, [&](auto lb) { ifft_out_kernel<decltype(lb)::value><<<2 * T, threads, smem, st>>>(p, T); });
}

// ----------------------------------------------------------------------------------------------
// carry_kernel: after a call, move the tail of every history-carrying row to its front
// (memmove semantics: destination precedes source, chunks processed in ascending order).
// grid: total rows; block 256.
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) carry_kernel(CarryPlan plan)
{
  if ((int)blockIdx.x >= plan.row_start[plan.n]) {
    // delay line: the last P spectra of every channel move to slots 0..P-1.  One thread per
    // (channel, bin): it reads its own column of slots before it writes it, so source and destination
    // may overlap (calls of fewer than P periods) and no barrier is needed.
    const long long t = (long long)(blockIdx.x - plan.row_start[plan.n]) * blockDim.x + threadIdx.x;
    if (t >= (long long)plan.fdl_channels * plan.fdl_B) return;
    float2 *col = plan.fdl + (t / plan.fdl_B) * plan.fdl_stride + (t % plan.fdl_B);
    for (int s0 = 0; s0 < plan.fdl_P; s0 += 8) {
      float2 v[8];
#pragma unroll
      for (int i = 0; i < 8; i++)
        if (s0 + i < plan.fdl_P) v[i] = col[(size_t)(s0 + i + plan.fdl_T) * plan.fdl_B];
#pragma unroll
      for (int i = 0; i < 8; i++)
        if (s0 + i < plan.fdl_P) col[(size_t)(s0 + i) * plan.fdl_B] = v[i];
    }
    return;
  }
  int e = 0;
  while (e + 1 < plan.n && (int)blockIdx.x >= plan.row_start[e + 1]) e++;
  const CarryRow r = plan.e[e];
  const long long first = (long long)(blockIdx.x - plan.row_start[e]) * r.stride;
  float *row = r.base + first;
  const int count = r.total > 0 ? (int)min((long long)r.count, r.total - first) : r.count;
  if (r.shift >= count) {   // source and destination apart: a plain copy (table slots, long calls)
    for (int i = threadIdx.x; i < count; i += blockDim.x) row[i] = row[i + r.shift];
    return;
  }
  for (int c0 = 0; c0 < count; c0 += blockDim.x) {
    const int i = c0 + threadIdx.x;
    float v = 0.f;
    if (i < count) v = row[i + r.shift];
    __syncthreads();
    if (i < count) row[i] = v;
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------------
// unpack_kernel: the single-period host path uploads audio and tables as ONE packed buffer; this
// scatters the audio rows into the staging rows and copies each table into its "current period"
// slot (or repeats the previous period's values for ports that did not change).
// grid: 32; block 256.
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) unpack_kernel(Layout p, UnpackPlan plan)
{
  const int stride = gridDim.x * blockDim.x, g = blockIdx.x * blockDim.x + threadIdx.x;
  const int na = plan.rows * p.B;
  for (int i = g; i < na; i += stride) {
    const int row = i / p.B, col = i - row * p.B;
    p.xin[(size_t)row * p.xs + p.hist + col] = plan.audio[i];
  }
  for (int it = 0; it < plan.n; it++) {
    const UnpackItem u = plan.item[it];
    for (int i = g; i < u.count; i += stride) u.dst[i] = u.src[i];
  }
}

void launch_unpack(const Layout &p, const UnpackPlan &plan, cudaStream_t st)
{
  unpack_kernel<<<32, 256, 0, st>>>(p, plan);
}

// ----------------------------------------------------------------------------------------------
// frames_kernel: interleaved file frames (float32 or PCM, frame = one sample of every file channel)
// -> the planar float input rows.  tracks[i] = file channel of input row i (objects, DirectSpeakers,
// HOA in that order), -1 = silence.  A CTA turns a tile of 64 frames x 32 rows through shared memory:
// reads run along the frame (consecutive rows are mostly consecutive file channels), writes along time.
// PCM scaling: sample / 2^(bits-1).   grid: (ceil(frames/64), ceil(rows/32)); block: 256.
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ float frame_sample(const unsigned char *frames, size_t index, int format)
{
  switch (format) {
    case 1: return (float)reinterpret_cast<const short *>(frames)[index] * (1.0f / 32768.0f);
    case 2: {
      const unsigned char *b = frames + 3 * index;
      const int v = (int)((unsigned)b[0] << 8 | (unsigned)b[1] << 16 | (unsigned)b[2] << 24) >> 8;   // sign-extended
      return (float)v * (1.0f / 8388608.0f);
    }
    case 3: return (float)reinterpret_cast<const int *>(frames)[index] * (1.0f / 2147483648.0f);
    default: return reinterpret_cast<const float *>(frames)[index];
  }
}

__global__ void __launch_bounds__(256) frames_kernel(Layout p, const unsigned char *frames, int format, int frame_channels,
                                                     const int *tracks, int rows, int first_frame, int nframes)
{
  __shared__ float tile[32][65];
  const int f0 = blockIdx.x * 64, r0 = blockIdx.y * 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = r0 + lane;
  const int ch = row < rows ? tracks[row] : -1;
  for (int f = warp; f < 64; f += 8) {
    float v = 0.f;
    if (ch >= 0 && f0 + f < nframes) v = frame_sample(frames, (size_t)(first_frame + f0 + f) * frame_channels + ch, format);
    tile[lane][f] = v;
  }
  __syncthreads();
  for (int rr = warp; rr < 32; rr += 8) {
    if (r0 + rr >= rows) break;
    float *dst = p.xin + (size_t)(r0 + rr) * p.xs + p.hist + first_frame + f0;
    for (int f = lane; f < 64; f += 32)
      if (f0 + f < nframes) dst[f] = tile[rr][f];
  }
}

void launch_frames(const Layout &p, const void *frames, int format, int frame_channels, const int *tracks, int rows,
                   int first_frame, int nframes, cudaStream_t st)
{
  if (rows <= 0 || nframes <= 0) return;
  dim3 grid((nframes + 63) / 64, (rows + 31) / 32);
  frames_kernel<<<grid, 256, 0, st>>>(p, static_cast<const unsigned char *>(frames), format, frame_channels, tracks, rows,
                                      first_frame, nframes);
}

void launch_carry(const CarryPlan &plan, cudaStream_t st)
{
  const long long cols = (long long)plan.fdl_channels * plan.fdl_B;   // 0: a plan without the delay line
  const int grid = plan.row_start[plan.n] + (int)((cols + 255) / 256);
  if (grid > 0) carry_kernel<<<grid, 256, 0, st>>>(plan);
}

// ----------------------------------------------------------------------------------------------
// control_kernel: the Control graph's table producers between the gain calculators and the DSP
// ports, one warp per (period, channel); all arithmetic in double like the reference, float at the
// ports.  References: Panner::get_delays panner.cpp:189-208, compensation_gain :277-316,:332-346,
// :364-380, compensation_gain_direct :318-330,:348-362,:391-404, GainNorm::process gain_norm.cpp:26-44,
// DirectDiffuseSplit direct_diffuse_split.cpp:24-32, StaticDelayCalc static_delay_calc.cpp:19-30.
// Lane j holds gain j (and j+32): direct gains 0..L-1, diffuse gains L..2L-1.  The quadratic form of
// the gain-normalisation table is evaluated only over the non-zero gains (panning gains are sparse).
// grid: (ceil((N+M+1)/4), T); block: 128 (4 warps).
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(128) control_kernel(Layout p, int T)
{
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * 4 + (threadIdx.x >> 5);     // channel: objects, DirectSpeakers, then "static delays"
  const int t = blockIdx.y;
  const int L = p.L, G = 2 * L, N = p.N, M = p.M;
  if (c > N + M) return;
  const int v = p.view[t + 1];
  const float *dl = p.delays_tab + (size_t)v * L * 2;
  if (c == N + M) {  // StaticDelayCalc
    for (int i = lane; i < G; i += 32) p.d_stat[(size_t)(t + 2) * G + i] = (float)((double)dl[i] / p.fs_d);
    return;
  }
  const bool obj = c < N;
  // gains of this channel: g0 = entry `lane`, g1 = entry `lane + 32`
  float f0 = 0.f, f1 = 0.f;
  if (obj) {
    const float *raw = p.raw_obj + (size_t)t * G * N + c;
    if (lane < G) f0 = raw[(size_t)lane * N];
    if (lane + 32 < G) f1 = raw[(size_t)(lane + 32) * N];
  } else {
    const float *raw = p.raw_ds + (size_t)t * L * M + (c - N);
    if (lane < L) f0 = raw[(size_t)lane * M];
  }
  const double g0 = (double)f0, g1 = (double)f1;
  // ---- delays: gain-weighted mean of the loudspeaker delays (direct + diffuse gains for objects) ----
  // total gain of loudspeaker l = g[l] + g[L + l]; entry L + l lives in lane (L + l) & 31, slot (L + l) >> 5
  const int jd = L + lane;                               // index of this loudspeaker's diffuse gain
  const double a_ = __shfl_sync(0xffffffffu, g0, jd & 31), b_ = __shfl_sync(0xffffffffu, g1, jd & 31);
  double tot = 0.0;
  if (lane < L) {
    tot = g0;
    if (obj) tot += (jd < 32) ? a_ : b_;
  }
  const double dL = lane < L ? (double)dl[2 * lane] : 0.0, dR = lane < L ? (double)dl[2 * lane + 1] : 0.0;
  const double total = warp_sum(tot), sl = warp_sum(dL * tot), sr = warp_sum(dR * tot);
  double ml, mr;
  if (total > 1e-6) { ml = sl / total; mr = sr / total; }
  else { ml = (double)dl[2 * p.front]; mr = (double)dl[2 * p.front + 1]; }
  const float del[2] = {(float)((ml + p.dec_delay_d) / p.fs_d), (float)((mr + p.dec_delay_d) / p.fs_d)};
  // ---- compensation gain ----
  double comp = 1.0;
  if (p.gn_factors) {
    const int Gq = obj ? G : L;                      // size of the quadratic form
    const size_t GG = (size_t)G * G * 2;             // floats per (sample, view)
    const float *F0 = p.gn_factors + (size_t)v * GG;  // sample 0
    const size_t sample_stride = (size_t)p.V * GG;
    double e_acc = 0.0;
    if (lane < Gq) e_acc += g0 * g0 * ((double)F0[((size_t)lane * G + lane) * 2] + (double)F0[((size_t)lane * G + lane) * 2 + 1]);
    if (lane + 32 < Gq) {
      const int j = lane + 32;
      e_acc += g1 * g1 * ((double)F0[((size_t)j * G + j) * 2] + (double)F0[((size_t)j * G + j) * 2 + 1]);
    }
    const double expected = sqrt(warp_sum(e_acc));
    const unsigned nz0 = __ballot_sync(0xffffffffu, f0 != 0.f), nz1 = __ballot_sync(0xffffffffu, f1 != 0.f);
    double r_acc = 0.0;
    for (int e = 0; e < 2; e++) {
      int s_lo = 0, ns = 1;
      double pw[2] = {1.0, 0.0};
      if (obj) {
        const double unclamped = (double)del[e] * p.fs_d - p.dec_delay_d;
        const double d = fmin(fmax(0.0, unclamped), (double)p.gnS - 1.0);
        s_lo = min((int)floor(d), p.gnS - 2);
        ns = 2;
        pw[0] = 1.0 - fabs(d - (double)s_lo);
        pw[1] = 1.0 - fabs(d - (double)(s_lo + 1));
      }
      for (int si = 0; si < ns; si++) {
        const float *Fs = F0 + (size_t)(s_lo + si) * sample_stride + e;
        for (int half = 0; half < 2; half++) {
          unsigned m = half ? nz1 : nz0;
          while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const int i = b + 32 * half;
            if (i >= Gq) continue;
            const double gi = half ? __shfl_sync(0xffffffffu, g1, b) : __shfl_sync(0xffffffffu, g0, b);
            const float *row = Fs + (size_t)i * G * 2;
            double part = 0.0;
            if (lane < Gq) part += g0 * (double)row[(size_t)lane * 2];
            if (lane + 32 < Gq) part += g1 * (double)row[(size_t)(lane + 32) * 2];
            r_acc += pw[si] * gi * part;
          }
        }
      }
    }
    const double real = sqrt(warp_sum(r_acc));
    comp = fabs(real) < 1e-6 ? 1.0 : expected / real;
  }
  // ---- write the tables of period t (slot t + 1) ----
  if (obj) {
    if (lane < 2) p.d_dir[(size_t)(t + 1) * 2 * N + 2 * c + lane] = del[lane];
    if (lane < G) {
      float *dst = lane < L ? p.g_dir + ((size_t)(t + 1) * L + lane) * N : p.g_dif + ((size_t)(t + 1) * L + lane - L) * N;
      dst[c] = (float)(comp * g0);
    }
    if (lane + 32 < G) p.g_dif[((size_t)(t + 1) * L + lane + 32 - L) * N + c] = (float)(comp * g1);
  } else {
    const int j = c - N;
    if (lane < 2) p.d_ds[(size_t)(t + 1) * 2 * M + 2 * j + lane] = del[lane];
    if (lane < L) p.g_ds[((size_t)(t + 1) * L + lane) * M + j] = (float)(comp * g0);
  }
}

void launch_control(const Layout &p, int T, cudaStream_t st)
{
  dim3 grid((p.N + p.M + 1 + 3) / 4, T);
  control_kernel<<<grid, 128, 0, st>>>(p, T);
}

// ----------------------------------------------------------------------------------------------
// Library-owned scratch for the partial spectra (see bear_cuda_process_device_static_scratch).
// One array per GPU that loads this library, held until the process exits.  Host code must not use
// the symbol itself as a pointer: on the host `g_static_ypart` only stands in for the device array,
// so the device address is looked up once per device and handed to the kernels as a plain pointer.
// ----------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_ypart[kStaticScratchBytes];

cudaError_t static_scratch_pointer(float2 **out)
{
  constexpr int kMaxDevices = 64;
  static float2 *resolved[kMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  if (!resolved[dev]) {
    void *ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_ypart);  // address on the CURRENT device
    if (e != cudaSuccess) return e;
    resolved[dev] = static_cast<float2 *>(ptr);
  }
  *out = resolved[dev];
  return cudaSuccess;
}

size_t mix_smem_bytes(int hist) { return (size_t)kMixStages * kMixChunk * (hist + kMixTile + kMixRecWords) * sizeof(float); }

static int ilog2(int v)
{
  int l = 0;
  while ((1 << l) < v) l++;
  return l;
}

cudaError_t configure_mix(int hist, int B)
{
  cudaError_t e = cudaFuncSetAttribute(mix_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mix_smem_bytes(hist));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(mix_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mix_smem_bytes(hist));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(mix_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mix_dense_smem_bytes(hist));
  if (e != cudaSuccess) return e;
  dispatch_logb(ilog2(B), [&](auto lb) {
    e = cudaFuncSetAttribute(bus_fdl_kernel<decltype(lb)::value>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)bus_fdl_smem_bytes(B, hist));
  });
  return e;
}

cudaError_t configure_kernels(int B)
{
  const int smem = fft_smem_floats2(B) * (int)sizeof(float2);
  cudaError_t e = cudaSuccess;
  if (smem > 48 * 1024) {
    if ((e = cudaFuncSetAttribute(filter_fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem))) return e;
    dispatch_logb(ilog2(B), [&](auto lb) {
      e = cudaFuncSetAttribute(decorrelate_kernel<decltype(lb)::value>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    });
    if (e) return e;
  }
  if (2 * smem > 48 * 1024)   // two transforms per CTA
    dispatch_logb(ilog2(B), [&](auto lb) {
      e = cudaFuncSetAttribute(ifft_out_kernel<decltype(lb)::value>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * smem);
    });
  return e;
}

}  // namespace bearcuda
