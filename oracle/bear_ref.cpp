// bear_ref -- CPU ORACLE for the BEAR render hot path.  TEST INFRASTRUCTURE ONLY.
//
// This file is a plain, single-threaded restatement of the audio DSP graph that sits behind
// bear::Renderer::process() in the reference (visr_bear/src/dsp.cpp:7-236, per_ear_delay.cpp:4-57,
// brir_interpolation_controller.cpp:18-28), at the Control->DSP parameter-port boundary
// (visr_bear/src/top.cpp:34-51).  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load it; the product (bear_b200/csrc) never does.
//
// The arithmetic of the reference lives in VISR (rcl::GainMatrix, rcl::DelayVector,
// rcl::FirFilterMatrix, rcl::InterpolatingFirFilterMatrix, rcl::Add; pinned rev
// 6907b24a250c9dadf1028fb5e8764eaa078a184f of the `visr` branch of ebu/bear, flake.lock:94-104),
// which is NOT in /root/reference.  Each atom below restates VISR's published algorithm and is
// anchored on the reference's own call sites.  PARITY STATUS (see DESIGN.md "Oracle pinning"):
//   * graph wiring, index maps, initial state, table semantics: pinned by the reference source;
//   * partitioned convolution (steady state): cross-checked against the reference's in-tree
//     libear BlockConvolver (oracle/_ref) and a naive time-domain convolution;
//   * path-delay equalities of visr_bear/test/test_delays.py:114-136 hold on the in-tree fixture;
//   * gain ramp (n+1)/B, delay ramp (n+1)/B, centred Lagrange-3 with method delay 1.5, initial state:
//     PINNED on the reference's golden vectors (oracle/identify_conventions.py ->
//     tests/golden/pinned_conventions.json), and the three static no-head-track goldens are replayed
//     end to end at 2e-5 (tests/test_golden_reference.py);
//   * view-crossfade ramp (n+1)/B and the diffuse path: "parity unpinned" (no golden that can be
//     replayed without default.tf exercises them); the ramp offsets are runtime switches.
//
// Build: see oracle/Makefile (REAL=float -> libbear_ref_f32.so, REAL=double -> libbear_ref_f64.so).
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#ifndef REAL
#define REAL float
#endif
typedef REAL real;
typedef std::complex<real> cplx;

namespace {

// ------------------------------------------------------------------------------------------
// Real FFT of even length n (power of two), n/2+1 bins.  Stands in for VISR's FFT wrapper
// ("default" = kissfft, bear/render_cli.py:32-39); textbook radix-2 on the half-length
// complex sequence.
// ------------------------------------------------------------------------------------------
class RealFFT {
 public:
  explicit RealFFT(int n) : n_(n), h_(n / 2), rev_(n / 2), w_(n / 2), wr_(n / 2 + 1), tmp_(n / 2), tmp2_(n / 2)
  {
    int bits = 0;
    while ((1 << bits) < h_) bits++;
    for (int i = 0; i < h_; i++) {
      int r = 0;
      for (int b = 0; b < bits; b++)
        if (i & (1 << b)) r |= 1 << (bits - 1 - b);
      rev_[i] = r;
    }
    const double pi = 3.14159265358979323846264338327950288;
    for (int k = 0; k < h_; k++) {
      double a = -2.0 * pi * k / h_;
      w_[k] = cplx((real)std::cos(a), (real)std::sin(a));
    }
    for (int k = 0; k <= h_; k++) {
      double a = -2.0 * pi * k / n_;
      wr_[k] = cplx((real)std::cos(a), (real)std::sin(a));
    }
  }
  int size() const { return n_; }
  int bins() const { return h_ + 1; }

  void forward(const real *x, cplx *X)
  {
    for (int i = 0; i < h_; i++) tmp_[rev_[i]] = cplx(x[2 * i], x[2 * i + 1]);
    butterflies(false);
    for (int k = 0; k <= h_; k++) {
      cplx a = tmp_[k % h_], b = std::conj(tmp_[(h_ - k) % h_]);
      cplx ev = (a + b) * (real)0.5;
      cplx od = (a - b) * cplx(0, (real)-0.5);
      X[k] = ev + wr_[k] * od;
    }
  }
  // unnormalised inverse: forward() followed by inverse() multiplies by n
  void inverse(const cplx *X, real *x)
  {
    for (int k = 0; k < h_; k++) {
      cplx a = X[k], b = std::conj(X[h_ - k]);
      cplx ev = a + b;
      cplx od = (a - b) * std::conj(wr_[k]);
      tmp2_[rev_[k]] = ev + cplx(0, 1) * od;
    }
    std::swap(tmp_, tmp2_);
    butterflies(true);
    for (int i = 0; i < h_; i++) {
      x[2 * i] = tmp_[i].real();
      x[2 * i + 1] = tmp_[i].imag();
    }
  }

 private:
  void butterflies(bool inv)
  {
    for (int len = 2; len <= h_; len <<= 1) {
      int half = len / 2, step = h_ / len;
      for (int s = 0; s < h_; s += len)
        for (int j = 0; j < half; j++) {
          cplx w = w_[j * step];
          if (inv) w = std::conj(w);
          cplx u = tmp_[s + j], v = tmp_[s + j + half] * w;
          tmp_[s + j] = u + v;
          tmp_[s + j + half] = u - v;
        }
    }
  }
  int n_, h_;
  std::vector<int> rev_;
  std::vector<cplx> w_, wr_, tmp_, tmp2_;
};

inline real ramp_value(int n, int steps, int offset) { return (real)(n + offset) / (real)steps; }

// ------------------------------------------------------------------------------------------
// rcl::GainMatrix with interpolationSteps == period (per_ear_delay.cpp:31-36, dsp.cpp:127-129,
// :178-180): out[o][n] = sum_i (prev[o,i] + ramp[n] (next[o,i]-prev[o,i])) in[i][n]; gains start
// at 0; new gains arrive every block (SharedData ports, dsp.hpp:48,51,55,66).
// ------------------------------------------------------------------------------------------
struct GainMatrix {
  int n_in = 0, n_out = 0, B = 0, off = 1;
  std::vector<real> prev, next, ramp;
  void setup(int in, int out, int period, int ramp_offset)
  {
    n_in = in; n_out = out; B = period; off = ramp_offset;
    prev.assign((size_t)in * out, 0);
    next.assign((size_t)in * out, 0);
    ramp.resize(B);
    for (int n = 0; n < B; n++) ramp[n] = ramp_value(n, B, off);
  }
  // gains: row-major [n_out][n_in], may be null (no change)
  void process(const float *gains, const real *const *in, real *const *out)
  {
    if (gains)
      for (size_t i = 0; i < next.size(); i++) next[i] = (real)gains[i];
    for (int o = 0; o < n_out; o++) {
      real *y = out[o];
      std::fill(y, y + B, (real)0);
      for (int i = 0; i < n_in; i++) {
        real g0 = prev[(size_t)o * n_in + i], dg = next[(size_t)o * n_in + i] - g0;
        const real *x = in[i];
        const real *rp = ramp.data();
        for (int n = 0; n < B; n++) y[n] += x[n] * (g0 + dg * rp[n]);
      }
    }
    prev = next;
  }
};

// Lagrange order-3 coefficients at position s in [1,2) relative to the first of 4 nodes 0..3.
inline void lagrange3(real s, real c[4])
{
  real s1 = s - 1, s2 = s - 2, s3 = s - 3;
  c[0] = -(s1 * s2 * s3) * (real)(1.0 / 6.0);
  c[1] = (s * s2 * s3) * (real)0.5;
  c[2] = -(s * s1 * s3) * (real)0.5;
  c[3] = (s * s1 * s2) * (real)(1.0 / 6.0);
}

// ------------------------------------------------------------------------------------------
// rcl::DelayVector, "lagrangeOrder3", interpolationSteps == period, MethodDelayPolicy::Add
// (per_ear_delay.cpp:22-29, dsp.cpp:137-144,190-197).  Delay in samples d = seconds*fs is
// ramped linearly across the block; the interpolator's method delay (order/2 = 1.5 samples) is
// added, so the realised delay is d + 1.5 and the four taps are centred on it:
//   i = floor(d + 0.5), mu = d - i in [-0.5, 0.5), y[n] = sum_j c_j(1.5 + mu) x[n - i - j].
// ------------------------------------------------------------------------------------------
struct DelayVector {
  int ch = 0, B = 0, off = 1, hist = 0;
  real fs = 48000;
  std::vector<real> buf, cur, next, ramp;  // buf: per channel [hist past samples | B current samples]
  void setup(int channels, int period, double sample_rate, double initial_seconds, int max_delay_samples,
             int ramp_offset)
  {
    ch = channels; B = period; off = ramp_offset; fs = (real)sample_rate;
    hist = max_delay_samples + 8;
    buf.assign((size_t)ch * (hist + B), 0);
    cur.assign(ch, (real)initial_seconds * fs);
    next = cur;
    ramp.resize(B);
    for (int n = 0; n < B; n++) ramp[n] = ramp_value(n, B, off);
  }
  // delays_seconds: [ch] or null (no change)
  void process(const float *delays_seconds, const real *const *in, real *const *out)
  {
    if (delays_seconds)
      for (int c = 0; c < ch; c++) next[c] = (real)delays_seconds[c] * fs;
    for (int c = 0; c < ch; c++) {
      real *rb = &buf[(size_t)c * (hist + B)];
      std::memmove(rb, rb + B, sizeof(real) * hist);
      std::memcpy(rb + hist, in[c], sizeof(real) * B);
      const real *now = rb + hist;
      real d0 = cur[c], dd = next[c] - d0;
      for (int n = 0; n < B; n++) {
        real d = d0 + dd * ramp[n];
        real fi = std::floor(d + (real)0.5);
        real coef[4];
        lagrange3((real)1.5 + (d - fi), coef);
        const real *q = now + n - (int)fi;
        out[c][n] = coef[0] * q[0] + coef[1] * q[-1] + coef[2] * q[-2] + coef[3] * q[-3];
      }
    }
    cur = next;
  }
};

// ------------------------------------------------------------------------------------------
// Uniformly partitioned overlap-save convolver core (rbbl::CoreConvolverUniform as used by
// rcl::FirFilterMatrix / rcl::InterpolatingFirFilterMatrix): DFT size 2B, B+1 bins, one
// frequency-domain delay line (FDL) per input, filters held as P zero-padded partitions that
// are pre-scaled by 1/(2B) so the inverse transform needs no normalisation; the output block is
// the last B samples of the inverse transform.
// ------------------------------------------------------------------------------------------
struct Routing { int in, out, filter; real gain; };

struct UpolsCore {
  int n_in = 0, n_out = 0, B = 0, P = 0, bins = 0, head = 0;
  std::unique_ptr<RealFFT> fft;
  std::vector<real> window;             // [n_in][2B]
  std::vector<cplx> fdl;                // [n_in][P][bins]
  std::vector<cplx> filters;            // [n_filters][P][bins]
  std::vector<cplx> acc;
  std::vector<real> td;
  void setup(int inputs, int outputs, int period, int filter_len, int n_filters)
  {
    n_in = inputs; n_out = outputs; B = period;
    P = std::max(1, (filter_len + B - 1) / B);
    fft.reset(new RealFFT(2 * B));
    bins = fft->bins();
    window.assign((size_t)n_in * 2 * B, 0);
    fdl.assign((size_t)n_in * P * bins, cplx(0, 0));
    filters.assign((size_t)n_filters * P * bins, cplx(0, 0));
    acc.resize(bins);
    td.resize(2 * B);
    head = 0;
  }
  void set_filter(int id, const float *taps, int len)
  {
    std::vector<real> pad(2 * B);
    for (int p = 0; p < P; p++) {
      std::fill(pad.begin(), pad.end(), (real)0);
      for (int n = 0; n < B; n++) {
        int t = p * B + n;
        if (t < len) pad[n] = (real)taps[t] / (real)(2 * B);
      }
      fft->forward(pad.data(), &filters[((size_t)id * P + p) * bins]);
    }
  }
  void push_inputs(const real *const *in)
  {
    head = (head + 1) % P;
    for (int c = 0; c < n_in; c++) {
      real *w = &window[(size_t)c * 2 * B];
      std::memmove(w, w + B, sizeof(real) * B);
      std::memcpy(w + B, in[c], sizeof(real) * B);
      fft->forward(w, &fdl[((size_t)c * P + head) * bins]);
    }
  }
  // y (B samples) = sum over routings to `out` of filter[map(r.filter)] applied to input r.in
  template <class Map>
  void render_output(const std::vector<Routing> &routings, int out, Map filter_row, real *y)
  {
    std::fill(acc.begin(), acc.end(), cplx(0, 0));
    for (const Routing &r : routings) {
      if (r.out != out) continue;
      const cplx *H = &filters[(size_t)filter_row(r.filter) * P * bins];
      for (int p = 0; p < P; p++) {
        const cplx *X = &fdl[((size_t)r.in * P + (head - p + P) % P) * bins];
        const cplx *Hp = H + (size_t)p * bins;
        for (int k = 0; k < bins; k++) acc[k] += X[k] * Hp[k] * r.gain;
      }
    }
    fft->inverse(acc.data(), td.data());
    std::memcpy(y, td.data() + B, sizeof(real) * B);
  }
};

// rcl::FirFilterMatrix (dsp.cpp:61-72 decorrelators, :98-110 hoa_irs): static filters and routings.
struct FirFilterMatrix {
  UpolsCore core;
  std::vector<Routing> routings;
  void process(const real *const *in, real *const *out)
  {
    core.push_inputs(in);
    for (int o = 0; o < core.n_out; o++) core.render_output(routings, o, [](int f) { return f; }, out[o]);
  }
};

// rcl::InterpolatingFirFilterMatrix with numberOfInterpolants == 1 and transitionSamples == period
// (dsp.cpp:73-87): every filter slot holds one row of the filter table with weight 1.0.  When new
// interpolants arrive, the block is rendered with the old and the new rows over the same FDL and
// the two time-domain results are crossfaded linearly across the block.
struct InterpolatingFirFilterMatrix {
  UpolsCore core;
  std::vector<Routing> routings;
  std::vector<int> slot_row, pending_row;
  bool pending = false;
  int off = 1;
  std::vector<real> y_old;
  void set_interpolants(const std::vector<int> &rows)
  {
    pending_row = rows;
    pending = true;
  }
  void process(const real *const *in, real *const *out)
  {
    core.push_inputs(in);
    const int B = core.B;
    for (int o = 0; o < core.n_out; o++) {
      if (!pending) {
        core.render_output(routings, o, [&](int f) { return slot_row[f]; }, out[o]);
      } else {
        y_old.resize(B);
        core.render_output(routings, o, [&](int f) { return slot_row[f]; }, y_old.data());
        core.render_output(routings, o, [&](int f) { return pending_row[f]; }, out[o]);
        for (int n = 0; n < B; n++) {
          real r = ramp_value(n, B, off);
          out[o][n] = ((real)1 - r) * y_old[n] + r * out[o][n];
        }
      }
    }
    if (pending) {
      slot_row = pending_row;
      pending = false;
    }
  }
};

struct ChannelBlock {
  int ch = 0, B = 0;
  std::vector<real> data;
  std::vector<real *> ptr;
  std::vector<const real *> cptr;
  void setup(int channels, int period)
  {
    ch = channels; B = period;
    data.assign((size_t)ch * B, 0);
    ptr.resize(ch); cptr.resize(ch);
    for (int c = 0; c < ch; c++) cptr[c] = ptr[c] = &data[(size_t)c * B];
  }
};

}  // namespace

extern "C" {

struct bear_ref_setup {
  int32_t num_objects, num_direct_speakers, num_hoa, period;
  double fs;
  int32_t num_views, num_speakers, brir_length, decorrelator_length, num_hoa_decoder_channels, hoa_ir_length;
  const float *brirs;          // [V][L][2][T]   (panner.cpp:258-263, C order)
  const float *decorrelators;  // [L][Td]        (panner.cpp:250-255)
  const float *hoa_irs;        // [A][2][Th]     (panner.cpp:408-412)
  double default_direct_delay;  // seconds, panner.cpp:172-178
  double default_static_delay;  // seconds, panner.cpp:180-186
  double hoa_delay;             // seconds, panner.hpp:106
  int32_t max_delay_samples;
  int32_t gain_ramp_offset, delay_ramp_offset, xfade_ramp_offset;  // ramp[n] = (n+offset)/B
};

}  // extern "C"

namespace {

// bear::PerEarDelay (per_ear_delay.cpp:4-57): channel i fans out to delay lines 2i (left) and
// 2i+1 (right); ear e's delayed channels feed GainMatrix e; both ears share one gain table;
// output channel 2l+e.
struct PerEarDelay {
  int n = 0, L = 0, B = 0;
  DelayVector delays;
  GainMatrix gains[2];
  ChannelBlock fan, delayed, ear_out[2];
  std::vector<const real *> ear_in[2];
  void setup(int inputs, int speakers, int period, double fs, double initial_delay, int max_delay,
             int gain_off, int delay_off)
  {
    n = inputs; L = speakers; B = period;
    delays.setup(2 * n, B, fs, initial_delay, max_delay, delay_off);
    for (int e = 0; e < 2; e++) {
      gains[e].setup(n, L, B, gain_off);
      ear_out[e].setup(L, B);
      ear_in[e].resize(n);
    }
    delayed.setup(2 * n, B);
    fan.ptr.resize(2 * n);
  }
  // out: 2L channel pointers (index 2l+e), overwritten
  void process(const real *const *in, const float *gain_table, const float *delay_table, real *const *out)
  {
    if (n == 0) {
      for (int c = 0; c < 2 * L; c++) std::fill(out[c], out[c] + B, (real)0);
      return;
    }
    std::vector<const real *> fan_in(2 * n);
    for (int i = 0; i < n; i++) fan_in[2 * i] = fan_in[2 * i + 1] = in[i];
    delays.process(delay_table, fan_in.data(), delayed.ptr.data());
    for (int e = 0; e < 2; e++) {
      for (int i = 0; i < n; i++) ear_in[e][i] = delayed.ptr[2 * i + e];
      std::vector<real *> o(L);
      for (int l = 0; l < L; l++) o[l] = out[2 * l + e];
      gains[e].process(gain_table, ear_in[e].data(), o.data());
    }
  }
};

// bear::DSP (dsp.cpp:7-206)
struct Dsp {
  bear_ref_setup s;
  int N, M, H, B, L, A;
  PerEarDelay objects_direct_path, direct_speakers_path;
  GainMatrix diffuse_gains, hoa_matrix;
  FirFilterMatrix decorrelators, hoa_irs;
  InterpolatingFirFilterMatrix brirs;
  DelayVector static_delays, hoa_delays;
  ChannelBlock in_obj, in_ds, in_hoa, bus_obj, bus_ds, bus_dif, dif, dec, bus, brir_out, ambi, hoa_out, hoa_del;

  explicit Dsp(const bear_ref_setup &cfg) : s(cfg)
  {
    N = s.num_objects; M = s.num_direct_speakers; H = s.num_hoa; B = s.period;
    L = s.num_speakers; A = s.num_hoa_decoder_channels;
    const int md = s.max_delay_samples;
    objects_direct_path.setup(N, L, B, s.fs, s.default_direct_delay, md, s.gain_ramp_offset, s.delay_ramp_offset);
    direct_speakers_path.setup(M, L, B, s.fs, s.default_direct_delay, md, s.gain_ramp_offset, s.delay_ramp_offset);
    diffuse_gains.setup(N, L, B, s.gain_ramp_offset);

    decorrelators.core.setup(L, L, B, s.decorrelator_length, L);
    for (int l = 0; l < L; l++) {
      decorrelators.core.set_filter(l, s.decorrelators + (size_t)l * s.decorrelator_length, s.decorrelator_length);
      decorrelators.routings.push_back({l, l, l, (real)1});
    }
    static_delays.setup(2 * L, B, s.fs, s.default_static_delay, md, s.delay_ramp_offset);

    // initial_brir_filters / initial_brir_interpolants / initial_brir_routings, dsp.cpp:208-236
    const int rows = s.num_views * 2 * L;
    brirs.core.setup(2 * L, 2, B, s.brir_length, rows);
    for (int r = 0; r < rows; r++) brirs.core.set_filter(r, s.brirs + (size_t)r * s.brir_length, s.brir_length);
    brirs.off = s.xfade_ramp_offset;
    brirs.slot_row.resize(2 * L);
    for (int l = 0; l < L; l++)
      for (int e = 0; e < 2; e++) {
        brirs.slot_row[2 * l + e] = (0 * L + l) * 2 + e;
        brirs.routings.push_back({2 * l + e, e, 2 * l + e, (real)1});
      }

    hoa_matrix.setup(H, A, B, s.gain_ramp_offset);
    hoa_irs.core.setup(A, 2, B, s.hoa_ir_length, 2 * A);
    for (int a = 0; a < A; a++)
      for (int e = 0; e < 2; e++) {
        hoa_irs.core.set_filter(2 * a + e, s.hoa_irs + ((size_t)a * 2 + e) * s.hoa_ir_length, s.hoa_ir_length);
        hoa_irs.routings.push_back({a, e, 2 * a + e, (real)1});
      }
    hoa_delays.setup(2, B, s.fs, s.hoa_delay, md, s.delay_ramp_offset);

    in_obj.setup(N, B); in_ds.setup(M, B); in_hoa.setup(H, B);
    bus_obj.setup(2 * L, B); bus_ds.setup(2 * L, B); bus_dif.setup(2 * L, B);
    dif.setup(L, B); dec.setup(L, B); bus.setup(2 * L, B); brir_out.setup(2, B);
    ambi.setup(A, B); hoa_out.setup(2, B); hoa_del.setup(2, B);
  }

  static void load(ChannelBlock &dst, const float *src)
  {
    for (size_t i = 0; i < dst.data.size(); i++) dst.data[i] = (real)src[i];
  }

  void process(const float *obj, const float *ds, const float *hoa, const float *direct_gains,
               const float *diffuse_gains_t, const float *direct_delays, const float *ds_gains,
               const float *ds_delays, const float *static_delays_t, int brir_index, const float *hoa_gains,
               real *out)
  {
    if (N) load(in_obj, obj);
    if (M) load(in_ds, ds);
    if (H) load(in_hoa, hoa);

    // BRIRInterpolationController::process, brir_interpolation_controller.cpp:18-28
    if (brir_index >= 0) {
      std::vector<int> rows(2 * L);
      for (int l = 0; l < L; l++)
        for (int e = 0; e < 2; e++) rows[2 * l + e] = (brir_index * L + l) * 2 + e;
      brirs.set_interpolants(rows);
    }

    objects_direct_path.process(in_obj.cptr.data(), direct_gains, direct_delays, bus_obj.ptr.data());
    direct_speakers_path.process(in_ds.cptr.data(), ds_gains, ds_delays, bus_ds.ptr.data());

    // diffuse path, dsp.cpp:127-158
    diffuse_gains.process(diffuse_gains_t, in_obj.cptr.data(), dif.ptr.data());
    decorrelators.process(dif.cptr.data(), dec.ptr.data());
    std::vector<const real *> fan(2 * L);
    for (int l = 0; l < L; l++) fan[2 * l] = fan[2 * l + 1] = dec.ptr[l];
    static_delays.process(static_delays_t, fan.data(), bus_dif.ptr.data());

    // add_brir_inputs, dsp.cpp:55-59
    for (size_t i = 0; i < bus.data.size(); i++) bus.data[i] = bus_obj.data[i] + bus_dif.data[i] + bus_ds.data[i];

    brirs.process(bus.cptr.data(), brir_out.ptr.data());

    // HOA, dsp.cpp:178-205
    hoa_matrix.process(hoa_gains, in_hoa.cptr.data(), ambi.ptr.data());
    hoa_irs.process(ambi.cptr.data(), hoa_out.ptr.data());
    hoa_delays.process(nullptr, hoa_out.cptr.data(), hoa_del.ptr.data());

    for (int e = 0; e < 2; e++)
      for (int n = 0; n < B; n++) out[(size_t)e * B + n] = brir_out.ptr[e][n] + hoa_del.ptr[e][n];
  }
};

}  // namespace

extern "C" {

int bear_ref_real_size(void) { return (int)sizeof(real); }

void *bear_ref_create(const bear_ref_setup *s) { return new Dsp(*s); }
void bear_ref_destroy(void *h) { delete static_cast<Dsp *>(h); }

// One period.  obj [N][B], ds [M][B], hoa [H][B] float32; tables as the Control->DSP ports
// (dsp.cpp:30-53,89-96): direct/diffuse gains [L][N], direct delays [2N] s, ds gains [L][M], ds delays
// [2M] s, static delays [2L] s or NULL (unchanged), brir_index or -1 (no listener change), hoa gains
// [A][H] or NULL.  out [2][B] of `real`.
void bear_ref_process(void *h, const float *obj, const float *ds, const float *hoa, const float *direct_gains,
                      const float *diffuse_gains, const float *direct_delays, const float *ds_gains,
                      const float *ds_delays, const float *static_delays, int brir_index, const float *hoa_gains,
                      real *out)
{
  static_cast<Dsp *>(h)->process(obj, ds, hoa, direct_gains, diffuse_gains, direct_delays, ds_gains, ds_delays,
                                 static_delays, brir_index, hoa_gains, out);
}

// ---- component-level entry points for the oracle's own cross-checks (tests/test_oracle_*.py) ----

// Partitioned convolution of x (len samples, multiple of period) with one filter; y same length.
void bear_ref_upols(const float *x, int len, const float *taps, int ntaps, int period, real *y)
{
  FirFilterMatrix f;
  f.core.setup(1, 1, period, ntaps, 1);
  f.core.set_filter(0, taps, ntaps);
  f.routings.push_back({0, 0, 0, (real)1});
  std::vector<real> in(period);
  for (int b = 0; b + period <= len; b += period) {
    for (int n = 0; n < period; n++) in[n] = (real)x[b + n];
    const real *ip = in.data();
    real *op = y + b;
    f.process(&ip, &op);
  }
}

// Delay line: x (len samples), per-block delay targets in seconds [len/period]; y same length.
void bear_ref_delay(const float *x, int len, const float *delay_seconds, int period, double fs,
                    double initial_seconds, int max_delay_samples, int ramp_offset, real *y)
{
  DelayVector d;
  d.setup(1, period, fs, initial_seconds, max_delay_samples, ramp_offset);
  std::vector<real> in(period);
  for (int b = 0, k = 0; b + period <= len; b += period, k++) {
    for (int n = 0; n < period; n++) in[n] = (real)x[b + n];
    const real *ip = in.data();
    real *op = y + b;
    d.process(delay_seconds + k, &ip, &op);
  }
}

void bear_ref_rfft(const float *x, int n, real *out_interleaved)
{
  RealFFT f(n);
  std::vector<real> xr(x, x + n);
  f.forward(xr.data(), reinterpret_cast<cplx *>(out_interleaved));
}

}  // extern "C"

// ================================================================================================
// Control tail: the table producers between the gain calculators and the DSP ports, restated from
// the in-tree reference source (all control math in double, float at the ports):
//   DirectDelayCalc::process          direct_delay_calc.cpp:25-42   -> Panner::get_delays panner.cpp:189-208
//   GainNorm::process                 gain_norm.cpp:26-44           -> Panner::compensation_gain :277-316,:332-346,:364-380
//   DirectDiffuseSplit::process       direct_diffuse_split.cpp:24-32
//   DirectSpeakersDelayCalc::process  direct_speakers_delay_calc.cpp:24-40
//   DirectSpeakersGainNorm::process   direct_speakers_gain_norm.cpp:19-32 -> compensation_gain_direct :318-330,:348-362,:391-404
//   StaticDelayCalc::process          static_delay_calc.cpp:19-30   -> Panner::get_diffuse_delays :224-235
// ================================================================================================
namespace {
struct ControlTables {
  int L, V, S;
  double fs, dec_delay;
  int front;
  const float *delays;   // [V][L][2]
  const float *factors;  // [S][V][2L][2L][2] or null
  double delay(int v, int l, int e) const { return (double)delays[((size_t)v * L + l) * 2 + e]; }
  double factor(int s, int v, int i, int j, int e) const
  {
    const size_t G = 2 * (size_t)L;
    return (double)factors[((((size_t)s * V + v) * G + i) * G + j) * 2 + e];
  }
  // Panner::get_delays
  void mean_delays(const double *g, int v, double out[2]) const
  {
    double total = 0;
    for (int l = 0; l < L; l++) total += g[l];
    double m[2];
    if (total > 1e-6) {
      double tl = 0, tr = 0;
      for (int l = 0; l < L; l++) {
        tl += delay(v, l, 0) * g[l];
        tr += delay(v, l, 1) * g[l];
      }
      m[0] = tl / total; m[1] = tr / total;
    } else {
      m[0] = delay(v, front, 0); m[1] = delay(v, front, 1);
    }
    out[0] = (m[0] + dec_delay) / fs;
    out[1] = (m[1] + dec_delay) / fs;
  }
  // Panner::compensation_gain (gains: 2L values, direct then diffuse; delays in seconds)
  double compensation_gain(const double *g, const double dly[2], int v) const
  {
    if (!factors) return 1.0;
    const int G = 2 * L;
    double expected = 0;
    for (int e = 0; e < 2; e++)
      for (int j = 0; j < G; j++) expected += g[j] * g[j] * factor(0, v, j, j, e);
    expected = std::sqrt(expected);
    double sum = 0;
    for (int e = 0; e < 2; e++) {
      double unclamped = dly[e] * fs - dec_delay;
      double d = std::min(std::max(0.0, unclamped), (double)S - 1);
      size_t di = std::min((size_t)std::floor(d), (size_t)S - 2);
      for (size_t si = 0; si < 2; si++) {
        size_t s = di + si;
        double p = 1 - std::abs(d - (double)s);
        for (int i = 0; i < G; i++) {
          double i_sum = 0;
          for (int j = 0; j < G; j++) i_sum += g[j] * factor((int)s, v, i, j, e);
          sum += p * g[i] * i_sum;
        }
      }
    }
    double real = std::sqrt(sum);
    if (std::abs(real) < 1e-6) return 1.0;
    return expected / real;
  }
  // Panner::compensation_gain_direct (gains: L values)
  double compensation_gain_direct(const double *g, int v) const
  {
    if (!factors) return 1.0;
    double expected = 0, sum = 0;
    for (int e = 0; e < 2; e++)
      for (int j = 0; j < L; j++) expected += g[j] * g[j] * factor(0, v, j, j, e);
    for (int e = 0; e < 2; e++)
      for (int i = 0; i < L; i++) {
        double i_sum = 0;
        for (int j = 0; j < L; j++) i_sum += g[j] * factor(0, v, i, j, e);
        sum += g[i] * i_sum;
      }
    double real = std::sqrt(sum);
    if (std::abs(real) < 1e-6) return 1.0;
    return std::sqrt(expected) / real;
  }
};
}  // namespace

extern "C" void bear_ref_control(int L, int V, int S, double fs, double decorrelation_delay, int front_loudspeaker,
                                 const float *delays, const float *factors, int view, int N, const float *gains /*[2L][N]*/,
                                 int M, const float *ds_gains /*[L][M]*/, float *direct_gains /*[L][N]*/,
                                 float *diffuse_gains /*[L][N]*/, float *direct_delays /*[2N]*/,
                                 float *ds_gains_out /*[L][M]*/, float *ds_delays /*[2M]*/, float *static_delays /*[2L]*/)
{
  ControlTables c{L, V, S, fs, decorrelation_delay, front_loudspeaker, delays, S > 0 ? factors : nullptr};
  std::vector<double> g(2 * L), tot(L);
  for (int i = 0; i < N; i++) {
    for (int l = 0; l < 2 * L; l++) g[l] = (double)gains[(size_t)l * N + i];
    for (int l = 0; l < L; l++) tot[l] = g[l] + g[L + l];
    double d[2];
    c.mean_delays(tot.data(), view, d);
    direct_delays[2 * i] = (float)d[0];
    direct_delays[2 * i + 1] = (float)d[1];
    // GainNorm reads the delays back from the float port (gain_norm.cpp:36-37)
    double df[2] = {(double)direct_delays[2 * i], (double)direct_delays[2 * i + 1]};
    double comp = c.compensation_gain(g.data(), df, view);
    for (int l = 0; l < L; l++) {
      float gd = (float)(comp * (double)gains[(size_t)l * N + i]);
      float gq = (float)(comp * (double)gains[(size_t)(L + l) * N + i]);
      direct_gains[(size_t)l * N + i] = gd;
      diffuse_gains[(size_t)l * N + i] = gq;
    }
  }
  for (int j = 0; j < M; j++) {
    for (int l = 0; l < L; l++) g[l] = (double)ds_gains[(size_t)l * M + j];
    double d[2];
    c.mean_delays(g.data(), view, d);
    ds_delays[2 * j] = (float)d[0];
    ds_delays[2 * j + 1] = (float)d[1];
    double comp = c.compensation_gain_direct(g.data(), view);
    for (int l = 0; l < L; l++) ds_gains_out[(size_t)l * M + j] = (float)(comp * (double)ds_gains[(size_t)l * M + j]);
  }
  for (int l = 0; l < L; l++)
    for (int e = 0; e < 2; e++) static_delays[2 * l + e] = (float)(c.delay(view, l, e) / fs);
}
