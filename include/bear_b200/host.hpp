// bear_b200/host.hpp -- host-side C++ layer over the C ABI (include/bear_cuda.h).
//
// the part of `bear::Renderer` (visr_bear/include/bear/api.hpp:155-256,
// visr_bear/src/api.cpp:117-234) and of the host Control graph that does not need libear or VISR:
//   * Config / Listener / Time with the reference's names and error messages,
//   * the metadata timing rules of add_*_block (api.cpp:156-196) and block_start_time (:198-210),
//   * per-channel gain scheduling: GainCalcObjects::PerObject (gain_calc_objects.cpp:60-120),
//     DirectSpeakersGainCalc::PerObject (direct_speakers_gain_calc.cpp:41-71), GainCalcHOA
//     (gain_calc_hoa.cpp:49-124) including the SH rotation of the sound field (sh_rotation.cpp),
//   * BRIR view selection with the residual listener (select_brir.cpp:12-53).
// What it cannot contain is the ADM-metadata -> gain-vector step (ear::GainCalculatorObjects /
// DirectSpeakers, libear, needs Boost): blocks therefore carry the GAIN VECTORS that
// `Panner::calc_objects_gains` / `calc_direct_speakers_gains` (panner.cpp:139-155,237-247) return for
// their type metadata, evaluated for the residual listener.  Everything downstream of those vectors
// -- interpolation, delays, gain normalisation, the whole DSP -- is done here / on the device.
//
// Header-only, C++17, no dependencies beyond the C ABI.
#pragma once

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../bear_cuda.h"

namespace bear_b200 {

using Sample = float;

// ---------------------------------------------------------------------------------------------
// Time in seconds as an exact rational (stands in for boost::rational<int64_t>, api.hpp:63-65)
// ---------------------------------------------------------------------------------------------
class Time {
 public:
  constexpr Time() = default;
  Time(int64_t n, int64_t d = 1) : n_(n), d_(d)
  {
    if (d_ == 0) throw std::domain_error("Time: zero denominator");
    if (d_ < 0) { n_ = -n_; d_ = -d_; }
    int64_t g = gcd(n_ < 0 ? -n_ : n_, d_);
    if (g > 1) { n_ /= g; d_ /= g; }
  }
  int64_t numerator() const { return n_; }
  int64_t denominator() const { return d_; }
  double to_double() const { return (double)n_ / (double)d_; }
  explicit operator bool() const { return n_ != 0; }

  friend Time operator+(const Time &a, const Time &b) { return from128((i128)a.n_ * b.d_ + (i128)b.n_ * a.d_, (i128)a.d_ * b.d_); }
  friend Time operator-(const Time &a, const Time &b) { return from128((i128)a.n_ * b.d_ - (i128)b.n_ * a.d_, (i128)a.d_ * b.d_); }
  friend Time operator*(const Time &a, const Time &b) { return from128((i128)a.n_ * b.n_, (i128)a.d_ * b.d_); }
  friend Time operator/(const Time &a, const Time &b)
  {
    if (b.n_ == 0) throw std::domain_error("Time: division by zero");
    return from128((i128)a.n_ * b.d_, (i128)a.d_ * b.n_);
  }
  Time &operator+=(const Time &b) { return *this = *this + b; }
  friend bool operator==(const Time &a, const Time &b) { return a.n_ == b.n_ && a.d_ == b.d_; }
  friend bool operator!=(const Time &a, const Time &b) { return !(a == b); }
  friend bool operator<(const Time &a, const Time &b) { return (i128)a.n_ * b.d_ < (i128)b.n_ * a.d_; }
  friend bool operator>(const Time &a, const Time &b) { return b < a; }
  friend bool operator<=(const Time &a, const Time &b) { return !(b < a); }
  friend bool operator>=(const Time &a, const Time &b) { return !(a < b); }

 private:
  using i128 = __int128;
  static int64_t gcd(int64_t a, int64_t b) { while (b) { int64_t t = a % b; a = b; b = t; } return a ? a : 1; }
  static Time from128(i128 n, i128 d)
  {
    if (d < 0) { n = -n; d = -d; }
    i128 a = n < 0 ? -n : n, b = d;
    while (b) { i128 t = a % b; a = b; b = t; }
    if (a > 1) { n /= a; d /= a; }
    Time t;
    t.n_ = (int64_t)n; t.d_ = (int64_t)d;
    return t;
  }
  int64_t n_ = 0, d_ = 1;
};

// ---------------------------------------------------------------------------------------------
// Config: same setters/getters and validation messages as bear::Config (api.hpp:16-61, api.cpp:41-72)
// ---------------------------------------------------------------------------------------------
class Config {
 public:
  void set_num_objects_channels(size_t n) { num_objects_ = n; }
  size_t get_num_objects_channels() const { return num_objects_; }
  void set_num_direct_speakers_channels(size_t n) { num_ds_ = n; }
  size_t get_num_direct_speakers_channels() const { return num_ds_; }
  void set_num_hoa_channels(size_t n) { num_hoa_ = n; }
  size_t get_num_hoa_channels() const { return num_hoa_; }
  void set_period_size(size_t n) { period_ = n; }
  size_t get_period_size() const { return period_; }
  void set_sample_rate(size_t fs) { sample_rate_ = fs; }
  size_t get_sample_rate() const { return sample_rate_; }
  void set_data_path(const std::string &p) { data_path_ = p; }
  const std::string &get_data_path() const { return data_path_; }
  // kept for source compatibility; there is one FFT here, "default" (a different name fails like
  // the reference does for an unknown wrapper, test_renderer.cpp:84-94)
  void set_fft_implementation(const std::string &f) { fft_ = f; }
  const std::string &get_fft_implementation() const { return fft_; }
  void set_device(int d) { device_ = d; }
  int get_device() const { return device_; }

  void validate() const
  {
    if (period_ == 0) throw std::invalid_argument("Config: period size must be set");
    if (data_path_.empty()) throw std::invalid_argument("Config: data path must be set");
  }

 private:
  size_t num_objects_ = 0, num_ds_ = 0, num_hoa_ = 0, period_ = 0, sample_rate_ = 48000;
  std::string data_path_, fft_ = "default";
  int device_ = 0;
};

// ---------------------------------------------------------------------------------------------
// Quaternion {w,x,y,z} helpers and Listener (api.hpp:118-153, listener_impl.hpp)
// ---------------------------------------------------------------------------------------------
struct Quat {
  double w = 1, x = 0, y = 0, z = 0;
  Quat conj() const { return {w, -x, -y, -z}; }
  Quat normalized() const
  {
    double n = std::sqrt(w * w + x * x + y * y + z * z);
    return {w / n, x / n, y / n, z / n};
  }
  friend Quat operator*(const Quat &a, const Quat &b)
  {
    return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
  }
  // q v q*  (active rotation of v)
  std::array<double, 3> rotate(const std::array<double, 3> &v) const
  {
    Quat p{0, v[0], v[1], v[2]};
    Quat r = (*this) * p * conj();
    return {r.x, r.y, r.z};
  }
};

class Listener {
 public:
  void set_position_cart(std::array<double, 3> p) { position_ = p; }
  std::array<double, 3> get_position_cart() const { return position_; }
  /// world -> listener coordinates, {w, x, y, z}, ADM Cartesian axes (x right, y front, z up)
  void set_orientation_quaternion(std::array<double, 4> q) { q_ = {q[0], q[1], q[2], q[3]}; }
  std::array<double, 4> get_orientation_quaternion() const { return {q_.w, q_.x, q_.y, q_.z}; }
  std::array<double, 3> look() const { return q_.conj().rotate({0.0, 1.0, 0.0}); }
  std::array<double, 3> right() const { return q_.conj().rotate({1.0, 0.0, 0.0}); }
  std::array<double, 3> up() const { return q_.conj().rotate({0.0, 0.0, 1.0}); }
  const Quat &orientation() const { return q_; }

 private:
  std::array<double, 3> position_{0, 0, 0};
  Quat q_{};
};

// ---------------------------------------------------------------------------------------------
// SelectBRIR (select_brir.cpp:12-53): nearest view by maximum dot product with the look vector, and
// the residual listener with that view's yaw rotation removed.
// ---------------------------------------------------------------------------------------------
class ViewSelector {
 public:
  ViewSelector() = default;
  /// views: [V][3] float (data file "views")
  ViewSelector(const float *views, size_t V)
  {
    for (size_t v = 0; v < V; v++) {
      double x = views[3 * v], y = views[3 * v + 1], z = views[3 * v + 2];
      double n = std::sqrt(x * x + y * y + z * z);
      x /= n; y /= n; z /= n;
      if (std::abs(z) >= 1e-6) throw std::logic_error("expected BRIR rotations around z only");
      views_.push_back({x, y, z});
      double qz = std::sin(-std::atan2(x, y) / 2.0);
      rotations_.push_back({std::sqrt(1.0 - qz * qz), 0.0, 0.0, qz});
    }
  }
  size_t num_views() const { return views_.size(); }
  struct Selection { unsigned view; Listener residual; };
  Selection select(const Listener &l) const
  {
    auto look = l.look();
    unsigned best = 0;
    double best_dp = -1e300;
    for (unsigned v = 0; v < views_.size(); v++) {
      double dp = views_[v][0] * look[0] + views_[v][1] * look[1] + views_[v][2] * look[2];
      if (dp > best_dp) { best_dp = dp; best = v; }
    }
    Listener out;
    out.set_position_cart(l.get_position_cart());
    Quat r = rotations_[best] * l.orientation();
    out.set_orientation_quaternion({r.w, r.x, r.y, r.z});
    return {best, out};
  }

 private:
  std::vector<std::array<double, 3>> views_;
  std::vector<Quat> rotations_;
};

// ---------------------------------------------------------------------------------------------
// Real spherical harmonics (BS.2076 section 10, ear_bits/hoa.hpp:66-124) and the rotation of a
// sound field given in ACN order (what quaternion_to_sh_rotation_matrix, sh_rotation.cpp:140-157,
// provides).  The matrix is obtained by projection instead of the Ivanic/Ruedenberg recursion the
// reference uses: encode a fixed set of directions, encode the rotated directions, solve the
// (well-conditioned, orthonormal-basis) least-squares problem.  Both give the unique linear map
// M with M Y(p) = Y(q p) for all p; the reference's own test checks exactly this property
// (test_sh_rotation.cpp:24-50).
// ---------------------------------------------------------------------------------------------
namespace sh {
inline int to_acn(int n, int m) { return n * n + n + m; }
inline std::pair<int, int> from_acn(int acn)
{
  int n = (int)std::sqrt((double)acn);
  while ((n + 1) * (n + 1) <= acn) n++;
  while (n * n > acn) n--;
  return {n, acn - n * n - n};
}
inline double factorial(int n) { double f = 1; for (int i = 2; i <= n; i++) f *= i; return f; }
inline double norm_N3D(int n, int am) { return std::sqrt((2.0 * n + 1.0) * factorial(n - am) / factorial(n + am)); }
inline double norm_SN3D(int n, int am) { return std::sqrt(factorial(n - am) / factorial(n + am)); }
inline double norm_FuMa(int n, int am)
{
  static const double f[4][4] = {{1.0 / std::sqrt(2.0), 0, 0, 0},
                                 {1.0, 1.0, 0, 0},
                                 {1.0, 2.0 / std::sqrt(3.0), 2.0 / std::sqrt(3.0), 0},
                                 {1.0, std::sqrt(45.0 / 32.0), 3.0 / std::sqrt(5.0), std::sqrt(8.0 / 5.0)}};
  if (n > 3) throw std::invalid_argument("FuMa normalisation is defined up to order 3");
  return f[n][am] * norm_SN3D(n, am);
}
using norm_fn = double (*)(int, int);
inline norm_fn norm_by_name(const std::string &name)
{
  if (name == "N3D") return norm_N3D;
  if (name == "SN3D") return norm_SN3D;
  if (name == "FuMa") return norm_FuMa;
  throw std::invalid_argument("unknown normalization type: '" + name + "'");
}
/// associated Legendre function P_n^m(x) without the Condon-Shortley phase
inline double legendre(int n, int m, double x)
{
  double pmm = 1.0;
  if (m > 0) {
    double s = std::sqrt(std::max(0.0, 1.0 - x * x));
    double f = 1.0;
    for (int i = 1; i <= m; i++) { pmm *= f * s; f += 2.0; }
  }
  if (n == m) return pmm;
  double pm1 = x * (2.0 * m + 1.0) * pmm;
  if (n == m + 1) return pm1;
  double p = 0;
  for (int l = m + 2; l <= n; l++) {
    p = ((2.0 * l - 1.0) * x * pm1 - (l + m - 1.0) * pmm) / (l - m);
    pmm = pm1; pm1 = p;
  }
  return p;
}
/// az, el: ADM azimuth / elevation in radians
inline double sph_harm(int n, int m, double az, double el, norm_fn norm)
{
  double scale = m > 0 ? std::sqrt(2.0) * std::cos(m * az) : (m < 0 ? -std::sqrt(2.0) * std::sin(m * az) : 1.0);
  return norm(n, std::abs(m)) * legendre(n, std::abs(m), std::sin(el)) * scale;
}
/// encode the ADM Cartesian direction pos (x right, y front, z up)
inline std::vector<double> encode(int order, const std::array<double, 3> &pos, norm_fn norm = norm_N3D)
{
  double az = -std::atan2(pos[0], pos[1]);
  double el = std::atan2(pos[2], std::hypot(pos[0], pos[1]));
  int nch = (order + 1) * (order + 1);
  std::vector<double> out(nch);
  for (int acn = 0; acn < nch; acn++) {
    auto nm = from_acn(acn);
    out[acn] = sph_harm(nm.first, nm.second, az, el, norm);
  }
  return out;
}
/// row-major [(order+1)^2][(order+1)^2] matrix M with M encode(p) == encode(q p q*)
inline std::vector<double> rotation_matrix(const Quat &q_in, int order)
{
  if (order <= 0) throw std::invalid_argument("order must be +ve");
  const Quat q = q_in.normalized();
  const int nch = (order + 1) * (order + 1);
  const int npts = 4 * nch + 8;  // spherical Fibonacci points
  // accumulate per order: A = sum Y_l(p) Y_l(p)^T, C = sum Y_l(q p) Y_l(p)^T, M_l = C A^-1
  std::vector<double> M((size_t)nch * nch, 0.0);
  M[0] = 1.0;
  std::vector<std::vector<double>> Y(npts), Yr(npts);
  const double golden = 3.14159265358979323846 * (3.0 - std::sqrt(5.0));
  for (int i = 0; i < npts; i++) {
    double z = 1.0 - 2.0 * (i + 0.5) / npts, r = std::sqrt(std::max(0.0, 1.0 - z * z)), a = golden * i;
    std::array<double, 3> p{r * std::cos(a), r * std::sin(a), z};
    Y[i] = encode(order, p);
    Yr[i] = encode(order, q.rotate(p));
  }
  for (int l = 1; l <= order; l++) {
    const int w = 2 * l + 1, o = l * l;
    std::vector<double> A((size_t)w * w, 0.0), C((size_t)w * w, 0.0);
    for (int i = 0; i < npts; i++)
      for (int r = 0; r < w; r++)
        for (int c = 0; c < w; c++) {
          A[(size_t)r * w + c] += Y[i][o + r] * Y[i][o + c];
          C[(size_t)r * w + c] += Yr[i][o + r] * Y[i][o + c];
        }
    // solve X A = C  (A symmetric positive definite): Gauss-Jordan on [A | C^T]
    std::vector<double> aug((size_t)w * 2 * w);
    for (int r = 0; r < w; r++)
      for (int c = 0; c < w; c++) {
        aug[(size_t)r * 2 * w + c] = A[(size_t)r * w + c];
        aug[(size_t)r * 2 * w + w + c] = C[(size_t)c * w + r];
      }
    for (int col = 0; col < w; col++) {
      int piv = col;
      for (int r = col + 1; r < w; r++)
        if (std::abs(aug[(size_t)r * 2 * w + col]) > std::abs(aug[(size_t)piv * 2 * w + col])) piv = r;
      if (piv != col)
        for (int c = 0; c < 2 * w; c++) std::swap(aug[(size_t)piv * 2 * w + c], aug[(size_t)col * 2 * w + c]);
      double d = aug[(size_t)col * 2 * w + col];
      for (int c = 0; c < 2 * w; c++) aug[(size_t)col * 2 * w + c] /= d;
      for (int r = 0; r < w; r++) {
        if (r == col) continue;
        double f = aug[(size_t)r * 2 * w + col];
        if (f != 0.0)
          for (int c = 0; c < 2 * w; c++) aug[(size_t)r * 2 * w + c] -= f * aug[(size_t)col * 2 * w + c];
      }
    }
    for (int r = 0; r < w; r++)
      for (int c = 0; c < w; c++) {
        double v = aug[(size_t)c * 2 * w + w + r];   // X[r][c] = (A^-1 C^T)[c][r]
        M[(size_t)(o + r) * nch + (o + c)] = std::abs(v) < 1e-5 ? 0.0 : v;   // same thresholding as sh_rotation.cpp:132-136
      }
  }
  return M;
}
}  // namespace sh

// ---------------------------------------------------------------------------------------------
// Metadata carriers.  rtime/duration/interpolationLength have the reference's meaning
// (api.hpp:94-112,178-226).
// ---------------------------------------------------------------------------------------------
struct MetadataInput {
  std::optional<Time> rtime;
  std::optional<Time> duration;
};
/// stands in for ObjectsInput: the gain vectors Panner::calc_objects_gains returns for the block's
/// ObjectsTypeMetadata (length num_gains each), evaluated for the residual listener
struct ObjectsGains : MetadataInput {
  std::vector<double> direct, diffuse;
  std::optional<Time> interpolationLength;
};
/// stands in for DirectSpeakersInput: Panner::calc_direct_speakers_gains' vector
struct DirectSpeakersGains : MetadataInput {
  std::vector<double> gains;
};
/// HOAInput with the fields GainCalcHOA reads (gain_calc_hoa.cpp:49-84)
struct HOAInput : MetadataInput {
  std::vector<size_t> channels;
  std::vector<int> orders, degrees;
  std::string normalization = "SN3D";
};

// GainCalcObjects::PerObject (gain_calc_objects.cpp:60-120)
class ObjectSchedule {
 public:
  explicit ObjectSchedule(size_t n_gains = 0) : n_(n_gains), a_(2 * n_gains, 0.0), b_(2 * n_gains, 0.0) {}
  void update(const ObjectsGains &block)
  {
    if (block.direct.size() != n_ || block.diffuse.size() != n_)
      throw std::invalid_argument("direct gains has wrong number of rows");
    std::vector<double> g(2 * n_);
    std::copy(block.direct.begin(), block.direct.end(), g.begin());
    std::copy(block.diffuse.begin(), block.diffuse.end(), g.begin() + n_);
    if (block.rtime && block.duration) {
      Time default_len = (first_block_ || last_block_end_ != *block.rtime) ? Time(0) : *block.duration;
      Time len = block.interpolationLength ? std::min(*block.interpolationLength, default_len) : default_len;
      if (len) a_ = b_;   // a is only looked at if there is some interpolation length
      b_ = std::move(g);
      a_time_ = *block.rtime;
      b_time_ = *block.rtime + len;
      last_block_end_ = *block.rtime + *block.duration;
      first_block_ = false;
      infinite_block_ = false;
    } else {
      b_ = std::move(g);
      first_block_ = true;
      infinite_block_ = true;
    }
  }
  /// gains at time t (the end of the period being rendered): 2*n values, direct then diffuse
  void gains_at(const Time &t, double *out) const
  {
    if (infinite_block_) {
      std::copy(b_.begin(), b_.end(), out);
    } else if (t > last_block_end_) {
      std::fill(out, out + 2 * n_, 0.0);
    } else if (t >= b_time_) {
      std::copy(b_.begin(), b_.end(), out);
    } else if (t >= a_time_) {
      double p = ((t - a_time_) / (b_time_ - a_time_)).to_double();
      for (size_t i = 0; i < 2 * n_; i++) out[i] = a_[i] * (1 - p) + b_[i] * p;
    } else {
      throw std::logic_error("block shouldn't have been processed yet");
    }
  }

 private:
  size_t n_;
  std::vector<double> a_, b_;
  Time a_time_, b_time_, last_block_end_;
  // a freshly constructed PerObject renders silence until the first block arrives: infinite_block is
  // false and last_block_end is 0, so any period end t > 0 gives zeros (gain_calc_objects.hpp defaults)
  bool first_block_ = true, infinite_block_ = false;
};

// DirectSpeakersGainCalc::PerObject (direct_speakers_gain_calc.cpp:41-71)
class DirectSpeakersSchedule {
 public:
  explicit DirectSpeakersSchedule(size_t n_gains = 0) : g_(n_gains, 0.0) {}
  void update(const DirectSpeakersGains &block)
  {
    if (block.gains.size() != g_.size()) throw std::invalid_argument("gains has wrong number of rows");
    g_ = block.gains;
    if (block.rtime && block.duration) {
      last_block_end_ = *block.rtime + *block.duration;
      infinite_block_ = false;
    } else {
      infinite_block_ = true;
    }
  }
  void gains_at(const Time &t, double *out) const
  {
    if (infinite_block_ || t <= last_block_end_) std::copy(g_.begin(), g_.end(), out);
    else std::fill(out, out + g_.size(), 0.0);
  }

 private:
  std::vector<double> g_;
  Time last_block_end_;
  bool infinite_block_ = false;
};

// GainCalcHOA (gain_calc_hoa.cpp:49-124): per input channel a column of the [A x H] decode-input
// matrix = SH rotation x (normalisation conversion placed at the channel's ACN)
class HOASchedule {
 public:
  HOASchedule() = default;
  HOASchedule(size_t n_decoder_channels, size_t n_input_channels)
      : A_(n_decoder_channels), ch_(n_input_channels), matrix_(n_decoder_channels * n_input_channels, 0.0f)
  {
    order_ = (int)std::llround(std::sqrt((double)A_)) - 1;
    rotation_.assign(A_ * A_, 0.0);
    for (size_t i = 0; i < A_; i++) rotation_[i * A_ + i] = 1.0;
    for (auto &c : ch_) c.to_hoa.assign(A_, 0.0);
  }
  void set_orientation(const Quat &q)
  {
    rotation_ = sh::rotation_matrix(q, order_);
    listener_changed_ = true;
  }
  void update(size_t stream, const HOAInput &in)
  {
    if (in.channels.size() != in.orders.size()) throw std::invalid_argument("number of HOA input channels must match orders");
    if (in.channels.size() != in.degrees.size()) throw std::invalid_argument("number of HOA input channels must match degrees");
    sh::norm_fn from = sh::norm_by_name(in.normalization);
    Time block_end(0);
    bool infinite = true;
    if (in.rtime && in.duration) { block_end = *in.rtime + *in.duration; infinite = false; }
    for (size_t i = 0; i < in.channels.size(); i++) {
      int n = in.orders[i], m = in.degrees[i];
      double convert = sh::norm_SN3D(n, std::abs(m)) / from(n, std::abs(m));
      size_t acn = (size_t)sh::to_acn(n, m);
      Channel &c = ch_.at(in.channels[i]);
      c.stream = stream; c.last_block_end = block_end; c.infinite = infinite; c.changed = true;
      std::fill(c.to_hoa.begin(), c.to_hoa.end(), 0.0);
      if (acn < A_) c.to_hoa[acn] = convert;   // orders beyond the decoder are ignored
    }
  }
  /// the [A x H] row-major float matrix for the period ending at block_end
  const std::vector<float> &matrix_at(const Time &block_end)
  {
    const size_t H = ch_.size();
    for (size_t i = 0; i < H; i++) {
      Channel &c = ch_[i];
      if (!c.infinite && c.last_block_end < block_end) {
        for (size_t j = 0; j < A_; j++) matrix_[j * H + i] = 0.0f;
      } else if (c.changed || listener_changed_) {
        for (size_t j = 0; j < A_; j++) {
          double acc = 0;
          for (size_t k = 0; k < A_; k++) acc += rotation_[j * A_ + k] * c.to_hoa[k];
          matrix_[j * H + i] = (float)acc;
        }
      }
      c.changed = false;
    }
    listener_changed_ = false;
    return matrix_;
  }

 private:
  // a new channel is silent until metadata arrives (infinite_block false, last_block_end 0)
  struct Channel { std::vector<double> to_hoa; size_t stream = 0; Time last_block_end; bool infinite = false, changed = false; };
  size_t A_ = 0;
  int order_ = 0;
  std::vector<Channel> ch_;
  std::vector<double> rotation_;
  std::vector<float> matrix_;
  bool listener_changed_ = false;
};

// ---------------------------------------------------------------------------------------------
// Renderer: bear::Renderer's process/add_*_block/time/listener interface (api.hpp:155-256) over
// libbear_cuda.so.  Single-threaded like the reference (dynamic_renderer.hpp:14-15).
// ---------------------------------------------------------------------------------------------
class Renderer {
 public:
  Renderer() = default;
  explicit Renderer(const Config &config) : cfg_(config)
  {
    config.validate();
    if (config.get_fft_implementation() != "default")
      throw std::runtime_error("The specified FFT wrapper does not exist: " + config.get_fft_implementation());
    bear_cuda_config c{};
    c.num_objects_channels = (uint32_t)config.get_num_objects_channels();
    c.num_direct_speakers_channels = (uint32_t)config.get_num_direct_speakers_channels();
    c.num_hoa_channels = (uint32_t)config.get_num_hoa_channels();
    c.period_size = (uint32_t)config.get_period_size();
    c.sample_rate = (uint32_t)config.get_sample_rate();
    c.max_batch_blocks = 1;
    c.device = config.get_device();
    c.gain_ramp_offset = c.delay_ramp_offset = c.xfade_ramp_offset = 1;
    int rc = bear_cuda_create_from_file(&c, config.get_data_path().c_str(), &h_);
    if (rc != BEAR_CUDA_OK) rethrow(rc, bear_cuda_last_error(nullptr));
    bear_cuda_data_file_info info{};
    check(bear_cuda_read_data_file_info(config.get_data_path().c_str(), &info), nullptr);
    L_ = info.num_virtual_loudspeakers;
    A_ = info.num_hoa_decoder_channels;
    std::vector<float> views((size_t)info.num_views * 3);
    size_t count = 0;
    uint32_t shape[8], ndim = 0;
    check(bear_cuda_tenf_read_f32(config.get_data_path().c_str(), "views", views.data(), views.size(), &count, shape, &ndim), nullptr);
    views_ = ViewSelector(views.data(), info.num_views);
    N_ = config.get_num_objects_channels(); M_ = config.get_num_direct_speakers_channels(); H_ = config.get_num_hoa_channels();
    objects_.assign(N_, ObjectSchedule(L_));
    direct_speakers_.assign(M_, DirectSpeakersSchedule(L_));
    hoa_ = HOASchedule(A_, H_);
    raw_obj_.assign(2 * L_ * std::max<size_t>(N_, 1), 0.0f);
    raw_ds_.assign(L_ * std::max<size_t>(M_, 1), 0.0f);
    tmp_.assign(2 * L_, 0.0);
    listener_changed_ = true;   // api.cpp:135: block 0 always sees a changed (default) listener
  }
  Renderer(Renderer &&o) noexcept { *this = std::move(o); }
  Renderer &operator=(Renderer &&o) noexcept
  {
    if (this != &o) {
      if (h_) bear_cuda_destroy(h_);
      cfg_ = o.cfg_; h_ = o.h_; o.h_ = nullptr;
      L_ = o.L_; A_ = o.A_; N_ = o.N_; M_ = o.M_; H_ = o.H_;
      views_ = std::move(o.views_); objects_ = std::move(o.objects_); direct_speakers_ = std::move(o.direct_speakers_);
      hoa_ = std::move(o.hoa_); raw_obj_ = std::move(o.raw_obj_); raw_ds_ = std::move(o.raw_ds_); tmp_ = std::move(o.tmp_);
      sample_count_ = o.sample_count_; time_offset_ = o.time_offset_; listener_ = o.listener_; residual_ = o.residual_;
      listener_changed_ = o.listener_changed_; view_ = o.view_;
    }
    return *this;
  }
  Renderer(const Renderer &) = delete;
  Renderer &operator=(const Renderer &) = delete;
  ~Renderer() { if (h_) bear_cuda_destroy(h_); }

  /// process period_size samples; same buffer contract as bear::Renderer::process (api.hpp:165-177)
  void process(const Sample *const *objects_input, const Sample *const *direct_speakers_input,
               const Sample *const *hoa_input, Sample *const *output)
  {
    const size_t B = cfg_.get_period_size(), fs = cfg_.get_sample_rate();
    if (listener_changed_) {   // SelectBRIR::process
      auto sel = views_.select(listener_);
      view_ = (int32_t)sel.view;
      residual_ = sel.residual;
      hoa_.set_orientation(listener_.orientation());   // GainCalcHOA sees the full listener (control.cpp)
      listener_changed_ = false;
    }
    const Time block_end((int64_t)(sample_count_ + B), (int64_t)fs);
    for (size_t i = 0; i < N_; i++) {   // GainCalcObjects::process
      objects_[i].gains_at(block_end, tmp_.data());
      for (size_t l = 0; l < 2 * L_; l++) raw_obj_[l * N_ + i] = (float)tmp_[l];
    }
    for (size_t j = 0; j < M_; j++) {   // DirectSpeakersGainCalc::process
      direct_speakers_[j].gains_at(block_end, tmp_.data());
      for (size_t l = 0; l < L_; l++) raw_ds_[l * M_ + j] = (float)tmp_[l];
    }
    const float *hoa_matrix = H_ ? hoa_.matrix_at(block_end).data() : nullptr;
    check(bear_cuda_control_tables(h_, 1, raw_obj_.data(), raw_ds_.data(), &view_, hoa_matrix), h_);
    check(bear_cuda_process(h_, objects_input, direct_speakers_input, hoa_input, nullptr, output), h_);
    sample_count_ += B;
  }

  bool add_objects_block(size_t channel, ObjectsGains metadata)
  {
    if (channel >= N_) throw std::invalid_argument("channel number out of range in add_objects_block");
    if (!accept(metadata)) return false;
    objects_[channel].update(metadata);
    return true;
  }
  bool add_direct_speakers_block(size_t channel, DirectSpeakersGains metadata)
  {
    if (channel >= M_) throw std::invalid_argument("channel number out of range in add_direct_speakers_block");
    if (!accept(metadata)) return false;
    direct_speakers_[channel].update(metadata);
    return true;
  }
  bool add_hoa_block(size_t stream, HOAInput metadata)
  {
    if (!accept(metadata)) return false;
    hoa_.update(stream, metadata);
    return true;
  }

  /// start time of the first sample in the next period
  Time get_block_start_time() const { return raw_block_start() - time_offset_; }
  void set_block_start_time(const Time &t) { time_offset_ = raw_block_start() - t; }
  /// interpolation_time is accepted and ignored, as in the reference (api.cpp:212-221)
  void set_listener(const Listener &l, const std::optional<Time> & = {})
  {
    listener_ = l;
    listener_changed_ = true;
  }
  /// the listener the gain calculators should use: position kept, the selected view's yaw removed
  /// (SelectBRIR's listener_out).  Valid after the period in which the listener was set.
  const Listener &residual_listener() const { return residual_; }
  int32_t selected_view() const { return view_; }
  size_t num_gains() const { return L_; }
  bear_cuda_renderer *handle() { return h_; }

 private:
  Time raw_block_start() const { return Time((int64_t)sample_count_, (int64_t)cfg_.get_sample_rate()); }
  Time raw_next_block_start() const
  {
    return Time((int64_t)(sample_count_ + cfg_.get_period_size()), (int64_t)cfg_.get_sample_rate());
  }
  // api.cpp:161-168: the offset is applied when the block is pushed; a block is taken only if it has
  // no rtime or starts before the end of the next period, otherwise the caller retries later
  bool accept(MetadataInput &m) const
  {
    if (m.rtime) *m.rtime += time_offset_;
    return !m.rtime || *m.rtime < raw_next_block_start();
  }
  static void rethrow(int rc, const char *msg)
  {
    if (rc == BEAR_CUDA_ERR_INVALID_ARGUMENT) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
  }
  static void check(int rc, bear_cuda_renderer *h)
  {
    if (rc != BEAR_CUDA_OK) rethrow(rc, bear_cuda_last_error(h));
  }

  Config cfg_;
  bear_cuda_renderer *h_ = nullptr;
  size_t L_ = 0, A_ = 0, N_ = 0, M_ = 0, H_ = 0;
  ViewSelector views_;
  std::vector<ObjectSchedule> objects_;
  std::vector<DirectSpeakersSchedule> direct_speakers_;
  HOASchedule hoa_;
  std::vector<float> raw_obj_, raw_ds_;
  std::vector<double> tmp_;
  uint64_t sample_count_ = 0;
  Time time_offset_;
  Listener listener_, residual_;
  bool listener_changed_ = true;
  int32_t view_ = 0;
};

}  // namespace bear_b200
