// Tests of include/bear_b200/host.hpp.  Run without arguments: host-only checks (no GPU).
// Run as `test_host gpu <data file>`: additionally drives a Renderer on cuda:0.
// The cases restate the reference's tests for the same behaviour:
//   Config validation            visr_bear/test/test_config.py, api.cpp:68-72
//   listener axis conventions    visr_bear/test/test_select_brir.py:84-120
//   view selection + residual    visr_bear/test/test_select_brir.py:52-76
//   object block timing          visr_bear/test/test_objects_gain_calc.py:82-131, api.hpp:178-215
//   SH rotation                  visr_bear/test/test_sh_rotation.cpp:24-50
//   renderer time API / errors   visr_bear/test/test_renderer.py:41-97, test_renderer.cpp:7-94
#include <bear_b200/host.hpp>

#include <cstdio>
#include <cstdlib>
#include <functional>
#include <random>

using namespace bear_b200;

static int failures = 0;
#define CHECK(cond)                                                                   \
  do {                                                                                \
    if (!(cond)) { std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); failures++; } \
  } while (0)
static bool close(double a, double b, double tol = 1e-6) { return std::abs(a - b) <= tol; }
template <class E, class F>
static bool throws(F &&f, const char *needle = nullptr)
{
  try { f(); } catch (const E &e) { return !needle || std::string(e.what()).find(needle) != std::string::npos; } catch (...) { return false; }
  return false;
}

static void test_time()
{
  CHECK(Time(512, 48000) == Time(4, 375));
  CHECK(Time(1, 2) + Time(1, 3) == Time(5, 6));
  CHECK(Time(1, 2) - Time(1, 2) == Time(0));
  CHECK(!(Time(0)));
  CHECK(Time(1, 3) < Time(1, 2) && Time(1, 2) > Time(1, 3) && Time(2, 4) <= Time(1, 2));
  CHECK(close(((Time(150, 1) - Time(100, 1)) / (Time(200, 1) - Time(100, 1))).to_double(), 0.5));
  CHECK(Time(-1, -2) == Time(1, 2) && Time(1, -2) == Time(-1, 2));
}

static void test_config()
{
  Config c;
  CHECK(c.get_sample_rate() == 48000 && c.get_fft_implementation() == "default");
  CHECK(throws<std::invalid_argument>([&] { c.validate(); }, "Config: period size must be set"));
  c.set_period_size(512);
  CHECK(throws<std::invalid_argument>([&] { c.validate(); }, "Config: data path must be set"));
  c.set_data_path("x.tf");
  c.validate();
  Config d = c;
  d.set_num_objects_channels(3);
  CHECK(c.get_num_objects_channels() == 0 && d.get_num_objects_channels() == 3 && d.get_data_path() == "x.tf");
}

static void test_listener_axes()
{
  const double s = std::sqrt(0.5);
  struct Row { double q[4], look[3], right[3], up[3]; };
  const Row rows[] = {
      {{1, 0, 0, 0}, {0, 1, 0}, {1, 0, 0}, {0, 0, 1}},
      {{s, -s, 0, 0}, {0, 0, 1}, {1, 0, 0}, {0, -1, 0}},
      {{s, s, 0, 0}, {0, 0, -1}, {1, 0, 0}, {0, 1, 0}},
      {{s, 0, -s, 0}, {0, 1, 0}, {0, 0, -1}, {1, 0, 0}},
      {{s, 0, s, 0}, {0, 1, 0}, {0, 0, 1}, {-1, 0, 0}},
      {{s, 0, 0, -s}, {-1, 0, 0}, {0, 1, 0}, {0, 0, 1}},
      {{s, 0, 0, s}, {1, 0, 0}, {0, -1, 0}, {0, 0, 1}},
  };
  for (const Row &r : rows) {
    Listener l;
    l.set_orientation_quaternion({r.q[0], r.q[1], r.q[2], r.q[3]});
    auto lk = l.look(), rt = l.right(), up = l.up();
    for (int i = 0; i < 3; i++) CHECK(close(lk[i], r.look[i]) && close(rt[i], r.right[i]) && close(up[i], r.up[i]));
  }
}

static Quat from_rotation_vector(double x, double y, double z)
{
  double a = std::sqrt(x * x + y * y + z * z);
  if (a == 0) return {};
  double sn = std::sin(a / 2) / a;
  return {std::cos(a / 2), x * sn, y * sn, z * sn};
}

static void test_view_selection()
{
  // views every 2 degrees of yaw, as in the data files
  std::vector<float> views;
  const double rad = 3.14159265358979323846 / 180.0;
  for (int v = 0; v < 180; v++) { views.push_back((float)-std::sin(2 * v * rad)); views.push_back((float)std::cos(2 * v * rad)); views.push_back(0.f); }
  ViewSelector sel(views.data(), 180);
  auto run = [&](Quat q) {
    Listener l;
    l.set_orientation_quaternion({q.w, q.x, q.y, q.z});
    return sel.select(l);
  };
  auto same = [&](const Listener &l, Quat q) {
    auto o = l.get_orientation_quaternion();
    return close(o[0], q.w) && close(o[1], q.x) && close(o[2], q.y) && close(o[3], q.z);
  };
  auto r = run({1, 0, 0, 0});
  CHECK(r.view == 0 && same(r.residual, {1, 0, 0, 0}));
  r = run(from_rotation_vector(0, 0, -2 * rad));          // rotate onto a view we have: no residual
  CHECK(r.view == 1 && same(r.residual, {1, 0, 0, 0}));
  Quat ud = from_rotation_vector(2 * rad, 0, 0);          // up/down: all in the residual
  r = run(ud);
  CHECK(r.view == 0 && same(r.residual, ud));
  r = run(from_rotation_vector(0, 0, -2.5 * rad));        // just past a view
  CHECK(r.view == 1 && same(r.residual, from_rotation_vector(0, 0, -0.5 * rad)));
}

static void test_object_schedule()
{
  const size_t n = 3;
  auto gv = [&](double a) { ObjectsGains g; g.direct = {a, 0, 1 - a}; g.diffuse = {0, a / 2, 0}; return g; };
  auto expect = [&](ObjectSchedule &s, int64_t end_sample, std::function<double(size_t)> want) {
    std::vector<double> out(2 * n);
    s.gains_at(Time(end_sample, 48000), out.data());
    for (size_t i = 0; i < 2 * n; i++) CHECK(close(out[i], want(i), 1e-12));
  };
  auto vec = [&](const ObjectsGains &g) { return [=](size_t i) { return i < n ? g.direct[i] : g.diffuse[i - n]; }; };
  auto zero = [](size_t) { return 0.0; };
  auto push = [&](ObjectSchedule &s, ObjectsGains g, int64_t rtime, int64_t dur) {
    g.rtime = Time(rtime, 48000); g.duration = Time(dur, 48000);
    s.update(g);
  };
  ObjectsGains g0 = gv(0.2), g1 = gv(0.5), g2 = gv(0.9);
  {  // test_single_block
    ObjectSchedule s(n);
    expect(s, 512, zero);                       // nothing pushed yet: silence
    push(s, g0, 0, 512);
    expect(s, 512, vec(g0));
    expect(s, 1024, zero);
  }
  {  // test_multiple_time_aligned_blocks
    ObjectSchedule s(n);
    push(s, g0, 0, 512);     expect(s, 512, vec(g0));
    push(s, g1, 512, 512);   expect(s, 1024, vec(g1));
    push(s, g2, 1024, 1024);
    expect(s, 1536, [&](size_t i) { return 0.5 * vec(g1)(i) + 0.5 * vec(g2)(i); });
    expect(s, 2048, vec(g2));
    expect(s, 2560, zero);
  }
  {  // test_silence_between_blocks
    ObjectSchedule s(n);
    const ObjectsGains *g[2] = {&g0, &g1};
    for (int i = 0; i < 2; i++) {
      push(s, *g[i], 512 * i * 3, 512 * 2);
      expect(s, 512 * (3 * i + 1), vec(*g[i]));   // no interpolation after a gap
      expect(s, 512 * (3 * i + 2), vec(*g[i]));
      expect(s, 512 * (3 * i + 3), zero);
    }
  }
  {  // test_multiple_blocks_in_round
    ObjectSchedule s(n);
    push(s, g0, 0, 256); push(s, g1, 256, 256);   expect(s, 512, vec(g1));
    push(s, g2, 512, 256); push(s, g0, 768, 256); expect(s, 1024, vec(g0));
    expect(s, 1536, zero);
  }
  {  // api.hpp:194-207: add(a,0..100) add(b,100..200) add(c,150..200) -> a, a, (a+b)/2 at 150, c at 200
    ObjectSchedule s(n);
    auto at = [&](int64_t t, std::function<double(size_t)> want) {
      std::vector<double> out(2 * n);
      s.gains_at(Time(t, 1), out.data());
      for (size_t i = 0; i < 2 * n; i++) CHECK(close(out[i], want(i), 1e-12));
    };
    ObjectsGains a = g0, b = g1, c = g2;
    a.rtime = Time(0); a.duration = Time(100); s.update(a);  at(100, vec(g0));
    b.rtime = Time(100); b.duration = Time(100); s.update(b); at(150, [&](size_t i) { return 0.5 * vec(g0)(i) + 0.5 * vec(g1)(i); });
    c.rtime = Time(150); c.duration = Time(50); s.update(c);  at(200, vec(g2));
  }
  {  // interpolationLength shorter than the block; no timing = held forever
    ObjectSchedule s(n);
    push(s, g0, 0, 512);
    ObjectsGains g = g1; g.rtime = Time(512, 48000); g.duration = Time(1024, 48000); g.interpolationLength = Time(256, 48000);
    s.update(g);
    expect(s, 640, [&](size_t i) { return 0.5 * vec(g0)(i) + 0.5 * vec(g1)(i); });
    expect(s, 768, vec(g1));
    expect(s, 1536, vec(g1));
    ObjectsGains inf = g2;
    s.update(inf);
    expect(s, 1 << 30, vec(g2));
    CHECK(throws<std::invalid_argument>([&] { ObjectsGains bad; bad.direct = {1}; bad.diffuse = {0}; s.update(bad); }));
  }
  {  // DirectSpeakers: constant inside [rtime, rtime+duration], silent after, no interpolation
    DirectSpeakersSchedule s(2);
    std::vector<double> out(2);
    s.gains_at(Time(512, 48000), out.data());
    CHECK(out[0] == 0 && out[1] == 0);
    DirectSpeakersGains g; g.gains = {0.25, 0.75}; g.rtime = Time(0); g.duration = Time(1024, 48000);
    s.update(g);
    s.gains_at(Time(1024, 48000), out.data());
    CHECK(out[0] == 0.25 && out[1] == 0.75);
    s.gains_at(Time(1025, 48000), out.data());
    CHECK(out[0] == 0 && out[1] == 0);
  }
}

static void test_sh_rotation()
{
  std::mt19937 gen(1);
  std::normal_distribution<double> nd;
  for (int order = 1; order < 4; order++) {
    for (int iter = 0; iter < 3; iter++) {
      Quat q = Quat{nd(gen), nd(gen), nd(gen), nd(gen)}.normalized();
      const int nch = (order + 1) * (order + 1);
      std::vector<double> M = sh::rotation_matrix(q, order);
      for (int sample = 0; sample < 100; sample++) {
        std::array<double, 3> pos{nd(gen), nd(gen), nd(gen)};
        double nrm = std::sqrt(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2]);
        for (double &p : pos) p /= nrm;
        auto shv = sh::encode(order, pos), want = sh::encode(order, q.rotate(pos));
        for (int r = 0; r < nch; r++) {
          double acc = 0;
          for (int c = 0; c < nch; c++) acc += M[(size_t)r * nch + c] * shv[c];
          CHECK(close(acc, want[r], 1e-4));     // entries below 1e-5 are zeroed, as in the reference
        }
      }
    }
  }
  // identity orientation -> identity matrix; normalisation conversion and ACN placement
  auto I = sh::rotation_matrix({1, 0, 0, 0}, 2);
  for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) CHECK(close(I[r * 9 + c], r == c ? 1.0 : 0.0, 1e-9));
  CHECK(close(sh::norm_N3D(1, 1) / sh::norm_SN3D(1, 1), std::sqrt(3.0)));
  CHECK(sh::to_acn(2, -1) == 5 && sh::from_acn(5).first == 2 && sh::from_acn(5).second == -1);
  HOASchedule h(16, 4);
  HOAInput in; in.channels = {0, 1, 2, 3}; in.orders = {0, 1, 1, 1}; in.degrees = {0, -1, 0, 1}; in.normalization = "N3D";
  h.update(0, in);
  auto &m = h.matrix_at(Time(512, 48000));
  CHECK(close(m[0 * 4 + 0], 1.0) && close(m[3 * 4 + 3], 1.0 / std::sqrt(3.0)) && m[3 * 4 + 0] == 0.f && m[5 * 4 + 1] == 0.f);
  in.rtime = Time(0); in.duration = Time(512, 48000);
  h.update(0, in);
  CHECK(close(h.matrix_at(Time(512, 48000))[0], 1.0) && h.matrix_at(Time(1024, 48000))[0] == 0.f);
  CHECK(throws<std::invalid_argument>([&] { HOAInput bad = in; bad.normalization = "XYZ"; h.update(0, bad); }, "unknown normalization type"));
}

static void test_renderer_gpu(const char *data_path)
{
  Config cfg;
  cfg.set_num_objects_channels(2);
  cfg.set_num_direct_speakers_channels(1);
  cfg.set_num_hoa_channels(4);
  cfg.set_period_size(512);
  cfg.set_data_path(data_path);
  CHECK(throws<std::runtime_error>([&] { Config bad = cfg; bad.set_fft_implementation("unavailable_fft"); Renderer r(bad); },
                                   "The specified FFT wrapper does not exist"));
  CHECK(throws<std::runtime_error>([&] { Config bad = cfg; bad.set_data_path("/nonexistent.tf"); Renderer r(bad); }));
  Renderer r(cfg);
  const size_t B = 512, L = r.num_gains();
  std::vector<float> in(B, 1.0f), outl(B), outr(B), hoa0(B, 0.5f);
  const float *obj[2] = {in.data(), in.data()}, *ds[1] = {in.data()}, *hoa[4] = {hoa0.data(), in.data(), in.data(), in.data()};
  float *out[2] = {outl.data(), outr.data()};
  // time API (test_renderer.py:41-62)
  CHECK(r.get_block_start_time() == Time(0));
  r.process(obj, ds, hoa, out);
  CHECK(r.get_block_start_time() == Time(512, 48000));
  r.set_block_start_time(Time(1, 10));
  CHECK(r.get_block_start_time() == Time(1, 10));
  r.process(obj, ds, hoa, out);
  CHECK(r.get_block_start_time() == Time(1, 10) + Time(512, 48000));
  for (size_t i = 0; i < B; i++) CHECK(outl[i] == 0.f && outr[i] == 0.f);   // no metadata yet: silence
  // channel range errors (test_renderer.py:84-97)
  ObjectsGains og; og.direct.assign(L, 0.0); og.diffuse.assign(L, 0.0); og.direct[2] = 1.0;
  DirectSpeakersGains dg; dg.gains.assign(L, 0.0); dg.gains[5] = 1.0;
  CHECK(throws<std::invalid_argument>([&] { r.add_objects_block(2, og); }, "channel number out of range in add_objects_block"));
  CHECK(throws<std::invalid_argument>([&] { r.add_direct_speakers_block(1, dg); }, "channel number out of range"));
  // future blocks are refused until their start lies before the end of the next period (test_renderer.py:65-82)
  Renderer r2(cfg);
  ObjectsGains fut = og; fut.rtime = Time(512, 48000); fut.duration = Time(512, 48000);
  DirectSpeakersGains dfut = dg; dfut.rtime = Time(512, 48000); dfut.duration = Time(512, 48000);
  CHECK(!r2.add_objects_block(0, fut) && !r2.add_direct_speakers_block(0, dfut));
  r2.set_block_start_time(Time(1, 48000));
  CHECK(r2.add_objects_block(0, fut) && r2.add_direct_speakers_block(0, dfut));
  // an untimed block renders from the next period on; output becomes non-zero and finite
  CHECK(r.add_objects_block(0, og));
  HOAInput hi; hi.channels = {0, 1, 2, 3}; hi.orders = {0, 1, 1, 1}; hi.degrees = {0, -1, 0, 1};
  CHECK(r.add_hoa_block(0, hi));
  Listener l;
  const double a = 10.0 * 3.14159265358979323846 / 180.0;
  l.set_orientation_quaternion({std::cos(a / 2), 0, 0, -std::sin(a / 2)});   // yaw: view 5 of a 2-degree grid
  r.set_listener(l);
  double energy = 0;
  for (int k = 0; k < 8; k++) {
    r.process(obj, ds, hoa, out);
    for (size_t i = 0; i < B; i++) { CHECK(std::isfinite(outl[i]) && std::isfinite(outr[i])); energy += outl[i] * outl[i]; }
  }
  CHECK(energy > 1e-3);
  CHECK(r.selected_view() == 5 || r.selected_view() == 0);   // 0 when the data file has a single view
  Renderer moved = std::move(r);
  moved.process(obj, ds, hoa, out);
  std::printf("gpu: renderer ok, energy %.4g, view %d\n", energy, moved.selected_view());
}

// `render <data file> <in.bin> <out.bin> <periods>`: a fixed scene through the Renderer; the Python
// test renders the same scene through the oracle and compares.
//   object 0: gains A on [0, 1024), then B on [1024, 2048) (interpolated over that block), then silence
//   object 1: untimed gains C;  DirectSpeakers 0: untimed, loudspeaker 5;  HOA: 4 SN3D channels
//   listener: yaw 0 for periods 0-1, then -4.6 degrees from period 2 on (nearest view: 4 degrees)
static int render_scene(const char *data_path, const char *in_path, const char *out_path, int periods)
{
  const size_t B = 256;
  Config cfg;
  cfg.set_num_objects_channels(2); cfg.set_num_direct_speakers_channels(1); cfg.set_num_hoa_channels(4);
  cfg.set_period_size(B); cfg.set_data_path(data_path);
  Renderer r(cfg);
  const size_t L = r.num_gains(), C = 7;
  std::vector<float> in(C * B * periods), out(2 * B * periods);
  FILE *f = std::fopen(in_path, "rb");
  if (!f || std::fread(in.data(), sizeof(float), in.size(), f) != in.size()) return 2;
  std::fclose(f);
  auto gains = [&](std::initializer_list<std::pair<int, double>> d, std::initializer_list<std::pair<int, double>> q) {
    ObjectsGains g; g.direct.assign(L, 0.0); g.diffuse.assign(L, 0.0);
    for (auto &p : d) g.direct[p.first] = p.second;
    for (auto &p : q) g.diffuse[p.first] = p.second;
    return g;
  };
  ObjectsGains A = gains({{2, 0.8}, {5, 0.3}}, {}), Bg = gains({{6, 0.5}, {9, 0.6}}, {{6, 0.2}, {9, 0.3}}), Cg = gains({{0, 0.4}}, {{0, 0.4}, {1, 0.1}});
  A.rtime = Time(0); A.duration = Time(1024, 48000);
  Bg.rtime = Time(1024, 48000); Bg.duration = Time(1024, 48000);
  bool pushedB = false;
  if (!r.add_objects_block(0, A)) return 3;
  if (!r.add_objects_block(1, Cg)) return 3;
  DirectSpeakersGains dg; dg.gains.assign(L, 0.0); dg.gains[5] = 1.0;
  r.add_direct_speakers_block(0, dg);
  HOAInput hi; hi.channels = {0, 1, 2, 3}; hi.orders = {0, 1, 1, 1}; hi.degrees = {0, -1, 0, 1};
  r.add_hoa_block(0, hi);
  for (int k = 0; k < periods; k++) {
    if (!pushedB) pushedB = r.add_objects_block(0, Bg);   // refused until its start is within reach
    if (k == 2) {
      Listener l;
      const double a = 4.6 * 3.14159265358979323846 / 180.0;
      l.set_orientation_quaternion({std::cos(a / 2), 0, 0, -std::sin(a / 2)});
      r.set_listener(l);
    }
    const float *ip[C];
    for (size_t c = 0; c < C; c++) ip[c] = in.data() + (c * periods + k) * B;
    float *op[2] = {out.data() + (size_t)k * B, out.data() + (size_t)(periods + k) * B};
    r.process(ip, ip + 2, ip + 3, op);
  }
  f = std::fopen(out_path, "wb");
  std::fwrite(out.data(), sizeof(float), out.size(), f);
  std::fclose(f);
  std::printf("rendered %d periods, view %d\n", periods, r.selected_view());
  return 0;
}

// `test_host hoa_schedule <quaternions> <out> <count> <periods>` (no GPU): what the facade's HOA gain
// calculation (HOASchedule = GainCalcHOA, gain_calc_hoa.cpp:86-124, with sh::rotation_matrix standing in
// for quaternion_to_sh_rotation_matrix, sh_rotation.cpp:140-157) emits per period for a first-order SN3D
// stream of 4 channels when listener i (orientation quaternion {w,x,y,z}, doubles) is set before period
// i -- the scenario of the reference's head-tracked HOA golden (test_changes.py:183-217,263-281).
// Writes [periods][16][4] floats.
static int dump_hoa_schedule(const char *quat_path, const char *out_path, int count, int periods)
{
  std::vector<double> q(4 * (size_t)count);
  FILE *f = std::fopen(quat_path, "rb");
  if (!f || std::fread(q.data(), sizeof(double), q.size(), f) != q.size()) return 2;
  std::fclose(f);
  const size_t B = 512, fs = 48000, A = 16, H = 4;
  HOASchedule sched(A, H);
  HOAInput hi; hi.channels = {0, 1, 2, 3}; hi.orders = {0, 1, 1, 1}; hi.degrees = {0, -1, 0, 1}; hi.normalization = "SN3D";
  hi.rtime = Time(0); hi.duration = Time(1);
  sched.update(0, hi);
  std::vector<float> out;
  for (int k = 0; k < periods; k++) {
    if (k < count) {
      Listener l;
      l.set_orientation_quaternion({q[4 * k], q[4 * k + 1], q[4 * k + 2], q[4 * k + 3]});
      sched.set_orientation(l.orientation());
    }
    const auto &m = sched.matrix_at(Time((int64_t)((k + 1) * B), (int64_t)fs));
    out.insert(out.end(), m.begin(), m.end());
  }
  f = std::fopen(out_path, "wb");
  std::fwrite(out.data(), sizeof(float), out.size(), f);
  std::fclose(f);
  return 0;
}

// `test_host render_hoa <data file> <audio> <quaternions> <out> <count> <periods>`: the same scenario
// through the whole Renderer facade on cuda:0 (add_hoa_block, set_listener before each period, process).
static int render_hoa_head_tracked(const char *data_path, const char *in_path, const char *quat_path, const char *out_path,
                                   int count, int periods)
{
  const size_t B = 512, H = 4;
  std::vector<double> q(4 * (size_t)count);
  FILE *f = std::fopen(quat_path, "rb");
  if (!f || std::fread(q.data(), sizeof(double), q.size(), f) != q.size()) return 2;
  std::fclose(f);
  std::vector<float> in(H * B * periods), out(2 * B * periods);
  f = std::fopen(in_path, "rb");
  if (!f || std::fread(in.data(), sizeof(float), in.size(), f) != in.size()) return 2;
  std::fclose(f);
  Config cfg;
  cfg.set_num_hoa_channels(H); cfg.set_period_size(B); cfg.set_data_path(data_path);
  Renderer r(cfg);
  HOAInput hi; hi.channels = {0, 1, 2, 3}; hi.orders = {0, 1, 1, 1}; hi.degrees = {0, -1, 0, 1}; hi.normalization = "SN3D";
  hi.rtime = Time(0); hi.duration = Time(1);
  if (!r.add_hoa_block(0, hi)) return 3;
  for (int k = 0; k < periods; k++) {
    if (k < count) {
      Listener l;
      l.set_orientation_quaternion({q[4 * k], q[4 * k + 1], q[4 * k + 2], q[4 * k + 3]});
      r.set_listener(l);
    }
    const float *ip[H];
    for (size_t c = 0; c < H; c++) ip[c] = in.data() + (c * periods + k) * B;
    float *op[2] = {out.data() + (size_t)k * B, out.data() + (size_t)(periods + k) * B};
    r.process(nullptr, nullptr, ip, op);
  }
  f = std::fopen(out_path, "wb");
  std::fwrite(out.data(), sizeof(float), out.size(), f);
  std::fclose(f);
  return 0;
}

int main(int argc, char **argv)
{
  if (argc >= 6 && std::string(argv[1]) == "render") return render_scene(argv[2], argv[3], argv[4], std::atoi(argv[5]));
  if (argc >= 6 && std::string(argv[1]) == "hoa_schedule")
    return dump_hoa_schedule(argv[2], argv[3], std::atoi(argv[4]), std::atoi(argv[5]));
  if (argc >= 8 && std::string(argv[1]) == "render_hoa")
    return render_hoa_head_tracked(argv[2], argv[3], argv[4], argv[5], std::atoi(argv[6]), std::atoi(argv[7]));
  test_time();
  test_config();
  test_listener_axes();
  test_view_selection();
  test_object_schedule();
  test_sh_rotation();
  if (argc >= 3 && std::string(argv[1]) == "gpu") test_renderer_gpu(argv[2]);
  std::printf("%s (%d failures)\n", failures ? "FAILED" : "OK", failures);
  return failures ? 1 : 0;
}
