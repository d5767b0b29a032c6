// bear_b200/dynamic.hpp -- SURVEY.md section 8 row f-3: the host-side wrappers that sit around a
// renderer in the reference's embedders.
//
//   VariableBlockSizeAdapter   visr_bear/include/bear/variable_block_size.hpp:19-38,
//                              visr_bear/src/variable_block_size.cpp:22-99 (+ libear's
//                              ear::dsp::VariableBlockSizeAdapter, variable_block_size_impl.cpp)
//   BackgroundBuilder          visr_bear/src/constructor_thread.{hpp,cpp}
//   DynamicRenderer            visr_bear/src/dynamic_renderer.{hpp,cpp}
//
// They are pure host control: buffering arbitrary host block sizes into fixed periods (one period
// of delay), building a new renderer on a background thread, and the fade-down / wait / pre-roll /
// fade-up sequence around a configuration change.  Everything is templated on the renderer type so
// that the logic can be exercised without a GPU; `bear_b200::Renderer` (host.hpp) is the default.
// Building a renderer on the GPU takes milliseconds instead of the reference's seconds (all BRIRs
// are transformed by one kernel launch), but the embedder-visible behaviour is kept: set_config()
// never blocks the audio thread.
#pragma once

#include <cmath>
#include <condition_variable>
#include <exception>
#include <map>
#include <memory>
#include <mutex>
#include <queue>
#include <thread>

#include "host.hpp"

namespace bear_b200 {

// ---------------------------------------------------------------------------------------------
// Fixed-period processing fed with arbitrary block sizes (adds one period of delay).
// ---------------------------------------------------------------------------------------------
template <class RendererT = Renderer>
class VariableBlockSizeAdapterT {
 public:
  VariableBlockSizeAdapterT(const Config &config, std::shared_ptr<RendererT> renderer)
      : cfg_(config), renderer_(std::move(renderer)), B_(config.get_period_size()),
        n_in_(config.get_num_objects_channels() + config.get_num_direct_speakers_channels() + config.get_num_hoa_channels()),
        in_buf_(n_in_ * B_, 0.0f), out_buf_(2 * B_, 0.0f), in_ptr_(n_in_), objects_queue_(config.get_num_objects_channels()),
        direct_speakers_queue_(config.get_num_direct_speakers_channels())
  {
    for (size_t c = 0; c < n_in_; c++) in_ptr_[c] = in_buf_.data() + c * B_;
  }

  /// Process nsamples samples of every channel.  Output lags the input by one period.
  void process(size_t nsamples, const Sample *const *objects_input, const Sample *const *direct_speakers_input,
               const Sample *const *hoa_input, Sample *const *output)
  {
    std::vector<const Sample *> in(n_in_);
    size_t j = 0;
    for (size_t i = 0; i < cfg_.get_num_objects_channels(); i++) in[j++] = objects_input[i];
    for (size_t i = 0; i < cfg_.get_num_direct_speakers_channels(); i++) in[j++] = direct_speakers_input[i];
    for (size_t i = 0; i < cfg_.get_num_hoa_channels(); i++) in[j++] = hoa_input[i];
    size_t done = 0;
    while (done < nsamples) {
      // the input buffer fills from the front while the previous period's output drains from the
      // same positions; a full input buffer triggers one fixed-size period
      const size_t n = std::min(nsamples - done, B_ - fill_);
      for (size_t c = 0; c < n_in_; c++) std::copy(in[c] + done, in[c] + done + n, in_buf_.begin() + c * B_ + fill_);
      for (size_t c = 0; c < 2; c++) std::copy(out_buf_.begin() + c * B_ + fill_, out_buf_.begin() + c * B_ + fill_ + n, output[c] + done);
      done += n;
      fill_ += n;
      if (fill_ == B_) {
        run_period();
        fill_ = 0;
      }
    }
    samples_processed_ += nsamples;
  }

  /// queue metadata; `nsamples` is the size of the process() call that will follow.  Returns false
  /// (try again after that call) when the block starts after it.
  bool add_objects_block(size_t nsamples, size_t channel, ObjectsGains metadata)
  {
    if (!in_reach(metadata, nsamples)) return false;
    objects_queue_.at(channel).push(std::move(metadata));
    return true;
  }
  bool add_direct_speakers_block(size_t nsamples, size_t channel, DirectSpeakersGains metadata)
  {
    if (!in_reach(metadata, nsamples)) return false;
    direct_speakers_queue_.at(channel).push(std::move(metadata));
    return true;
  }
  bool add_hoa_block(size_t nsamples, size_t stream, HOAInput metadata)
  {
    if (!in_reach(metadata, nsamples)) return false;
    hoa_queue_[stream].push(std::move(metadata));
    return true;
  }
  /// delay added by the adapter, in samples
  size_t get_delay() const { return B_; }

 private:
  bool in_reach(const MetadataInput &m, size_t nsamples) const
  {
    return !m.rtime || *m.rtime < Time((int64_t)(samples_processed_ + nsamples), (int64_t)cfg_.get_sample_rate());
  }
  void run_period()
  {
    for (size_t i = 0; i < objects_queue_.size(); i++)
      while (!objects_queue_[i].empty() && renderer_->add_objects_block(i, objects_queue_[i].front())) objects_queue_[i].pop();
    for (size_t i = 0; i < direct_speakers_queue_.size(); i++)
      while (!direct_speakers_queue_[i].empty() && renderer_->add_direct_speakers_block(i, direct_speakers_queue_[i].front()))
        direct_speakers_queue_[i].pop();
    for (auto &kv : hoa_queue_)
      while (!kv.second.empty() && renderer_->add_hoa_block(kv.first, kv.second.front())) kv.second.pop();
    Sample *out[2] = {out_buf_.data(), out_buf_.data() + B_};
    const size_t N = cfg_.get_num_objects_channels(), M = cfg_.get_num_direct_speakers_channels();
    renderer_->process(in_ptr_.data(), in_ptr_.data() + N, in_ptr_.data() + N + M, out);
  }

  Config cfg_;
  std::shared_ptr<RendererT> renderer_;
  size_t B_, n_in_, fill_ = 0, samples_processed_ = 0;
  std::vector<Sample> in_buf_, out_buf_;
  std::vector<const Sample *> in_ptr_;
  std::vector<std::queue<ObjectsGains>> objects_queue_;
  std::vector<std::queue<DirectSpeakersGains>> direct_speakers_queue_;
  std::map<size_t, std::queue<HOAInput>> hoa_queue_;
};
using VariableBlockSizeAdapter = VariableBlockSizeAdapterT<Renderer>;

// ---------------------------------------------------------------------------------------------
// Builds (and destroys) renderers on a background thread; the audio thread only ever try-locks.
// ---------------------------------------------------------------------------------------------
template <class RendererT = Renderer>
class BackgroundBuilderT {
 public:
  BackgroundBuilderT() : thread_([this] { run(); }) {}
  ~BackgroundBuilderT()
  {
    {
      std::lock_guard<std::mutex> lk(mu_);
      exit_ = true;
    }
    cv_.notify_one();
    thread_.join();
  }
  BackgroundBuilderT(const BackgroundBuilderT &) = delete;
  BackgroundBuilderT &operator=(const BackgroundBuilderT &) = delete;

  /// ask for a renderer of this configuration; false = busy, try again later.  Never blocks.
  bool start(const Config &config)
  {
    std::unique_lock<std::mutex> lk(mu_, std::try_to_lock);
    if (!lk) return false;
    request_ = config;
    result_.reset();        // a result or error of an older request must not be picked up now
    error_ = nullptr;
    cv_.notify_one();
    return true;
  }
  /// the finished renderer, if there is one and the builder is idle.  Never blocks.
  std::optional<RendererT> get_result()
  {
    std::unique_lock<std::mutex> lk(mu_, std::try_to_lock);
    if (!lk || !result_) return std::nullopt;
    std::optional<RendererT> out = std::move(result_);
    result_.reset();
    return out;
  }
  /// exchange `slot` with the finished renderer; the old content of `slot` is destroyed on the
  /// background thread.  False (and no effect) if there is no result or the previous old renderer
  /// has not been disposed of yet.  Never blocks.
  bool try_swap_result(std::optional<RendererT> &slot)
  {
    std::unique_lock<std::mutex> lk(mu_, std::try_to_lock);
    if (!lk || !result_ || (slot && retired_)) return false;
    if (slot) retired_ = std::move(slot);
    slot = std::move(result_);
    result_.reset();
    cv_.notify_one();
    return true;
  }
  /// the exception of the last request's construction, once
  std::exception_ptr get_construction_error()
  {
    std::unique_lock<std::mutex> lk(mu_, std::try_to_lock);
    std::exception_ptr e = nullptr;
    if (lk) std::swap(e, error_);
    return e;
  }

 private:
  void run()
  {
    std::unique_lock<std::mutex> lk(mu_);
    for (;;) {
      cv_.wait(lk, [&] { return request_ || retired_ || exit_; });
      if (exit_) return;
      retired_.reset();
      if (request_) {
        // the lock stays held while building: nothing useful can be done with the builder meanwhile,
        // and every caller only try-locks
        try {
          result_.emplace(*request_);
        } catch (...) {
          error_ = std::current_exception();
        }
        request_.reset();
      }
    }
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::optional<Config> request_;
  std::optional<RendererT> result_, retired_;
  std::exception_ptr error_ = nullptr;
  bool exit_ = false;
  std::thread thread_;
};

// ---------------------------------------------------------------------------------------------
// Renderer whose configuration can change while audio runs.  After set_config() the output fades
// down over one period, stays silent while the new renderer is built and pre-rolled (long enough
// to flush the BRIR, decorrelator and delay tails), then fades up over one period.
// ---------------------------------------------------------------------------------------------
template <class RendererT = Renderer>
class DynamicRendererT {
 public:
  /// reference constants (dynamic_renderer.cpp:120-123): BRIR 2976 + decorrelators 128 + delays 38 + 3
  static constexpr int kPrerollSamples = 2976 + 128 + 38 + 3;

  DynamicRendererT(size_t period_size, size_t max_size)
      : B_(period_size), preroll_periods_((int)std::ceil((double)kPrerollSamples / (double)period_size)),
        ramp_up_(period_size), ramp_down_(period_size), zeros_(period_size, 0.0f)
  {
    objects_recent_.reserve(max_size); direct_speakers_recent_.reserve(max_size); hoa_recent_.reserve(max_size);
    for (size_t n = 0; n < B_; n++) {   // start-inclusive ramps (dynamic_renderer.cpp:187-188)
      ramp_up_[n] = (Sample)n / (Sample)B_;
      ramp_down_[n] = (Sample)1 - (Sample)n / (Sample)B_;
    }
  }

  /// non-blocking reconfiguration; metadata for the new configuration can be pushed right away
  void set_config(const Config &config)
  {
    adopt(config);
    if (!builder_.start(config)) deferred_ = config;
    state_ = (state_ == State::Running || state_ == State::FadeDown) ? State::FadeDown : State::Wait;
  }
  /// blocking reconfiguration without fades (before audio starts); construction errors propagate
  void set_config_blocking(const Config &config)
  {
    renderer_.emplace(config);
    adopt(config);
    current_ = next_;
    prime_renderer();
    state_ = State::Running;
  }
  std::exception_ptr get_error() { return builder_.get_construction_error(); }
  bool is_running() const { return state_ == State::Running; }

  /// one period.  Channels beyond the configured count are ignored, missing ones are silent.
  void process(size_t num_objects, const Sample *const *objects_input, size_t num_direct_speakers,
               const Sample *const *direct_speakers_input, size_t num_hoa, const Sample *const *hoa_input,
               Sample *const *output)
  {
    if (deferred_) {
      (void)builder_.get_result();   // a stale result would block the builder
      if (builder_.start(*deferred_)) deferred_.reset();
    }
    if ((state_ == State::Start || state_ == State::Wait) && builder_.try_swap_result(renderer_)) {
      current_ = next_;
      prime_renderer();
      if (state_ == State::Wait) {
        state_ = State::Preroll;
        preroll_left_ = preroll_periods_;
      }
    }
    const bool render = state_ == State::Running || state_ == State::FadeDown || state_ == State::Preroll || state_ == State::FadeUp;
    if (render) {
      if (!renderer_) throw std::logic_error("expected renderer to be set");
      pad(tmp_obj_, current_.get_num_objects_channels(), num_objects, objects_input);
      pad(tmp_ds_, current_.get_num_direct_speakers_channels(), num_direct_speakers, direct_speakers_input);
      pad(tmp_hoa_, current_.get_num_hoa_channels(), num_hoa, hoa_input);
      renderer_->process(tmp_obj_.data(), tmp_ds_.data(), tmp_hoa_.data(), output);
    }
    if (state_ == State::FadeDown)
      for (int c = 0; c < 2; c++)
        for (size_t n = 0; n < B_; n++) output[c][n] *= ramp_down_[n];
    if (state_ == State::FadeUp)
      for (int c = 0; c < 2; c++)
        for (size_t n = 0; n < B_; n++) output[c][n] *= ramp_up_[n];
    if (state_ == State::Start || state_ == State::Wait || state_ == State::Preroll)
      for (int c = 0; c < 2; c++) std::fill(output[c], output[c] + B_, (Sample)0);
    // end-of-period transitions
    if (state_ == State::FadeDown) state_ = State::Wait;
    else if (state_ == State::Preroll) { if (--preroll_left_ == 0) state_ = State::FadeUp; }
    else if (state_ == State::FadeUp) state_ = State::Running;
    periods_++;
  }

  bool add_objects_block(size_t channel, ObjectsGains m)
  {
    return add(objects_recent_, 2, channel, std::move(m), channel < next_.get_num_objects_channels(),
               channel < current_.get_num_objects_channels(), [&](const ObjectsGains &x) { return renderer_->add_objects_block(channel, x); });
  }
  bool add_direct_speakers_block(size_t channel, DirectSpeakersGains m)
  {
    return add(direct_speakers_recent_, 1, channel, std::move(m), channel < next_.get_num_direct_speakers_channels(),
               channel < current_.get_num_direct_speakers_channels(),
               [&](const DirectSpeakersGains &x) { return renderer_->add_direct_speakers_block(channel, x); });
  }
  bool add_hoa_block(size_t stream, HOAInput m)
  {
    auto fits = [&](const Config &c) { for (size_t ch : m.channels) if (ch >= c.get_num_hoa_channels()) return false; return true; };
    const bool fn = fits(next_), fc = fits(current_);
    return add(hoa_recent_, 1, stream, std::move(m), fn, fc, [&](const HOAInput &x) { return renderer_->add_hoa_block(stream, x); });
  }
  Time get_block_start_time() const { return raw_start() - time_offset_; }
  void set_block_start_time(const Time &t) { time_offset_ = raw_start() - t; }
  void set_listener(const Listener &l, const std::optional<Time> &interpolation_time = {})
  {
    if (renderer_) renderer_->set_listener(l, interpolation_time);
    listener_ = l;
    listener_interp_ = interpolation_time;
  }

 private:
  enum class State { Start, FadeDown, Wait, Preroll, FadeUp, Running };   // what the NEXT period does

  void adopt(const Config &config)
  {
    if (config.get_period_size() != B_) throw std::logic_error("config has incorrect period size");
    sample_rate_ = config.get_sample_rate();
    next_ = config;
  }
  Time raw_start() const { return Time((int64_t)(periods_ * B_), (int64_t)std::max<size_t>(sample_rate_, 1)); }
  Time raw_next_start() const { return Time((int64_t)((periods_ + 1) * B_), (int64_t)std::max<size_t>(sample_rate_, 1)); }
  bool rendering() const { return state_ == State::Running || state_ == State::FadeDown || state_ == State::Preroll || state_ == State::FadeUp; }

  template <class T, class Push>
  bool add(std::vector<std::vector<T>> &recent, size_t keep, size_t channel, T m, bool fits_next, bool fits_current, Push &&push)
  {
    if (!fits_next) throw std::invalid_argument("not enough channels configured to add block");
    if (m.rtime) *m.rtime += time_offset_;
    if (m.rtime && !(*m.rtime < raw_next_start())) return false;
    // always remembered, so that a renderer built later can be brought up to date; passed on only
    // while the renderer actually runs (otherwise its clock and ours would drift apart)
    if (channel >= recent.size()) recent.resize(channel + 1);
    if (recent[channel].size() == keep) recent[channel].erase(recent[channel].begin());
    recent[channel].push_back(m);
    if (rendering() && fits_current) {
      if (!renderer_) throw std::logic_error("expected renderer to be set");
      if (!push(m)) throw std::logic_error("could not push");
    }
    return true;
  }
  /// bring a freshly built renderer up to date: time, listener, the remembered metadata
  void prime_renderer()
  {
    renderer_->set_block_start_time(Time((int64_t)(B_ * periods_), (int64_t)sample_rate_));
    renderer_->set_listener(listener_, listener_interp_);
    for (size_t ch = 0; ch < objects_recent_.size(); ch++)
      if (ch < current_.get_num_objects_channels())
        for (auto &m : objects_recent_[ch])
          if (!renderer_->add_objects_block(ch, m)) throw std::logic_error("could not push");
    for (size_t ch = 0; ch < direct_speakers_recent_.size(); ch++)
      if (ch < current_.get_num_direct_speakers_channels())
        for (auto &m : direct_speakers_recent_[ch])
          if (!renderer_->add_direct_speakers_block(ch, m)) throw std::logic_error("could not push");
    for (size_t st = 0; st < hoa_recent_.size(); st++)
      for (auto &m : hoa_recent_[st]) {
        bool fits = true;
        for (size_t ch : m.channels) fits = fits && ch < current_.get_num_hoa_channels();
        if (fits && !renderer_->add_hoa_block(st, m)) throw std::logic_error("could not push");
      }
  }
  void pad(std::vector<const Sample *> &out, size_t want, size_t have, const Sample *const *in)
  {
    out.resize(want);
    for (size_t i = 0; i < want; i++) out[i] = i < have ? in[i] : zeros_.data();
  }

  size_t B_, periods_ = 0, sample_rate_ = 0;
  int preroll_periods_, preroll_left_ = 0;
  State state_ = State::Start;
  std::vector<Sample> ramp_up_, ramp_down_, zeros_;
  std::vector<std::vector<ObjectsGains>> objects_recent_;            // last 2 blocks per channel
  std::vector<std::vector<DirectSpeakersGains>> direct_speakers_recent_;  // last block per channel
  std::vector<std::vector<HOAInput>> hoa_recent_;                    // last block per stream
  std::optional<RendererT> renderer_;
  std::optional<Config> deferred_;
  Config current_, next_;
  BackgroundBuilderT<RendererT> builder_;
  std::vector<const Sample *> tmp_obj_, tmp_ds_, tmp_hoa_;
  Time time_offset_;
  Listener listener_;
  std::optional<Time> listener_interp_;
};
using DynamicRenderer = DynamicRendererT<Renderer>;

}  // namespace bear_b200
