// Host side of libbear_cuda.so: device state of one renderer and the C ABI of include/bear_cuda.h.
// Replaces the construction and per-period work of bear::DSP (visr_bear/src/dsp.cpp:7-236).
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/bear_cuda.h"
#include "kernels.cuh"
#include "tenf.hpp"

using namespace bearcuda;

namespace {

thread_local std::string g_create_error;

struct CudaFail {
  std::string msg;
};
#define CK(expr)                                                                                          \
  do {                                                                                                    \
    cudaError_t e__ = (expr);                                                                             \
    if (e__ != cudaSuccess)                                                                               \
      throw CudaFail{std::string(#expr) + ": " + cudaGetErrorName(e__) + " (" + cudaGetErrorString(e__) + ")"}; \
  } while (0)

struct InvalidArgument { std::string msg; };
struct Unsupported { std::string msg; };

size_t round_up(size_t v, size_t m) { return (v + m - 1) / m * m; }
bool mix_hist_too_large(int hist)
{
  return mix_smem_bytes(hist) > 200 * 1024;
}

// ---- NCCL, bound at run time -------------------------------------------------------------------
// The library has no link-time dependency on NCCL: a renderer that is never connected to peers needs
// none, and an embedder that already has NCCL loaded (PyTorch brings its own copy) must not get a
// second one.  dlopen("libnccl.so.2") returns the copy that is already in the process, else the
// system's.  Only what the render path uses is bound; the constants are NCCL 2.x ABI (nccl.h).
struct NcclUniqueId { char internal[128]; };
typedef void *NcclComm;
constexpr int kNcclFloat32 = 7, kNcclSum = 0;
struct NcclApi {
  void *handle = nullptr;
  int (*GetVersion)(int *) = nullptr;
  int (*GetUniqueId)(NcclUniqueId *) = nullptr;
  int (*CommInitRank)(NcclComm *, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*Reduce)(const void *, void *, size_t, int, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};

const NcclApi &nccl()
{
  static NcclApi api = [] {
    NcclApi a;
    a.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!a.handle) throw CudaFail{std::string("NCCL is not available: ") + dlerror()};
    auto sym = [&](const char *name) {
      void *f = dlsym(a.handle, name);
      if (!f) throw CudaFail{std::string("NCCL symbol missing: ") + name};
      return f;
    };
    a.GetVersion = reinterpret_cast<decltype(a.GetVersion)>(sym("ncclGetVersion"));
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
    a.Reduce = reinterpret_cast<decltype(a.Reduce)>(sym("ncclReduce"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
    return a;
  }();
  return api;
}
#define NCK(expr)                                                                                         \
  do {                                                                                                    \
    int e__ = (expr);                                                                                     \
    if (e__ != 0) throw CudaFail{std::string(#expr) + ": NCCL error " + nccl().GetErrorString(e__)};      \
  } while (0)

// What the ranks of one scene tell each other before the first period (bear_cuda_comm_blob).
struct CommBlob {
  uint32_t magic, B, Tmax, L;
  uint64_t cs;
  uint32_t N, M, H, rank;
  cudaIpcMemHandle_t bus;
  NcclUniqueId id;       // rank 0's is the one that counts
};
static_assert(sizeof(CommBlob) <= BEAR_CUDA_COMM_BLOB_BYTES, "blob too large");
constexpr uint32_t kCommMagic = 0x42454152u;   // "BEAR"

}  // namespace

// Everything derived from the data file for one (device, period size): filter spectra, twiddles, the
// delay and gain-normalisation tables.  Immutable after construction and shared (reference counted)
// between renderers created with bear_cuda_create_shared, so N instances cost one 212 MB table.
struct FilterBank {
  int device = 0, B = 0, logB = 0, V = 0, L = 0, A = 0, P = 0, gnS = 0, front = 0;
  uint32_t auto_max_delay = 0;
  double fs = 48000, dec_delay = 0, hoa_delay = 0, default_direct_delay = 0, default_static_delay = 0;
  const float2 *h_brir = nullptr, *h_dec = nullptr, *h_hoa = nullptr, *tw = nullptr;
  const float *delays_tab = nullptr, *gn_factors = nullptr;
  std::vector<void *> allocs;
  uint64_t bytes = 0;
  uint64_t launches = 0;
  template <class T>
  T *alloc(size_t count)
  {
    void *p = nullptr;
    size_t n = (count * sizeof(T) + 16 + 255) / 256 * 256;
    cudaError_t e = cudaMalloc(&p, n);
    if (e != cudaSuccess) throw CudaFail{std::string("cudaMalloc: ") + cudaGetErrorString(e)};
    allocs.push_back(p);
    bytes += n;
    return static_cast<T *>(p);
  }
  ~FilterBank()
  {
    cudaSetDevice(device);
    for (void *p : allocs) cudaFree(p);
  }
};

constexpr int kBusSets = 3;   // sets of bus rows a shared scene cycles through (see run_pipeline)

struct bear_cuda_renderer {
  std::shared_ptr<FilterBank> bank;
  bear_cuda_config cfg{};
  Layout lay{};
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;      // host->device uploads of the host-buffer path
  cudaStream_t d2h_stream = nullptr;       // device->host reads of finished chunks (host-buffer path)
  int ks_cfg = 1;                          // configured cMAC splits (the minimum; short launches use more)
  int ks_units = 1;                        // cMAC work units (loudspeakers + HOA decode)
  bool ks_dynamic = true;                  // false while the caller provides the partial-spectrum scratch
  std::vector<cudaEvent_t> h2d_events;     // one per upload chunk
  std::vector<cudaEvent_t> done_events;    // one per rendered chunk
  bool tables_on_device = false;           // bear_cuda_control_tables filled the staging tables
  int s_cfg = 0;                           // configured object splits (0 = per launch)
  float *bus[kBusSets] = {};               // direct + diffuse bus rows [3L][cs]; [1...] only while sharing a scene
  uint64_t shared_calls = 0;               // calls made as part of a shared scene
  int parity = 0;                          // which of the sets the current call writes (kBusSets once shared)
  int *h_err = nullptr;                    // pinned copy of the kernels' error flags
  cudaEvent_t ev_order = nullptr;          // orders the copy stream behind the render stream
  float2 *static_scratch = nullptr;        // device address of the library's built-in scratch array
  uint64_t calls = 0;                      // process calls since creation / reset
  // one scene shared with other GPUs (bear_cuda_comm_connect)
  NcclComm comm = nullptr;
  int comm_root = -1;                      // rank that receives the mix, -1 = every rank
  float *peer_bus[kMaxRanks][kBusSets] = {};   // every rank's bus rows (cudaIpc mappings), [rank] = own
  float *bus_sum = nullptr;                // [3L][cs] rows of this rank's loudspeakers summed over the ranks
  size_t comm_bytes_per_period = 0;        // largest upload per period among the ranks (same chunking everywhere)
  cudaStream_t conv_stream = nullptr;      // shared scene: exchange + convolution of chunk c run here, under the mix of chunk c+1
  cudaStream_t result_stream = nullptr;    // the stream a chunk's output is complete on (render stream, or conv_stream)
  std::vector<cudaEvent_t> mix_events;     // one per chunk: its bus rows are written
  cudaEvent_t ev_call = nullptr;
  int shared_chunks = 1;                   // chunks a shared-scene call is cut into (BEAR_CUDA_SHARED_CHUNKS overrides)
  // pipelined calls of a shared scene (bear_cuda_process_device_pipelined): the exchange + convolution of
  // call s stay queued on conv_stream while the mix of call s+1 runs on the render stream
  bool pipelined_pending = false;          // conv_stream holds work the render stream has not waited for
  cudaEvent_t ev_barrier[2] = {nullptr, nullptr};   // the ordering collective of pipelined call c has completed: [c & 1]
  int32_t *sh_view[kBusSets] = {};         // the convolution side's own copies of the view / static-delay tables
  float *sh_dstat[kBusSets] = {};
  bool shared_bus = false;                 // the bus rows are the kBusSets sets that peers may map
  // the rest of the allocation the peers map: an inbox for the ranks' stereo partial mixes (used on the
  // rank that receives the result) and two arrays of counters (rows complete / partial mix stored)
  size_t inbox_off = 0, flags_off = 0;     // in floats from bus[0]
  float *peer_base[kMaxRanks] = {};        // = peer_bus[w][0]
  int bar_count = 0, out_count = 0;        // synchronisation points passed so far (same on every rank)
  // development aid (BEAR_CUDA_TIMELINE=<path prefix>): when each side of the pipelined calls reached
  // each kernel, in ms since the first mark, written to <prefix>.rank<r> by bear_cuda_synchronize
  struct TimelineMark { cudaEvent_t e; int kernel; int side; uint64_t call; };
  std::vector<TimelineMark> timeline;
  std::string timeline_path;
  cudaEvent_t timeline_base = nullptr;
  int chunk = 0;                           // periods per launch group (0 = whole call)
  int h2d_chunk = 0;                       // periods per upload chunk of the host-buffer path
  std::vector<void *> allocs;
  std::vector<std::pair<void *, size_t>> state_allocs;
  uint64_t device_bytes = 0;
  uint64_t launches = 0;
  uint32_t kernels_per_call = 0;
  std::string error;
  double default_direct_delay = 0, default_static_delay = 0;

  // optional per-kernel timing: a chain of events, kernel i runs between events i and i+1
  bool profiling = false;
  std::vector<cudaEvent_t> ev_free;
  std::vector<std::vector<cudaEvent_t>> ev_pending;  // one chain per profiled call
  std::vector<std::vector<int>> ev_kernel;           // kernel id of each link
  double prof_ms[BEAR_CUDA_NUM_KERNELS] = {0};
  uint64_t prof_n[BEAR_CUDA_NUM_KERNELS] = {0};

  // staging (device) for the T blocks of the current call: slot 1 of each table
  float *audio_out = nullptr;
  // pinned host staging for the single-block API: one packed buffer [audio | tables] and its
  // device mirror, so that a period costs one upload
  float *h_in = nullptr, *h_out = nullptr, *d_pack = nullptr;
  // single-period host path as one CUDA graph (upload, scatter, the kernels, read-back): every
  // parameter of a one-period call is constant, only the contents of h_graph change.  h_graph holds
  // the audio and ALL tables of the period at fixed offsets; a port passed as NULL keeps last period's
  // values there, which is valid as long as the last call went through this path (shadow_valid).
  // offline rendering from interleaved file frames: raw frames of one call and the track map (grown on demand)
  unsigned char *d_frames = nullptr;
  size_t d_frames_bytes = 0;
  int *d_tracks = nullptr;
  float *h_graph = nullptr;
  struct StreamSlot { size_t off, count; float *dev; size_t slot; };
  std::vector<StreamSlot> stream_slots;    // direct gains, diffuse gains, ..., view (order of bear_cuda_process)
  size_t stream_floats = 0;
  bool shadow_valid = false, graph_ok = true;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  uint64_t graph_launch_count = 0;         // kernels in the graph
  size_t h_in_floats = 0;

  template <class T>
  T *dalloc(size_t count, bool zero = true)
  {
    void *p = nullptr;
    size_t bytes = round_up(count * sizeof(T) + 16, 256);
    CK(cudaMalloc(&p, bytes));
    allocs.push_back(p);
    device_bytes += bytes;
    if (zero) {
      CK(cudaMemsetAsync(p, 0, bytes, stream));
      state_allocs.emplace_back(p, bytes);   // render state: cleared again by bear_cuda_reset
    }
    return static_cast<T *>(p);
  }

  void dfree(void *p)
  {
    allocs.erase(std::remove(allocs.begin(), allocs.end(), p), allocs.end());
    for (auto it = state_allocs.begin(); it != state_allocs.end(); ++it)
      if (it->first == p) { device_bytes -= it->second; state_allocs.erase(it); break; }
    cudaFree(p);
  }

  ~bear_cuda_renderer()
  {
    cudaSetDevice(device);
    if (stream) cudaStreamSynchronize(stream);
    if (conv_stream) cudaStreamSynchronize(conv_stream);
    for (void *p : allocs) cudaFree(p);
    for (auto &c : ev_pending)
      for (cudaEvent_t e : c) cudaEventDestroy(e);
    for (cudaEvent_t e : ev_free) cudaEventDestroy(e);
    if (comm) nccl().CommDestroy(comm);
    for (cudaEvent_t e : mix_events) cudaEventDestroy(e);
    if (ev_call) cudaEventDestroy(ev_call);
    for (cudaEvent_t e : ev_barrier) if (e) cudaEventDestroy(e);
    if (conv_stream) cudaStreamDestroy(conv_stream);
    for (int w = 0; w < kMaxRanks; w++)
      if (peer_bus[w][0] && w != lay.rank) cudaIpcCloseMemHandle(peer_bus[w][0]);
    if (h_err) cudaFreeHost(h_err);
    if (ev_order) cudaEventDestroy(ev_order);
    if (d_frames) cudaFree(d_frames);
    if (d_tracks) cudaFree(d_tracks);
    if (graph_exec) cudaGraphExecDestroy(graph_exec);
    if (graph) cudaGraphDestroy(graph);
    if (h_graph) cudaFreeHost(h_graph);
    if (h_in) cudaFreeHost(h_in);
    if (h_out) cudaFreeHost(h_out);
    for (cudaEvent_t e : h2d_events) cudaEventDestroy(e);
    for (cudaEvent_t e : done_events) cudaEventDestroy(e);
    if (copy_stream) cudaStreamDestroy(copy_stream);
    if (d2h_stream) cudaStreamDestroy(d2h_stream);
    if (stream) cudaStreamDestroy(stream);
  }
};

namespace {

int ilog2(uint32_t v)
{
  int l = 0;
  while ((1u << l) < v) l++;
  return l;
}

void fill_device(bear_cuda_renderer *r, float *dst, size_t count, float value)
{
  std::vector<float> h(count, value);
  CK(cudaMemcpyAsync(dst, h.data(), count * sizeof(float), cudaMemcpyHostToDevice, r->stream));
  CK(cudaStreamSynchronize(r->stream));
}

// Initial state of the DSP graph (dsp.cpp:24-29,41-46,137-144; panner.cpp:172-186): gains 0 (buffers are
// zeroed), object / DirectSpeakers delays = default direct delay, static delays = default static delay,
// view 0.
// The tables of "the period before the first one" in the fixed-offset host buffer of the graph path.
void init_stream_shadow(bear_cuda_renderer *r)
{
  if (!r->h_graph) return;
  Layout &p = r->lay;
  std::fill(r->h_graph, r->h_graph + r->stream_floats, 0.0f);
  auto fill = [&](float *dev, float v) {
    for (auto &sl : r->stream_slots)
      if (sl.dev == dev) std::fill(r->h_graph + sl.off, r->h_graph + sl.off + sl.count, v);
  };
  fill(p.d_dir, (float)r->default_direct_delay);
  fill(p.d_ds, (float)r->default_direct_delay);
  fill(p.d_stat, (float)r->default_static_delay);
  r->shadow_valid = true;   // view 0 = all-zero bits, gains 0
}

void set_initial_state(bear_cuda_renderer *r)
{
  Layout &p = r->lay;
  if (p.N) fill_device(r, p.d_dir, 2 * p.N, (float)r->default_direct_delay);
  if (p.M) fill_device(r, p.d_ds, 2 * p.M, (float)r->default_direct_delay);
  fill_device(r, p.d_stat, 2 * 2 * p.L, (float)r->default_static_delay);
  r->tables_on_device = false;
  init_stream_shadow(r);
}

std::shared_ptr<FilterBank> build_bank(const bear_cuda_config &cfg, const bear_cuda_data &d)
{
  const uint32_t B = cfg.period_size;
  if (B == 0) throw InvalidArgument{"Config: period size must be set"};  // api.cpp:70
  if (!d.brirs || !d.delays || !d.decorrelation_filters || !d.hoa_irs)
    throw InvalidArgument{"data tables must not be NULL"};
  if ((B & (B - 1)) != 0 || B < 128 || B > 4096)
    throw Unsupported{"period size must be a power of two between 128 and 4096"};
  if (d.num_virtual_loudspeakers == 0 || d.num_virtual_loudspeakers > (uint32_t)kMaxSpeakers)
    throw Unsupported{"number of virtual loudspeakers must be 1..24"};
  if (d.decorrelator_length > B || d.hoa_ir_length > B)
    throw Unsupported{"decorrelator and HOA decode filters must not be longer than the period"};
  if (d.num_views == 0 || d.brir_length == 0) throw InvalidArgument{"empty BRIR table"};
  if (d.front_loudspeaker >= d.num_virtual_loudspeakers)
    throw InvalidArgument{"front_loudspeaker out of range"};

  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    throw CudaFail{std::string("no usable CUDA device (") + cudaGetErrorString(e) +
                   "); this library has no CPU fallback"};
  if (cfg.device < 0 || cfg.device >= ndev) throw InvalidArgument{"device ordinal out of range"};

  auto bank = std::make_shared<FilterBank>();
  FilterBank &k = *bank;
  k.device = cfg.device;
  CK(cudaSetDevice(k.device));
  CK(configure_kernels((int)B));
  k.B = (int)B; k.logB = ilog2(B);
  k.P = (int)((d.brir_length + B - 1) / B);
  k.L = (int)d.num_virtual_loudspeakers; k.A = (int)d.num_hoa_decoder_channels; k.V = (int)d.num_views;
  k.fs = d.fs; k.dec_delay = d.decorrelation_delay; k.hoa_delay = d.hoa_delay; k.front = (int)d.front_loudspeaker;
  // Longest table delay the data file can produce (panner.cpp:189-235): the largest loudspeaker delay
  // plus the decorrelation delay, or the HOA delay.
  float dm = 0.f;
  for (size_t i = 0; i < (size_t)d.num_views * d.num_virtual_loudspeakers * 2; i++) dm = std::max(dm, d.delays[i]);
  k.auto_max_delay = (uint32_t)std::ceil(std::max((double)dm + d.decorrelation_delay, d.hoa_delay)) + 4;
  // initial delays (panner.cpp:172-186)
  const double front = ((double)d.delays[(0 * k.L + d.front_loudspeaker) * 2 + 0] +
                        (double)d.delays[(0 * k.L + d.front_loudspeaker) * 2 + 1]) / 2.0;
  k.default_static_delay = front / d.fs;
  k.default_direct_delay = (front + d.decorrelation_delay) / d.fs;

  cudaStream_t st = nullptr;
  CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{st};
  {  // twiddles, computed in double: [0, B) exp(-i pi q / B); [B, 2B) the per-pass tables of the
     // radix-8 FFT passes, exp(-2 pi i r j / (8 Ns)) at B + (Ns - 8) + (r - 1) Ns + j  (kernels.cu fft_shared)
    std::vector<float2> tw(2 * (size_t)B, make_float2(0.f, 0.f));
    const double pi = 3.14159265358979323846264338327950288;
    for (uint32_t q = 0; q < B; q++) tw[q] = make_float2((float)std::cos(pi * q / B), (float)-std::sin(pi * q / B));
    for (uint32_t Ns = 8; Ns * 8 <= B; Ns *= 8)
      for (uint32_t r = 1; r < 8; r++)
        for (uint32_t j = 0; j < Ns; j++) {
          const double a = 2.0 * pi * (double)r * (double)j / (8.0 * Ns);
          tw[B + (Ns - 8) + (r - 1) * Ns + j] = make_float2((float)std::cos(a), (float)-std::sin(a));
        }
    float2 *dtw = k.alloc<float2>(2 * (size_t)B);
    CK(cudaMemcpyAsync(dtw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    k.tw = dtw;
  }
  {  // data-file delay table and optional gain-normalisation factors (control tail)
    const size_t nd = (size_t)k.V * k.L * 2;
    float *dt = k.alloc<float>(nd);
    CK(cudaMemcpyAsync(dt, d.delays, nd * sizeof(float), cudaMemcpyHostToDevice, st));
    k.delays_tab = dt;
    if (d.gain_norm_factors && d.num_gain_norm_delays) {
      if (d.num_gain_norm_delays < 2) throw InvalidArgument{"gain_norm factors need at least 2 delay samples"};
      const size_t nf = (size_t)d.num_gain_norm_delays * k.V * 4 * k.L * k.L * 2;
      float *gf = k.alloc<float>(nf);
      CK(cudaMemcpyAsync(gf, d.gain_norm_factors, nf * sizeof(float), cudaMemcpyHostToDevice, st));
      k.gn_factors = gf;
      k.gnS = (int)d.num_gain_norm_delays;
    }
    CK(cudaStreamSynchronize(st));
  }
  {  // filter spectra, pre-scaled by 1/(2B)
    const float scale = 1.0f / (float)(2 * B);
    auto transform = [&](const float *host, size_t n_filters, int len, int parts) {
      float *tmp = nullptr;
      const size_t bytes = n_filters * len * sizeof(float);
      CK(cudaMalloc(&tmp, bytes));
      CK(cudaMemcpyAsync(tmp, host, bytes, cudaMemcpyHostToDevice, st));
      float2 *spec = k.alloc<float2>(n_filters * parts * B);
      const size_t slice = 16384;   // launch in slices to keep grid sizes moderate
      for (size_t f0 = 0; f0 < n_filters; f0 += slice) {
        size_t nf = std::min(slice, n_filters - f0);
        launch_filter_fft(tmp + f0 * len, (int)nf, len, (int)B, k.logB, parts, k.tw, spec + f0 * parts * B, scale, st);
        k.launches++;
      }
      cudaError_t le = cudaGetLastError();
      cudaError_t se = cudaStreamSynchronize(st);
      cudaFree(tmp);
      CK(le);
      CK(se);
      return spec;
    };
    k.h_brir = transform(d.brirs, (size_t)k.V * 2 * k.L, (int)d.brir_length, k.P);
    k.h_dec = transform(d.decorrelation_filters, (size_t)k.L, (int)d.decorrelator_length, 1);
    k.h_hoa = transform(d.hoa_irs, (size_t)std::max(1, k.A) * 2, (int)d.hoa_ir_length, 1);
  }
  return bank;
}

int cmac_splits_for(const bear_cuda_renderer *r, int n);
int object_splits_for(const bear_cuda_renderer *r, int n);

bear_cuda_renderer *create_with_bank(const bear_cuda_config &cfg, std::shared_ptr<FilterBank> bank)
{
  const FilterBank &k = *bank;
  const uint32_t B = cfg.period_size;
  if ((int)B != k.B) throw InvalidArgument{"period size differs from the shared filter bank's"};
  if (cfg.device != k.device) throw InvalidArgument{"device differs from the shared filter bank's"};
  if (cfg.gain_ramp_offset > 1 || cfg.delay_ramp_offset > 1 || cfg.xfade_ramp_offset > 1)
    throw InvalidArgument{"ramp offsets must be 0 or 1"};

  std::unique_ptr<bear_cuda_renderer> r(new bear_cuda_renderer);
  r->bank = bank;
  r->cfg = cfg;
  r->device = cfg.device;
  CK(cudaSetDevice(r->device));
  CK(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&r->copy_stream, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&r->d2h_stream, cudaStreamNonBlocking));
  r->result_stream = r->stream;

  Layout &p = r->lay;
  p.N = cfg.num_objects_channels; p.M = cfg.num_direct_speakers_channels; p.H = cfg.num_hoa_channels;
  p.B = B; p.logB = ilog2(B);
  p.P = k.P;
  p.L = k.L; p.A = k.A;
  p.CH = 2 * p.L + p.A;
  p.V = k.V;
  p.Tmax = cfg.max_batch_blocks ? cfg.max_batch_blocks : 1;
  // max_delay_samples == 0: the longest table delay the data file can produce
  const uint32_t max_delay = cfg.max_delay_samples ? cfg.max_delay_samples : k.auto_max_delay;
  p.hist = (int)round_up(max_delay + 8, 4);
  if (mix_hist_too_large(p.hist)) throw Unsupported{"max_delay_samples too large for the mix kernel's shared-memory tile"};
  CK(configure_mix(p.hist, p.B));
  p.goff = cfg.gain_ramp_offset; p.doff = cfg.delay_ramp_offset; p.xoff = cfg.xfade_ramp_offset;
  p.fs = (float)k.fs;
  p.hoa_delay = (float)k.hoa_delay;
  if (k.hoa_delay + 4 > p.hist) throw Unsupported{"hoa delay exceeds max_delay_samples"};

  const int C = p.N + p.M;
  if (C > kMaxSplitChannels * kMaxObjectSplits) throw Unsupported{"more than 131072 object + DirectSpeakers channels"};
  r->s_cfg = (int)cfg.object_splits;
  p.S = 1;
  p.W = 1; p.rank = 0; p.l0 = 0; p.l1 = p.L; p.dif = p.N > 0;
  {
    // cMAC work units: one per loudspeaker (P partition terms each) + one for the HOA decode terms
    const int nunits = p.L + (p.H > 0 ? 1 : 0);
    const int threads = std::min(p.B / 2, 128);
    const int base = (p.B / 2 / threads) * 2 * (p.Tmax + 1);
    int KS = cfg.cmac_splits;
    if (KS == 0) KS = (4 * 148 + base - 1) / base;
    p.KS = std::max(1, std::min(KS, nunits));
    r->ks_cfg = p.KS;
    r->ks_units = nunits;
  }
  p.R = p.Tmax + p.P;
  r->chunk = (int)cfg.chunk_blocks;
  // host-buffer path: upload in ~12 chunks (at least 16 periods each) so copies overlap the kernels
  // host-buffer paths upload in chunks: chunk_blocks periods if configured, else sized per call by
  // host_chunk_periods() (at least 16 periods)
  r->h2d_chunk = cfg.chunk_blocks ? (int)cfg.chunk_blocks : 0;
  r->h2d_events.resize((p.Tmax + (r->h2d_chunk ? r->h2d_chunk : 16) - 1) / (r->h2d_chunk ? r->h2d_chunk : 16));
  r->done_events.resize(r->h2d_events.size());
  for (cudaEvent_t &e : r->h2d_events) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (cudaEvent_t &e : r->done_events) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));

  // ---- allocations -----------------------------------------------------------------------
  const size_t TB = (size_t)p.Tmax * p.B;
  p.xs = round_up(p.hist + TB, 4);
  p.xin = r->dalloc<float>((size_t)std::max(1, p.N + p.M + p.H) * p.xs);
  const size_t T1 = p.Tmax + 1;
  p.g_dir = r->dalloc<float>(T1 * p.L * std::max(1, p.N));
  p.g_dif = r->dalloc<float>(T1 * p.L * std::max(1, p.N));
  p.g_ds = r->dalloc<float>(T1 * p.L * std::max(1, p.M));
  p.g_hoa = r->dalloc<float>(T1 * std::max(1, p.A) * std::max(1, p.H));
  p.d_dir = r->dalloc<float>(T1 * 2 * std::max(1, p.N));
  p.d_ds = r->dalloc<float>(T1 * 2 * std::max(1, p.M));
  p.d_stat = r->dalloc<float>((T1 + 1) * 2 * p.L);
  p.view = r->dalloc<int32_t>(T1);
  p.cs = round_up(p.B + TB, 4);
  // both bus-row sets in one allocation: a renderer that shares a scene alternates between them from
  // call to call (its peers read the rows over NVLink), a renderer on its own uses the first
  r->bus[0] = r->dalloc<float>((size_t)3 * p.L * p.cs);   // (bear_cuda_comm_blob replaces it by kBusSets sets)
  p.cdir = r->bus[0];
  p.qdif = r->bus[0] + (size_t)2 * p.L * p.cs;
  p.peer_cdir[0] = p.cdir; p.peer_qdif[0] = p.qdif;
  p.cin = p.cdir; p.qin = p.qdif;
  {  // partial sums of the object splits of one launch, [S - 1][3L][n B]; sized for the worst launch
    size_t cap = 1;
    for (int n = 1; n <= p.Tmax; n++) cap = std::max(cap, (size_t)(object_splits_for(r.get(), n) - 1) * n);
    p.bps = 0;   // row stride, set per call
    p.bus_part = r->dalloc<float>((size_t)3 * p.L * cap * p.B, false);
  }
  p.mix_recs = r->dalloc<float>((size_t)p.Tmax * std::max(1, C) * kMixRecWords, false);
  p.mix_keys = r->dalloc<unsigned>((size_t)p.Tmax * std::max(1, C), false);
  p.mix_order = r->dalloc<unsigned>((size_t)p.Tmax * std::max(1, C), false);
  p.mix_counts = r->dalloc<int>((size_t)p.Tmax * kMaxObjectSplits * 2, false);
  p.err_flag = r->dalloc<int>(4);
  CK(cudaMallocHost(&r->h_err, 4 * sizeof(int)));
  r->h_err[0] = 0;
  CK(cudaEventCreateWithFlags(&r->ev_order, cudaEventDisableTiming));
  {  // the library's built-in scratch array, looked up once per renderer with its device current
    float2 *ss = nullptr;
    CK(static_scratch_pointer(&ss));
    r->static_scratch = ss;
  }
  p.ws = round_up(p.B + p.hist + TB, 4);
  p.w = r->dalloc<float>((size_t)p.L * p.ws);
  p.amb = r->dalloc<float>((size_t)std::max(1, p.A) * p.ws);
  p.fdl = r->dalloc<float2>((size_t)p.CH * p.R * p.B);
  p.grp_sorted = r->dalloc<int>((size_t)p.Tmax + 1);
  p.grp_items = r->dalloc<int2>((size_t)p.Tmax + 1);
  p.grp_count = r->dalloc<int>(2);
  {  // partial spectra of one launch, [periods of the launch][2][2][splits][B]; sized for the worst launch
    size_t cap = 0;
    for (int n = 1; n <= p.Tmax; n++) cap = std::max(cap, (size_t)n * cmac_splits_for(r.get(), n));
    p.ypart = r->dalloc<float2>(cap * 2 * 2 * p.B);
  }
  p.os = round_up(TB, 4);
  p.out = r->dalloc<float>(2 * p.os);
  r->audio_out = p.out;

  // control tail: shared tables + this renderer's raw-gain staging
  p.delays_tab = k.delays_tab;
  p.gn_factors = k.gn_factors;
  p.gnS = k.gnS;
  p.raw_obj = r->dalloc<float>((size_t)p.Tmax * 2 * p.L * std::max(1, p.N));
  p.raw_ds = r->dalloc<float>((size_t)p.Tmax * p.L * std::max(1, p.M));
  p.fs_d = k.fs;
  p.dec_delay_d = k.dec_delay;
  p.front = k.front;
  p.tw = k.tw;
  p.h_brir = k.h_brir;
  p.h_dec = k.h_dec;
  p.h_hoa = k.h_hoa;

  r->default_static_delay = k.default_static_delay;
  r->default_direct_delay = k.default_direct_delay;
  if (k.default_direct_delay * k.fs + 4 > p.hist) throw Unsupported{"default delay exceeds max_delay_samples"};
  set_initial_state(r.get());

  r->h_in_floats = (size_t)std::max(1, p.N + p.M + p.H) * p.B + (size_t)2 * p.L * p.N + 2 * p.N + (size_t)p.L * p.M +
                   2 * p.M + 2 * p.L + (size_t)p.A * p.H + 8;
  CK(cudaMallocHost(&r->h_in, r->h_in_floats * sizeof(float)));
  r->d_pack = r->dalloc<float>(r->h_in_floats);
  CK(cudaMallocHost(&r->h_out, 2 * (size_t)p.B * sizeof(float)));
  {
    size_t off = (size_t)(p.N + p.M + p.H) * p.B;
    auto slot = [&](float *dev, size_t per_block, size_t sl) {
      if (per_block == 0) return;
      r->stream_slots.push_back({off, per_block, dev, sl});
      off += per_block;
    };
    slot(p.g_dir, (size_t)p.L * p.N, 1);
    slot(p.g_dif, (size_t)p.L * p.N, 1);
    slot(p.d_dir, (size_t)2 * p.N, 1);
    slot(p.g_ds, (size_t)p.L * p.M, 1);
    slot(p.d_ds, (size_t)2 * p.M, 1);
    slot(p.d_stat, (size_t)2 * p.L, 2);
    slot(p.g_hoa, (size_t)p.A * p.H, 1);
    slot(reinterpret_cast<float *>(p.view), 1, 1);
    r->stream_floats = off;
    CK(cudaMallocHost(&r->h_graph, std::max<size_t>(off, 1) * sizeof(float)));
    init_stream_shadow(r.get());
  }
  CK(cudaStreamSynchronize(r->stream));

  r->kernels_per_call = 4 * (C > 0) + (p.H > 0) + (p.N > 0) + 4;   // long launch: one object split, ungrouped cMAC
  r->error.reserve(512);   // so that reporting an argument error allocates nothing
  return r.release();
}

bear_cuda_renderer *create_impl(const bear_cuda_config &cfg, const bear_cuda_data &d)
{
  return create_with_bank(cfg, build_bank(cfg, d));
}

// cMAC splits of a launch of n periods: the configured number, or as many as it takes to put about
// four CTAs on every SM when the launch is short (a single period is otherwise 4 CTAs that walk all
// 24 loudspeakers x P partitions one after the other).
int cmac_splits_for(const bear_cuda_renderer *r, int n)
{
  const Layout &p = r->lay;
  if (!r->ks_dynamic) return r->ks_cfg;
  const int threads = std::min(p.B / 2, 128);
  if (cmac_groups_by_view(p, n)) {
    // grouped launch: an item is up to kCmacGroup periods of one view.  How many items there are
    // depends on the views (device data): this is the upper bound (fewest possible items, ~12 CTAs
    // per SM) that sizes the grid and the scratch; view_group_kernel picks the number actually used
    const int base = (p.B / 2 / threads) * 2 * ((n + kCmacGroup) / kCmacGroup);
    return std::max(r->ks_cfg, std::min((12 * 148 + base - 1) / base, r->ks_units));
  }
  const int base = (p.B / 2 / threads) * 2 * (n + 1);
  return std::max(r->ks_cfg, std::min((4 * 148 + base - 1) / base, r->ks_units));
}

// Object splits of a launch of n periods: mix_kernel runs one CTA per (tile of kMixTile samples, split),
// each walking its share of the channels; a launch with few tiles splits the channels so that there
// is about one wave of CTAs (5 per SM) -- a 16-period upload chunk of a 1024-object scene is
// otherwise 32 CTAs that each walk 1024 channels.  The partial sums cost one small extra kernel.
int object_splits_for(const bear_cuda_renderer *r, int n)
{
  const Layout &p = r->lay;
  const int C = p.N + p.M;
  const int maxS = std::min(kMaxObjectSplits, C > 0 ? (C + 2 * kMixChunk - 1) / (2 * kMixChunk) : 1);
  const int minS = (C + kMaxSplitChannels - 1) / kMaxSplitChannels;   // a split is sorted in shared memory
  if (r->s_cfg) return std::max(std::max(1, minS), std::min(r->s_cfg, maxS));
  const int tiles = std::max(1, n * p.B / kMixTile);
  // (long launches keep one split: measured at 1024 objects x 1872 periods, two splits cost more in partial
  // sums and per-CTA set-up than their fuller last wave of CTAs gains -- 2.09 against 2.03 ms)
  return std::max(std::max(1, minS), std::min((5 * 148 + tiles / 2) / tiles, maxS));   // about one wave of CTAs
}
// Periods per upload chunk of a host-buffer call whose period takes bytes_per_period on the bus.
// A chunk of any size needs ~0.15 ms to pass through the six kernels, so a chunk should take at least
// that long to upload (~11 MB at the 55 GB/s measured on PCIe 5 x16), or the kernels, not the bus,
// pace the call.
static int host_chunk_periods(const bear_cuda_renderer *r, int T, size_t bytes_per_period)
{
  if (r->h2d_chunk) return r->h2d_chunk;
  // ranks that share a scene must cut a call into the same chunks (one barrier per chunk)
  if (r->lay.W > 1) bytes_per_period = r->comm_bytes_per_period;
  const size_t want = (11u << 20) / std::max<size_t>(bytes_per_period, 1) + 1;
  return (int)std::min<size_t>((size_t)T, std::max<size_t>(16, want));
}
static size_t table_bytes_per_period(const Layout &p)
{
  return sizeof(float) * ((size_t)2 * p.L * p.N + 2 * p.N + (size_t)p.L * p.M + 2 * p.M + 2 * p.L + (size_t)p.A * p.H + 1);
}

constexpr unsigned all_ranks = 0xffu;   // (masked by W in the kernel)
// A synchronisation point of a shared scene: which counter array (offset in ints), which value, whom to
// tell and whom to wait for.
PeerSync peer_sync(const bear_cuda_renderer *r, int array, int value, unsigned signal, unsigned wait)
{
  PeerSync s{};
  for (int w = 0; w < r->lay.W; w++) s.flags[w] = reinterpret_cast<int *>(r->peer_base[w] + r->flags_off) + array;
  s.W = r->lay.W; s.rank = r->lay.rank; s.value = value; s.signal = signal; s.wait = wait;
  s.err = r->lay.err_flag;
  return s;
}

// After pipelined calls the convolution side may still be running on conv_stream: every other entry
// point first makes the render stream wait for it, and is then what it always was.
void join_pipelined(bear_cuda_renderer *r)
{
  if (!r->pipelined_pending) return;
  CK(cudaEventRecord(r->ev_call, r->conv_stream));
  CK(cudaStreamWaitEvent(r->stream, r->ev_call, 0));
  r->pipelined_pending = false;
  r->result_stream = r->stream;
}

// Queue the kernels for T blocks whose audio and tables are already in the staging areas.
// A call may be issued in chunks of `chunk` periods (all kernels of chunk c, then chunk c+1, ...;
// chunk <= 0: all at once): every intermediate buffer is indexed by the period number within the
// call, so later chunks just read earlier columns.  `before_chunk(c, k0, n)` runs on the host before
// chunk c is queued (the host-buffer path uses it to make the stream wait for that chunk's upload),
// `after_chunk(c, k0, n)` after it.  Nothing here allocates unless per-kernel profiling is on.
template <class Before, class After>
void run_pipeline(bear_cuda_renderer *r, int T, int chunk, Before &&before_chunk, After &&after_chunk,
                  bool device_tables = false, bool pipelined = false)
{
  Layout &p = r->lay;
  cudaStream_t st = r->stream;
  // pipelined: the convolution side of this call is left running when the call returns (one chunk, no HOA
  // bus on this rank -- its history would be shared by both sides --, no per-kernel profiling)
  const bool pipe = pipelined && p.W > 1 && p.H == 0 && !r->profiling && (chunk <= 0 || chunk >= T);
  if (!pipe) join_pipelined(r);
  std::vector<cudaEvent_t> chain;
  std::vector<int> ids;
  auto mark = [&](int kernel_id, int side = 0) {
    if (pipe && !r->timeline_path.empty()) {
      cudaEvent_t e;
      if (!r->timeline_base) { CK(cudaEventCreate(&r->timeline_base)); CK(cudaEventRecord(r->timeline_base, st)); }
      if (r->timeline.size() >= 800) {
        for (size_t i = 0; i < 400; i++) r->ev_free.push_back(r->timeline[i].e);
        r->timeline.erase(r->timeline.begin(), r->timeline.begin() + 400);
      }
      if (r->ev_free.empty()) CK(cudaEventCreate(&e));
      else { e = r->ev_free.back(); r->ev_free.pop_back(); }
      CK(cudaEventRecord(e, side ? r->conv_stream : st));
      r->timeline.push_back({e, kernel_id, side, r->shared_calls});
      return;
    }
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
  // A shared scene that is issued in chunks (chunk_blocks, the upload chunks of the host-buffer calls, or
  // BEAR_CUDA_SHARED_CHUNKS) runs the exchange (ordering collective, gather of the peers' rows over
  // NVLink) and the convolution of chunk c on a second stream under the mix of chunk c+1 -- the mix keeps
  // the SMs busy, the gather the NVLink ports.  A device-resident call is NOT cut into chunks by default:
  // measured at 1024 objects on 2 and 8 GPUs, every kernel loses more to the shorter launches (and the
  // convolution to fewer periods per BRIR view) than the overlap returns (1.71 -> 1.78 / 2.00 ms for 2 / 4
  // chunks on 2 GPUs; 0.847 -> 0.845 / 0.987 ms on 8).  (Per-kernel profiling keeps everything on one
  // stream so that its event pairs mean something.)
  if (shared_scene && chunk == T && T >= 64 && r->shared_chunks > 1)
    chunk = (int)round_up((size_t)(T + r->shared_chunks - 1) / r->shared_chunks, 8);
  const int nchunks_total = (T + chunk - 1) / chunk;
  const bool overlap = pipe || (shared_scene && !r->profiling && r->conv_stream && nchunks_total > 1 &&
                                nchunks_total <= (int)r->mix_events.size());
  cudaStream_t sb = overlap ? r->conv_stream : st;   // exchange + convolution side
  r->result_stream = sb;
  Layout pc = p;   // what the convolution side sees (filled in below)
  // bus rows of this call; a renderer that shares a scene cycles through kBusSets sets, so that a peer
  // still reading an earlier call's rows is never overwritten (one barrier per chunk suffices)
  float *const bus = r->bus[r->parity];
  p.cdir = bus;
  p.qdif = bus + (size_t)2 * p.L * p.cs;
  for (int w = 0; w < p.W; w++) {
    const float *pb = shared_scene ? r->peer_bus[w][r->parity] : bus;
    p.peer_cdir[w] = pb;
    p.peer_qdif[w] = pb + (size_t)2 * p.L * p.cs;
  }
  const float *in = shared_scene ? r->bus_sum : bus;
  p.cin = in;
  p.qin = in + (size_t)2 * p.L * p.cs;
  p.S = object_splits_for(r, chunk);     // one value per call
  p.bps = (size_t)chunk * p.B;
  uint64_t launches = 0;
  if (device_tables) { launch_validate_views(p, T, st); launches++; }
  if (pipe) {
    // the bus rows this call writes were read by the peers two calls ago: their gathers are over once the
    // ordering collective of the previous call has completed here
    // (three sets: the set written now was last read in call c-3, and every peer had finished that
    // gather when it entered the ordering collective of call c-2 -- the event recorded two calls ago,
    // long complete, so the mix never waits for its peers)
    CK(cudaStreamWaitEvent(st, r->ev_barrier[r->shared_calls & 1], 0));
    // the convolution side works on its own copies of the tables it reads, so that the caller may write
    // the next call's tables as soon as this call has been queued
    const int par = r->parity;
    CK(cudaMemcpyAsync(r->sh_view[par], p.view, (size_t)(T + 1) * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(r->sh_dstat[par], p.d_stat, (size_t)(T + 2) * 2 * p.L * sizeof(float), cudaMemcpyDeviceToDevice, st));
  } else if (overlap) {   // the convolution side starts after everything queued so far (the previous call's carry)
    CK(cudaEventRecord(r->ev_call, st));
    CK(cudaStreamWaitEvent(sb, r->ev_call, 0));
  }
  int nchunks = 0;
  for (int k0 = 0; k0 < T; nchunks++) {
    const int n = std::min(chunk, T - k0);
    before_chunk(nchunks, k0, n);
    p.k0 = k0;
    if (p.N + p.M > 0) {
      mark(BEAR_CUDA_KERNEL_MIX_AUX); launch_mix_prep(p, n, st);
      mark(BEAR_CUDA_KERNEL_MIX); launch_mix(p, n, st);
      launches += 4;   // prep + order, mix + dense
      if (p.S > 1) { mark(BEAR_CUDA_KERNEL_MIX_AUX); launch_bus_sum(p, n, st); launches++; }
    }
    if (p.H > 0) { mark(BEAR_CUDA_KERNEL_HOA_MIX); launch_hoa_mix(p, n, st); launches++; }
    if (overlap) {
      CK(cudaEventRecord(r->mix_events[nchunks], st));
      CK(cudaStreamWaitEvent(sb, r->mix_events[nchunks], 0));
    }
    pc = p;
    if (pipe) { pc.view = r->sh_view[r->parity]; pc.d_stat = r->sh_dstat[r->parity]; }
    if (shared_scene) {
      // every rank's bus rows of this chunk are complete before anyone reads them
      mark(BEAR_CUDA_KERNEL_EXCHANGE, 1);
      launch_peer_sync(peer_sync(r, 0, ++r->bar_count, all_ranks, all_ranks), sb);
      launches++;
      if (pipe) CK(cudaEventRecord(r->ev_barrier[r->shared_calls & 1], sb));
      mark(BEAR_CUDA_KERNEL_GATHER, 1);
      launch_bus_gather(pc, n, sb, pipe && p.W > 4);   // (measured: the thin form pays from 8 ranks; at 2 and 4 the wide one is faster)
      if (p.l1 > p.l0) launches++;
    }
    if (p.dif && p.l1 > p.l0) { mark(BEAR_CUDA_KERNEL_DECORRELATE, 1); launch_decorrelate(pc, n, sb); launches++; }
    if (p.l1 > p.l0 || p.H > 0) { mark(BEAR_CUDA_KERNEL_BUS_FDL, 1); launch_bus_fdl(pc, n, sb); launches++; }
    pc.KS = cmac_splits_for(r, n);
    pc.ks_dev = cmac_groups_by_view(pc, n);
    mark(BEAR_CUDA_KERNEL_CMAC, 1); launches += launch_cmac(pc, n, sb);
    // with a receiving rank, every rank stores its stereo partial mix straight into that rank's inbox
    // (ordinary stores over NVLink), tells it so, and the receiver adds the W partial mixes up
    const bool to_inbox = shared_scene && r->comm_root >= 0;
    if (to_inbox) pc.out = r->peer_base[r->comm_root] + r->inbox_off + (size_t)p.rank * 2 * p.os;
    mark(BEAR_CUDA_KERNEL_IFFT_OUT, 1); launch_ifft_out(pc, n, sb); launches++;
    if (to_inbox) {
      mark(BEAR_CUDA_KERNEL_EXCHANGE, 1);
      const bool receiver = p.rank == r->comm_root;
      launch_peer_sync(peer_sync(r, 32, ++r->out_count, 1u << r->comm_root, receiver ? all_ranks : 0u), sb);
      launches++;
      if (receiver) {
        launch_out_sum(p.out, r->bus[0] + r->inbox_off, p.os, p.W, (size_t)k0 * p.B, (size_t)n * p.B, sb);
        launches++;
      }
    } else if (shared_scene) {
      // no receiving rank: the stereo partial mixes are all-reduced (north_star: "only per-block stereo
      // partial mixes are summed with NCCL"), chunk by chunk so that the sum of chunk c runs under chunk c+1
      mark(BEAR_CUDA_KERNEL_EXCHANGE, 1);
      NCK(nccl().GroupStart());
      for (int e = 0; e < 2; e++) {
        float *row = p.out + (size_t)e * p.os + (size_t)k0 * p.B;
        NCK(nccl().AllReduce(row, row, (size_t)n * p.B, kNcclFloat32, kNcclSum, r->comm, sb));
      }
      NCK(nccl().GroupEnd());
    }
    after_chunk(nchunks, k0, n);
    k0 += n;
  }
  p.k0 = 0;
  if (overlap && !pipe) {   // the carry moves state both sides use: it follows both
    CK(cudaEventRecord(r->ev_call, sb));
    CK(cudaStreamWaitEvent(st, r->ev_call, 0));
  }
  mark(BEAR_CUDA_KERNEL_CARRY);

  // Two groups of state: what the mix side reads (input history, the tables' last period) and what the
  // convolution side reads (bus prefix, decorrelator outputs, delay line).  Normally one kernel moves
  // both; a pipelined call moves each on its own stream.
  CarryPlan plan{}, plan_conv{};
  int n = 0, rows = 0, nc = 0, rows_c = 0;
  auto add_to = [&](CarryPlan &pl, int &cnt, int &nrows_total, float *base, int nrows, long long stride, int count, long long shift) {
    if (nrows <= 0 || count <= 0) return;
    constexpr int piece = 1024;
    if (nrows == 1 && shift >= count && count > 2 * piece) {
      // one long run whose source and destination do not overlap (a table's last slot): cut into pieces,
      // one CTA each, instead of one CTA walking it 256 floats at a time
      nrows = (count + piece - 1) / piece;
      pl.e[cnt] = CarryRow{base, nrows, piece, piece, shift, count};
    } else {
      pl.e[cnt] = CarryRow{base, nrows, stride, count, shift, 0};
    }
    pl.row_start[cnt] = nrows_total;
    nrows_total += nrows;
    cnt++;
  };
  auto add = [&](float *base, int nrows, long long stride, int count, long long shift) {
    add_to(plan, n, rows, base, nrows, stride, count, shift);
  };
  auto add_conv = [&](float *base, int nrows, long long stride, int count, long long shift) {
    if (pipe) add_to(plan_conv, nc, rows_c, base, nrows, stride, count, shift);
    else add_to(plan, n, rows, base, nrows, stride, count, shift);
  };
  const long long TB = (long long)T * p.B;
  add(p.xin, p.N + p.M + p.H, (long long)p.xs, p.hist, TB);
  if (shared_scene) {
    // the last period of the gathered rows is the overlap-save prefix of the next call (the rows the
    // peers read have no history: every call's columns are gathered once)
    add_conv(r->bus_sum + (size_t)2 * p.l0 * p.cs, 2 * (p.l1 - p.l0), (long long)p.cs, p.B, TB);
    if (p.dif) add_conv(r->bus_sum + (size_t)(2 * p.L + p.l0) * p.cs, p.l1 - p.l0, (long long)p.cs, p.B, TB);
  } else {
    if (p.N + p.M > 0) add(bus, 2 * p.L, (long long)p.cs, p.B, TB);
    if (p.N > 0) add(bus + (size_t)2 * p.L * p.cs, p.L, (long long)p.cs, p.B, TB);
  }
  if (p.dif) add_conv(p.w, p.L, (long long)p.ws, p.B + p.hist, TB);
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
  CarryPlan &fdl_plan = pipe ? plan_conv : plan;
  fdl_plan.fdl = p.fdl;
  fdl_plan.fdl_stride = (long long)p.R * p.B;
  fdl_plan.fdl_channels = p.CH;
  fdl_plan.fdl_B = p.B;
  fdl_plan.fdl_P = p.P;
  fdl_plan.fdl_T = T;
  launch_carry(plan, st);
  launches++;
  if (pipe) {
    plan_conv.n = nc;
    plan_conv.row_start[nc] = rows_c;
    mark(BEAR_CUDA_KERNEL_CARRY, 1);
    launch_carry(plan_conv, sb);
    mark(-1, 1);
    launches++;
    r->pipelined_pending = true;
  }
  mark(-1);
  if (r->profiling) {
    r->ev_pending.push_back(std::move(chain));
    r->ev_kernel.push_back(std::move(ids));
  }
  CK(cudaGetLastError());
  r->launches += launches;
  r->calls++;
  if (shared_scene) { r->parity = (r->parity + 1) % kBusSets; r->shared_calls++; }
}

void run_pipeline(bear_cuda_renderer *r, int T, bool device_tables = false, bool pipelined = false)
{
  run_pipeline(r, T, r->chunk, [](int, int, int) {}, [](int, int, int) {}, device_tables, pipelined);
}

// Error bits raised by the kernels of the calls queued so far (a table delay out of range, a view
// index out of range on the device-resident path); synchronises the render stream.
void dump_timeline(bear_cuda_renderer *r)
{
  if (r->timeline.empty()) return;
  static const char *names[] = {"mix", "hoa_mix", "decorrelate", "bus_fdl", "cmac", "ifft_out", "carry", "control",
                                "mix_aux", "exchange", "gather"};
  if (FILE *f = std::fopen((r->timeline_path + ".rank" + std::to_string(r->lay.rank)).c_str(), "a")) {
    std::fprintf(f, "# %zu marks: call side kernel-that-follows ms\n", r->timeline.size());
    for (auto &m : r->timeline) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, r->timeline_base, m.e) != cudaSuccess) { cudaGetLastError(); continue; }
      std::fprintf(f, "%llu %d %s %.4f\n", (unsigned long long)m.call, m.side, m.kernel < 0 ? "end" : names[m.kernel], ms);
    }
    std::fclose(f);
  }
  for (auto &m : r->timeline) r->ev_free.push_back(m.e);
  r->timeline.clear();
}

void fetch_kernel_errors(bear_cuda_renderer *r)
{
  join_pipelined(r);
  CK(cudaMemcpyAsync(r->h_err, r->lay.err_flag, sizeof(int), cudaMemcpyDeviceToHost, r->stream));
  CK(cudaStreamSynchronize(r->stream));
  dump_timeline(r);
  const int bits = r->h_err[0];
  if (!bits) return;
  CK(cudaMemsetAsync(r->lay.err_flag, 0, sizeof(int), r->stream));
  r->h_err[0] = 0;
  if (bits & kErrPeerTimeout) throw CudaFail{"shared scene: a peer GPU did not reach a synchronisation point within ~20 s "
                                             "(every rank must make the same sequence of process calls)"};
  if (bits & kErrViewRange) throw InvalidArgument{"brir_index out of range (device tables); view 0 was rendered instead"};
  throw InvalidArgument{"a table delay is negative or longer than history_samples - 4 samples (" +
                        std::to_string(r->lay.hist - 4) + "); it was clamped for this call -- create the renderer "
                        "with a larger max_delay_samples"};
}

template <class F>
int guarded(bear_cuda_renderer *r, F &&f, bool join = true)
{
  std::string *err = r ? &r->error : &g_create_error;
  try {
    if (r) {
      cudaError_t e = cudaSetDevice(r->device);
      if (e != cudaSuccess) throw CudaFail{std::string("cudaSetDevice: ") + cudaGetErrorString(e)};
      if (join) join_pipelined(r);   // no entry point but the pipelined one leaves or finds work on the second stream
    }
    f();
    return BEAR_CUDA_OK;
  } catch (const InvalidArgument &e) {
    *err = e.msg;
    return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  } catch (const Unsupported &e) {
    *err = e.msg;
    return BEAR_CUDA_ERR_UNSUPPORTED;
  } catch (const CudaFail &e) {
    *err = e.msg;
    return BEAR_CUDA_ERR_CUDA;
  } catch (const std::exception &e) {
    *err = e.what();
    return BEAR_CUDA_ERR_DATA_FILE;
  }
}

void upload_tables(bear_cuda_renderer *r, int T, const bear_cuda_block_tables *t, bool allow_unchanged,
                   cudaStream_t st)
{
  Layout &p = r->lay;
  auto up = [&](float *dev, size_t per_block, size_t slot0_blocks, const float *host, const char *name) {
    if (per_block == 0) return;
    float *dst = dev + slot0_blocks * per_block;
    if (host) {
      CK(cudaMemcpyAsync(dst, host, (size_t)T * per_block * sizeof(float), cudaMemcpyHostToDevice, st));
    } else if (allow_unchanged && T == 1) {
      CK(cudaMemcpyAsync(dst, dst - per_block, per_block * sizeof(float), cudaMemcpyDeviceToDevice, st));
    } else {
      throw InvalidArgument{std::string("table must not be NULL: ") + name};
    }
  };
  if (!t) {  // tables of this call were prepared on the device by bear_cuda_control_tables
    if (!r->tables_on_device) throw InvalidArgument{"tables must not be NULL (no bear_cuda_control_tables call pending)"};
    r->tables_on_device = false;
    return;
  }
  up(p.g_dir, (size_t)p.L * p.N, 1, t->direct_gains, "direct_gains");
  up(p.g_dif, (size_t)p.L * p.N, 1, t->diffuse_gains, "diffuse_gains");
  up(p.d_dir, (size_t)2 * p.N, 1, t->direct_delays, "direct_delays");
  up(p.g_ds, (size_t)p.L * p.M, 1, t->direct_speakers_gains, "direct_speakers_gains");
  up(p.d_ds, (size_t)2 * p.M, 1, t->direct_speakers_delays, "direct_speakers_delays");
  up(p.d_stat, (size_t)2 * p.L, 2, t->static_delays, "static_delays");
  up(p.g_hoa, (size_t)p.A * p.H, 1, t->hoa_gains, "hoa_gains");
  if (t->brir_index) {
    for (int k = 0; k < T; k++) {
      int32_t v = t->brir_index[k];
      if (allow_unchanged && T == 1 && v < 0) {
        CK(cudaMemcpyAsync(p.view + 1, p.view, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
        return;
      }
      if (v < 0 || v >= p.V) throw InvalidArgument{"brir_index out of range"};
    }
    CK(cudaMemcpyAsync(p.view + 1, t->brir_index, (size_t)T * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  } else if (allow_unchanged && T == 1) {
    CK(cudaMemcpyAsync(p.view + 1, p.view, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  } else {
    throw InvalidArgument{"table must not be NULL: brir_index"};
  }
}

// The tables Panner::Panner unpacks from a data file, with its checks (panner.cpp:52-137).
struct LoadedData {
  tenf::File tf;
  tenf::FloatTensor views, brirs, delays, dec, hoa, gn;
  bear_cuda_data d{};
  LoadedData(const char *path, bool need_tables) : tf(path)
  {
    (void)need_tables;
    const tenf::Json &m = tf.metadata();
    auto check = [](bool ok, const char *msg) {
      if (!ok) throw std::runtime_error(std::string("BEAR data file error: ") + msg);               // panner.cpp:43-46
    };
    if (const tenf::Json *v = m.find("bear_data_version")) {                                         // data_file.cpp:6-20
      if (v->kind != tenf::Json::Number || !v->is_int) throw std::runtime_error("version must be an integer");
      if (v->inum > 0)
        throw std::runtime_error("data file version mismatch: expected 0, got " + std::to_string(v->inum));
    }
    views = tf.unpack_f32(m.at("views")); brirs = tf.unpack_f32(m.at("brirs"));
    delays = tf.unpack_f32(m.at("delays")); dec = tf.unpack_f32(m.at("decorrelation_filters"));
    check(views.shape.size() == 2, "views must have 2 dimensions");                                  // panner.cpp:67-92
    check(brirs.shape.size() == 4, "brirs must have 4 dimensions");
    check(delays.shape.size() == 3, "delays must have 3 dimensions");
    check(dec.shape.size() == 2, "decorrelation_filters must have 2 dimensions");
    const size_t V = brirs.shape[0], L = brirs.shape[1];
    check(views.shape[0] == V, "views axis 0 is wrong size");
    check(views.shape[1] == 3, "views axis 1 is wrong size");
    check(brirs.shape[2] == 2, "brirs axis 2 is wrong size");
    check(delays.shape[0] == V, "delays axis 0 is wrong size");
    check(delays.shape[1] == L, "delays axis 1 is wrong size");
    check(delays.shape[2] == 2, "delays axis 2 is wrong size");
    check(dec.shape[0] == L, "decorrelation filters axis 0 is wrong size");
    if (const tenf::Json *q = m.find("gain_norm_quick")) {                                           // panner.cpp:113-123
      gn = tf.unpack_f32(q->at("factors"));
      check(gn.shape.size() == 5, "gain comp factors must have 5 dimensions");
      check(gn.shape[1] == V, "gain comp factors axis 1 is wrong size");
      check(gn.shape[2] == L * 2, "gain comp factors axis 2 is wrong size");
      check(gn.shape[3] == L * 2, "gain comp factors axis 3 is wrong size");
      check(gn.shape[4] == 2, "gain comp factors axis 4 is wrong size");
      d.gain_norm_factors = gn.data.data();
      d.num_gain_norm_delays = (uint32_t)gn.shape[0];
    }
    check(m.has("hoa"), "HOA decoder not present");                                                  // panner.cpp:125
    hoa = tf.unpack_f32(m.at("hoa").at("irs"));
    check(hoa.shape.size() == 3, "HOA decoder must have 3 dimensions");
    check(hoa.shape[1] == 2, "HOA decoder axis 1 is wrong size");
    const size_t order = (size_t)std::llround(std::sqrt((double)hoa.shape[0])) - 1;
    if ((order + 1) * (order + 1) != hoa.shape[0]) throw std::logic_error("bad number of hoa channels");

    d.num_views = (uint32_t)V; d.num_virtual_loudspeakers = (uint32_t)L;
    d.brir_length = (uint32_t)brirs.shape[3]; d.decorrelator_length = (uint32_t)dec.shape[1];
    d.num_hoa_decoder_channels = (uint32_t)hoa.shape[0]; d.hoa_ir_length = (uint32_t)hoa.shape[2];
    d.brirs = brirs.data.data(); d.delays = delays.data.data();
    d.decorrelation_filters = dec.data.data(); d.hoa_irs = hoa.data.data();
    d.fs = m.at("fs").num;
    d.decorrelation_delay = (double)(size_t)m.at("decorrelation_delay").num;                         // panner.cpp:62
    d.front_loudspeaker = (uint32_t)m.at("front_loudspeaker").num;
    const tenf::Json *hd = m.at("hoa").find("delay");
    d.hoa_delay = hd ? hd->num : 0.0;
  }
};

// On the ranks of a shared scene that do not receive the mix, the process calls take no output buffer.
bool output_optional(const bear_cuda_renderer *r) { return r->lay.W > 1 && r->comm_root >= 0 && r->comm_root != r->lay.rank; }

}  // namespace

// ================================================================================================
extern "C" {

int bear_cuda_create(const bear_cuda_config *config, const bear_cuda_data *data, bear_cuda_renderer **out)
{
  if (out) *out = nullptr;
  return guarded(nullptr, [&] {
    if (!config || !data || !out) throw InvalidArgument{"NULL argument"};
    *out = create_impl(*config, *data);
  });
}

int bear_cuda_create_shared(const bear_cuda_config *config, const bear_cuda_renderer *share_with, bear_cuda_renderer **out)
{
  if (out) *out = nullptr;
  return guarded(nullptr, [&] {
    if (!config || !share_with || !out) throw InvalidArgument{"NULL argument"};
    *out = create_with_bank(*config, share_with->bank);
  });
}

int bear_cuda_create_from_file(const bear_cuda_config *config, const char *data_path, bear_cuda_renderer **out)
{
  if (out) *out = nullptr;
  return guarded(nullptr, [&] {
    if (!config || !out) throw InvalidArgument{"NULL argument"};
    if (config->period_size == 0) throw InvalidArgument{"Config: period size must be set"};          // api.cpp:70
    if (!data_path || !*data_path) throw InvalidArgument{"Config: data path must be set"};           // api.cpp:71
    LoadedData ld(data_path, true);
    *out = create_impl(*config, ld.d);
  });
}

int bear_cuda_read_data_file_info(const char *data_path, bear_cuda_data_file_info *info)
{
  return guarded(nullptr, [&] {
    if (!data_path || !info) throw InvalidArgument{"NULL argument"};
    std::memset(info, 0, sizeof(*info));
    LoadedData ld(data_path, false);
    const bear_cuda_data &d = ld.d;
    info->num_views = d.num_views; info->num_virtual_loudspeakers = d.num_virtual_loudspeakers;
    info->brir_length = d.brir_length; info->decorrelator_length = d.decorrelator_length;
    info->num_hoa_decoder_channels = d.num_hoa_decoder_channels; info->hoa_ir_length = d.hoa_ir_length;
    info->front_loudspeaker = d.front_loudspeaker;
    info->fs = d.fs; info->decorrelation_delay = d.decorrelation_delay; info->hoa_delay = d.hoa_delay;
    // DataFileMetadata::read_from_file, api.cpp:330-353
    if (const tenf::Json *md = ld.tf.metadata().find("metadata")) {
      if (md->kind != tenf::Json::Object) throw std::runtime_error("data file metadata should be object");
      info->has_metadata = 1;
      const tenf::Json *label = md->find("label");
      if (!label) throw std::runtime_error("expected label");
      if (label->kind != tenf::Json::String) throw std::runtime_error("label should be a string");
      std::snprintf(info->label, sizeof(info->label), "%s", label->str.c_str());
      if (const tenf::Json *desc = md->find("description")) {
        if (desc->kind != tenf::Json::String) throw std::runtime_error("description should be a string");
        std::snprintf(info->description, sizeof(info->description), "%s", desc->str.c_str());
      }
      const tenf::Json *rel = md->find("released");
      if (!rel) throw std::runtime_error("expected released");
      if (rel->kind != tenf::Json::Bool) throw std::runtime_error("released should be a bool");
      info->is_released = rel->b;
    }
  });
}

int bear_cuda_tenf_read_f32(const char *path, const char *key, float *out, size_t capacity, size_t *count,
                            uint32_t *shape, uint32_t *ndim)
{
  return guarded(nullptr, [&] {
    if (!path || !key) throw InvalidArgument{"NULL argument"};
    tenf::File tf(path);
    const tenf::Json *node = &tf.metadata();
    std::string k(key);
    size_t pos = 0;
    while (pos <= k.size()) {
      size_t next = k.find('/', pos);
      if (next == std::string::npos) next = k.size();
      if (next > pos) node = &node->at(k.substr(pos, next - pos));
      pos = next + 1;
    }
    tenf::FloatTensor t = tf.unpack_f32(*node);
    if (count) *count = t.size();
    if (ndim) *ndim = (uint32_t)t.shape.size();
    if (shape)
      for (size_t i = 0; i < t.shape.size() && i < 8; i++) shape[i] = (uint32_t)t.shape[i];
    if (out && capacity) std::memcpy(out, t.data.data(), std::min(capacity, t.size()) * sizeof(float));
  });
}

void bear_cuda_destroy(bear_cuda_renderer *r) { delete r; }

const char *bear_cuda_last_error(const bear_cuda_renderer *r) { return r ? r->error.c_str() : g_create_error.c_str(); }

int bear_cuda_process_batch(bear_cuda_renderer *r, uint32_t T, const float *objects, size_t objects_stride,
                            const float *direct_speakers, size_t direct_speakers_stride, const float *hoa,
                            size_t hoa_stride, const bear_cuda_block_tables *tables, float *output, size_t out_stride)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    r->shadow_valid = false;   // the single-period graph path has to see all tables again
    Layout &p = r->lay;
    if (T == 0 || (int)T > p.Tmax) throw InvalidArgument{"T must be in 1..max_batch_blocks"};
    if (!output && !output_optional(r)) throw InvalidArgument{"output must not be NULL"};
    if (p.N && !objects) throw InvalidArgument{"objects input must not be NULL"};
    if (p.M && !direct_speakers) throw InvalidArgument{"direct_speakers input must not be NULL"};
    if (p.H && !hoa) throw InvalidArgument{"hoa input must not be NULL"};
    // Uploads go to their own stream in chunks of periods, the render stream waits for
    // chunk c before it queues the kernels of chunk c: copies of later chunks overlap the kernels.
    cudaStream_t cs = r->copy_stream;
    // the staging areas may still be in use by an earlier asynchronous call (bear_cuda_process_device)
    CK(cudaEventRecord(r->ev_order, r->stream));
    CK(cudaStreamWaitEvent(cs, r->ev_order, 0));
    upload_tables(r, (int)T, tables, false, cs);
    // (Shorter chunks towards the end do not pay: a chunk of a few periods still takes ~0.1 ms to
    // go through the six kernels, longer than its upload.)
    const int chunk = host_chunk_periods(r, (int)T, (size_t)(p.N + p.M + p.H) * p.B * sizeof(float) + table_bytes_per_period(p));
    for (int k0 = 0, c = 0; k0 < (int)T; k0 += chunk, c++) {
      const size_t col = (size_t)k0 * p.B, width = (size_t)std::min(chunk, (int)T - k0) * p.B * sizeof(float);
      auto up_audio = [&](int row0, int rows, const float *host, size_t stride) {
        if (rows)
          CK(cudaMemcpy2DAsync(p.xin + (size_t)row0 * p.xs + p.hist + col, p.xs * sizeof(float), host + col,
                               stride * sizeof(float), width, rows, cudaMemcpyHostToDevice, cs));
      };
      up_audio(0, p.N, objects, objects_stride);
      up_audio(p.N, p.M, direct_speakers, direct_speakers_stride);
      up_audio(p.N + p.M, p.H, hoa, hoa_stride);
      CK(cudaEventRecord(r->h2d_events[c], cs));
    }
    // finished chunks are read back on a third stream while later chunks still upload and render
    run_pipeline(
        r, (int)T, chunk, [&](int c, int, int) { CK(cudaStreamWaitEvent(r->stream, r->h2d_events[c], 0)); },
        [&](int c, int k0, int n) {
          if (!output) return;
          CK(cudaEventRecord(r->done_events[c], r->result_stream));
          CK(cudaStreamWaitEvent(r->d2h_stream, r->done_events[c], 0));
          CK(cudaMemcpy2DAsync(output + (size_t)k0 * p.B, out_stride * sizeof(float), p.out + (size_t)k0 * p.B,
                               p.os * sizeof(float), (size_t)n * p.B * sizeof(float), 2, cudaMemcpyDeviceToHost,
                               r->d2h_stream));
        });
    CK(cudaStreamSynchronize(r->d2h_stream));
    fetch_kernel_errors(r);
  });
}

int bear_cuda_process_batch_frames(bear_cuda_renderer *r, uint32_t T, const void *frames, uint32_t format,
                                   uint32_t frame_channels, const int32_t *tracks,
                                   const bear_cuda_block_tables *tables, float *output, size_t out_stride)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    r->shadow_valid = false;
    Layout &p = r->lay;
    const int rows = p.N + p.M + p.H;
    if (T == 0 || (int)T > p.Tmax) throw InvalidArgument{"T must be in 1..max_batch_blocks"};
    if (!output && !output_optional(r)) throw InvalidArgument{"output must not be NULL"};
    if (format > BEAR_CUDA_FRAMES_S32) throw InvalidArgument{"unknown frame format"};
    if (rows && (!frames || !tracks || frame_channels == 0)) throw InvalidArgument{"frames and tracks must not be NULL"};
    for (int i = 0; i < rows; i++)
      if (tracks[i] < -1 || tracks[i] >= (int)frame_channels) throw InvalidArgument{"track index out of range"};
    static const size_t bytes_per_sample[4] = {4, 2, 3, 4};
    const size_t frame_bytes = (size_t)frame_channels * bytes_per_sample[format];
    const size_t total = (size_t)T * p.B * frame_bytes;
    if (rows) {
      if (total > r->d_frames_bytes) {   // first call (or a wider file than before): grow the staging buffer
        CK(cudaStreamSynchronize(r->stream));
        if (r->d_frames) { cudaFree(r->d_frames); r->device_bytes -= r->d_frames_bytes; }
        r->d_frames = nullptr;
        r->d_frames_bytes = 0;
        CK(cudaMalloc(&r->d_frames, total));
        r->d_frames_bytes = total;
        r->device_bytes += total;
      }
      if (!r->d_tracks) CK(cudaMalloc(&r->d_tracks, (size_t)rows * sizeof(int)));
    }
    cudaStream_t cs = r->copy_stream;
    CK(cudaEventRecord(r->ev_order, r->stream));
    CK(cudaStreamWaitEvent(cs, r->ev_order, 0));
    if (rows) CK(cudaMemcpyAsync(r->d_tracks, tracks, (size_t)rows * sizeof(int), cudaMemcpyHostToDevice, cs));
    upload_tables(r, (int)T, tables, false, cs);
    const int chunk = host_chunk_periods(r, (int)T, (size_t)p.B * frame_bytes + table_bytes_per_period(p));
    for (int k0 = 0, c = 0; k0 < (int)T; k0 += chunk, c++) {
      const size_t off = (size_t)k0 * p.B * frame_bytes, bytes = (size_t)std::min(chunk, (int)T - k0) * p.B * frame_bytes;
      if (rows)
        CK(cudaMemcpyAsync(r->d_frames + off, static_cast<const unsigned char *>(frames) + off, bytes,
                           cudaMemcpyHostToDevice, cs));
      CK(cudaEventRecord(r->h2d_events[c], cs));
    }
    run_pipeline(
        r, (int)T, chunk,
        [&](int c, int k0, int n) {
          CK(cudaStreamWaitEvent(r->stream, r->h2d_events[c], 0));
          launch_frames(p, r->d_frames, (int)format, (int)frame_channels, r->d_tracks, rows, k0 * p.B, n * p.B, r->stream);
          if (rows) r->launches++;
        },
        [&](int c, int k0, int n) {
          if (!output) return;
          CK(cudaEventRecord(r->done_events[c], r->result_stream));
          CK(cudaStreamWaitEvent(r->d2h_stream, r->done_events[c], 0));
          CK(cudaMemcpy2DAsync(output + (size_t)k0 * p.B, out_stride * sizeof(float), p.out + (size_t)k0 * p.B,
                               p.os * sizeof(float), (size_t)n * p.B * sizeof(float), 2, cudaMemcpyDeviceToHost,
                               r->d2h_stream));
        });
    CK(cudaStreamSynchronize(r->d2h_stream));
    fetch_kernel_errors(r);
  });
}

// One period through the captured graph.  Returns false when the call has to take the plain path:
// tables prepared on the device, profiling on, a NULL port while the host copy of the tables is stale,
// or graphs unavailable.  Argument errors throw before anything is queued.
static bool stream_call_as_graph(bear_cuda_renderer *r, const float *const *objects_input,
                                 const float *const *direct_speakers_input, const float *const *hoa_input,
                                 const bear_cuda_block_tables *tables, float *const *output)
{
  Layout &p = r->lay;
  const size_t B = p.B;
  if (!r->graph_ok || !tables || r->profiling || !r->h_graph || r->lay.W > 1) return false;
  const float *host[8] = {tables->direct_gains, tables->diffuse_gains, tables->direct_delays,
                          tables->direct_speakers_gains, tables->direct_speakers_delays, tables->static_delays,
                          tables->hoa_gains, nullptr};
  float *dev[8] = {p.g_dir, p.g_dif, p.d_dir, p.g_ds, p.d_ds, p.d_stat, p.g_hoa, reinterpret_cast<float *>(p.view)};
  const bool have_view = tables->brir_index && tables->brir_index[0] >= 0;
  if (have_view && tables->brir_index[0] >= p.V) throw InvalidArgument{"brir_index out of range"};
  host[7] = have_view ? reinterpret_cast<const float *>(tables->brir_index) : nullptr;
  if (!r->shadow_valid)
    for (auto &sl : r->stream_slots)
      for (int i = 0; i < 8; i++)
        if (dev[i] == sl.dev && !host[i]) return false;
  // audio
  int row = 0;
  auto gather = [&](const float *const *ptrs, int count, const char *name) {
    if (count && !ptrs) throw InvalidArgument{std::string(name) + " input must not be NULL"};
    for (int i = 0; i < count; i++)
      if (!ptrs[i]) throw InvalidArgument{std::string(name) + " channel pointer is NULL"};
    for (int i = 0; i < count; i++, row++) std::memcpy(r->h_graph + (size_t)row * B, ptrs[i], B * sizeof(float));
  };
  gather(objects_input, p.N, "objects");
  gather(direct_speakers_input, p.M, "direct_speakers");
  gather(hoa_input, p.H, "hoa");
  for (auto &sl : r->stream_slots)
    for (int i = 0; i < 8; i++)
      if (dev[i] == sl.dev && host[i]) std::memcpy(r->h_graph + sl.off, host[i], sl.count * sizeof(float));
  r->shadow_valid = true;

  if (!r->graph_exec) {
    // capture once: upload, scatter, the kernels of one period, read-back
    UnpackPlan plan{};
    plan.audio = r->d_pack;
    plan.rows = row;
    for (auto &sl : r->stream_slots) {
      float *dst = sl.dev + sl.slot * sl.count;
      plan.item[plan.n++] = UnpackItem{r->d_pack + sl.off, dst, (int)sl.count};
    }
    const uint64_t launches0 = r->launches;
    cudaGraph_t g = nullptr;
    bool ok = cudaStreamBeginCapture(r->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
    if (ok) {
      try {
        CK(cudaMemcpyAsync(r->d_pack, r->h_graph, r->stream_floats * sizeof(float), cudaMemcpyHostToDevice, r->stream));
        launch_unpack(p, plan, r->stream);
        r->launches++;
        run_pipeline(r, 1);
        CK(cudaMemcpy2DAsync(r->h_out, B * sizeof(float), p.out, p.os * sizeof(float), B * sizeof(float), 2,
                             cudaMemcpyDeviceToHost, r->stream));
        CK(cudaMemcpyAsync(r->h_err, p.err_flag, sizeof(int), cudaMemcpyDeviceToHost, r->stream));
      } catch (...) {
        ok = false;
      }
      if (cudaStreamEndCapture(r->stream, &g) != cudaSuccess) ok = false;
    }
    if (ok && cudaGraphInstantiate(&r->graph_exec, g, 0) != cudaSuccess) ok = false;
    r->graph_launch_count = r->launches - launches0;
    r->launches = launches0;
    if (!ok) {   // no graphs here: stay on the plain path for good
      if (g) cudaGraphDestroy(g);
      r->graph_exec = nullptr;
      r->graph_ok = false;
      r->shadow_valid = false;
      cudaGetLastError();
      return false;
    }
    r->graph = g;
  }
  CK(cudaGraphLaunch(r->graph_exec, r->stream));
  r->launches += r->graph_launch_count;
  r->calls++;
  CK(cudaStreamSynchronize(r->stream));
  if (r->h_err[0]) fetch_kernel_errors(r);   // the graph ends with a copy of the kernels' error bits
  std::memcpy(output[0], r->h_out, B * sizeof(float));
  std::memcpy(output[1], r->h_out + B, B * sizeof(float));
  return true;
}

int bear_cuda_process(bear_cuda_renderer *r, const float *const *objects_input,
                      const float *const *direct_speakers_input, const float *const *hoa_input,
                      const bear_cuda_block_tables *tables, float *const *output)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    Layout &p = r->lay;
    if (!output || !output[0] || !output[1]) throw InvalidArgument{"output must not be NULL"};
    const size_t B = p.B;
    if (stream_call_as_graph(r, objects_input, direct_speakers_input, hoa_input, tables, output)) return;
    r->shadow_valid = false;   // this call goes the plain way: the graph path's host copy of the tables goes stale
    int row = 0;
    auto gather = [&](const float *const *ptrs, int count, const char *name) {
      if (count && !ptrs) throw InvalidArgument{std::string(name) + " input must not be NULL"};
      for (int i = 0; i < count; i++, row++) {
        if (!ptrs[i]) throw InvalidArgument{std::string(name) + " channel pointer is NULL"};
        std::memcpy(r->h_in + (size_t)row * B, ptrs[i], B * sizeof(float));
      }
    };
    gather(objects_input, p.N, "objects");
    gather(direct_speakers_input, p.M, "direct_speakers");
    gather(hoa_input, p.H, "hoa");
    // tables: packed behind the audio; a port that did not change repeats the previous period's slot
    UnpackPlan plan{};
    plan.audio = r->d_pack;
    plan.rows = row;
    size_t off = (size_t)row * B;
    auto item = [&](const float *host, float *dev, size_t per_block, size_t slot, const char *name) {
      if (per_block == 0) return;
      float *dst = dev + slot * per_block;
      UnpackItem u{dst - per_block, dst, (int)per_block};      // default: unchanged
      if (host) {
        std::memcpy(r->h_in + off, host, per_block * sizeof(float));
        u.src = r->d_pack + off;
        off += per_block;
      }
      (void)name;
      plan.item[plan.n++] = u;
    };
    if (tables) {
      item(tables->direct_gains, p.g_dir, (size_t)p.L * p.N, 1, "direct_gains");
      item(tables->diffuse_gains, p.g_dif, (size_t)p.L * p.N, 1, "diffuse_gains");
      item(tables->direct_delays, p.d_dir, (size_t)2 * p.N, 1, "direct_delays");
      item(tables->direct_speakers_gains, p.g_ds, (size_t)p.L * p.M, 1, "direct_speakers_gains");
      item(tables->direct_speakers_delays, p.d_ds, (size_t)2 * p.M, 1, "direct_speakers_delays");
      item(tables->static_delays, p.d_stat, (size_t)2 * p.L, 2, nullptr);
      item(tables->hoa_gains, p.g_hoa, (size_t)p.A * p.H, 1, nullptr);
      const bool have_view = tables->brir_index && tables->brir_index[0] >= 0;
      if (have_view && tables->brir_index[0] >= p.V) throw InvalidArgument{"brir_index out of range"};
      item(have_view ? reinterpret_cast<const float *>(tables->brir_index) : nullptr, reinterpret_cast<float *>(p.view), 1, 1,
           nullptr);
    } else {
      if (!r->tables_on_device) throw InvalidArgument{"tables must not be NULL (no bear_cuda_control_tables call pending)"};
      r->tables_on_device = false;
    }
    if (off) CK(cudaMemcpyAsync(r->d_pack, r->h_in, off * sizeof(float), cudaMemcpyHostToDevice, r->stream));
    launch_unpack(p, plan, r->stream);
    r->launches++;
    run_pipeline(r, 1);
    CK(cudaMemcpy2DAsync(r->h_out, B * sizeof(float), p.out, p.os * sizeof(float), B * sizeof(float), 2,
                         cudaMemcpyDeviceToHost, r->stream));
    fetch_kernel_errors(r);
    std::memcpy(output[0], r->h_out, B * sizeof(float));
    std::memcpy(output[1], r->h_out + B, B * sizeof(float));
  });
}

int bear_cuda_control_tables(bear_cuda_renderer *r, uint32_t T, const float *object_gains,
                             const float *direct_speakers_gains, const int32_t *brir_index, const float *hoa_gains)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    r->shadow_valid = false;   // the single-period graph path has to see all tables again
    Layout &p = r->lay;
    cudaStream_t st = r->stream;
    if (T == 0 || (int)T > p.Tmax) throw InvalidArgument{"T must be in 1..max_batch_blocks"};
    if (p.N && !object_gains) throw InvalidArgument{"object_gains must not be NULL"};
    if (p.M && !direct_speakers_gains) throw InvalidArgument{"direct_speakers_gains must not be NULL"};
    if (!brir_index) throw InvalidArgument{"brir_index must not be NULL"};
    if (p.H && !hoa_gains) throw InvalidArgument{"hoa_gains must not be NULL"};
    for (uint32_t k = 0; k < T; k++)
      if (brir_index[k] < 0 || brir_index[k] >= p.V) throw InvalidArgument{"brir_index out of range"};
    CK(cudaMemcpyAsync(p.view + 1, brir_index, (size_t)T * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (p.N)
      CK(cudaMemcpyAsync(p.raw_obj, object_gains, (size_t)T * 2 * p.L * p.N * sizeof(float), cudaMemcpyHostToDevice, st));
    if (p.M)
      CK(cudaMemcpyAsync(p.raw_ds, direct_speakers_gains, (size_t)T * p.L * p.M * sizeof(float), cudaMemcpyHostToDevice, st));
    if (p.H)
      CK(cudaMemcpyAsync(p.g_hoa + (size_t)p.A * p.H, hoa_gains, (size_t)T * p.A * p.H * sizeof(float),
                         cudaMemcpyHostToDevice, st));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (r->profiling) {
      auto get = [&]() { cudaEvent_t e; if (r->ev_free.empty()) CK(cudaEventCreate(&e)); else { e = r->ev_free.back(); r->ev_free.pop_back(); } return e; };
      e0 = get(); e1 = get();
      CK(cudaEventRecord(e0, st));
    }
    launch_control(p, (int)T, st);
    CK(cudaGetLastError());
    if (r->profiling) {
      CK(cudaEventRecord(e1, st));
      r->ev_pending.push_back({e0, e1});
      r->ev_kernel.push_back({BEAR_CUDA_KERNEL_CONTROL});
    }
    r->launches++;
    r->tables_on_device = true;
  });
}

int bear_cuda_get_device_io(bear_cuda_renderer *r, bear_cuda_device_io *io)
{
  if (!r || !io) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  Layout &p = r->lay;
  std::memset(io, 0, sizeof(*io));
  io->audio_in = p.xin; io->audio_stride = p.xs; io->audio_offset = p.hist;
  io->direct_gains = p.g_dir + (size_t)p.L * p.N;
  io->diffuse_gains = p.g_dif + (size_t)p.L * p.N;
  io->direct_delays = p.d_dir + (size_t)2 * p.N;
  io->direct_speakers_gains = p.g_ds + (size_t)p.L * p.M;
  io->direct_speakers_delays = p.d_ds + (size_t)2 * p.M;
  io->static_delays = p.d_stat + (size_t)2 * 2 * p.L;
  io->brir_index = p.view + 1;
  io->hoa_gains = p.g_hoa + (size_t)p.A * p.H;
  io->audio_out = p.out; io->out_stride = p.os;
  io->stream = r->stream;
  return BEAR_CUDA_OK;
}

int bear_cuda_process_device(bear_cuda_renderer *r, uint32_t T)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    r->shadow_valid = false;   // the single-period graph path has to see all tables again
    if (T == 0 || (int)T > r->lay.Tmax) throw InvalidArgument{"T must be in 1..max_batch_blocks"};
    run_pipeline(r, (int)T, true);
  });
}

int bear_cuda_process_device_pipelined(bear_cuda_renderer *r, uint32_t T)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    r->shadow_valid = false;
    if (T == 0 || (int)T > r->lay.Tmax) throw InvalidArgument{"T must be in 1..max_batch_blocks"};
    if (r->conv_stream && !r->pipelined_pending) {
      // reads of the previous output that the caller queued on the result stream come first
      CK(cudaEventRecord(r->ev_call, r->conv_stream));
      CK(cudaStreamWaitEvent(r->stream, r->ev_call, 0));
    }
    run_pipeline(r, (int)T, true, true);
    if (!r->pipelined_pending && r->conv_stream) {
      // the call ran unpipelined (see the header for when): its output is still promised on the result stream
      CK(cudaEventRecord(r->ev_call, r->stream));
      CK(cudaStreamWaitEvent(r->conv_stream, r->ev_call, 0));
    }
  }, false);
}

int bear_cuda_result_stream(bear_cuda_renderer *r, void **stream)
{
  if (!r || !stream) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  *stream = r->conv_stream ? r->conv_stream : r->stream;
  return BEAR_CUDA_OK;
}

int bear_cuda_process_device_static_scratch(bear_cuda_renderer *r, uint32_t T)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  Layout &p = r->lay;
  // Argument errors are reported without the exception path of the other entry points (an exception
  // object owns a heap string): the text goes into the buffer reserved when the renderer was created.
  char msg[384];
  if (T == 0 || (int)T > p.Tmax) {
    r->error.assign("T must be in 1..max_batch_blocks");
    return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  }
  if (p.W > 1) {
    r->error.assign("a renderer that shares a scene with other GPUs cannot use the static scratch variant");
    return BEAR_CUDA_ERR_UNSUPPORTED;
  }
  const size_t need = (size_t)T * 2 * 2 * r->ks_cfg * p.B * sizeof(float2);
  if (need > kStaticScratchBytes) {
    std::snprintf(msg, sizeof msg,
                  "static scratch too small: %u periods need %zu bytes of partial spectra (2 ears x 2 views x %d splits x "
                  "%d bins), the built-in array holds %zu; use fewer periods per call or bear_cuda_process_device",
                  T, need, r->ks_cfg, p.B, (size_t)kStaticScratchBytes);
    r->error.assign(msg);
    return BEAR_CUDA_ERR_UNSUPPORTED;
  }
  return guarded(r, [&] {
    r->shadow_valid = false;   // the single-period graph path has to see all tables again
    // Same kernels, same state; only the partial spectra go to the built-in array.  Nothing in it has
    // to start at zero: cmac_kernel writes every partial spectrum that ifft_out_kernel reads in the
    // same call, so whatever an earlier call (of any renderer) left behind is never read.
    float2 *own = p.ypart;
    const bool prof = r->profiling;
    p.ypart = r->static_scratch;   // looked up at creation (cudaGetSymbolAddress with the device current)
    r->profiling = false;  // the event pool may grow; this entry point must not create anything
    r->ks_dynamic = false; // the size check above is for the configured number of splits
    try {
      run_pipeline(r, (int)T, true);
    } catch (...) {
      p.ypart = own;
      r->profiling = prof;
      r->ks_dynamic = true;
      throw;
    }
    p.ypart = own;
    r->profiling = prof;
    r->ks_dynamic = true;
  });
}

int bear_cuda_reset(bear_cuda_renderer *r)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    CK(cudaStreamSynchronize(r->stream));
    for (auto &a : r->state_allocs) CK(cudaMemsetAsync(a.first, 0, a.second, r->stream));
    r->parity = 0;
    r->calls = 0;
    set_initial_state(r);
    CK(cudaStreamSynchronize(r->stream));
  });
}

int bear_cuda_synchronize(bear_cuda_renderer *r)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] { fetch_kernel_errors(r); });
}

int bear_cuda_get_info(const bear_cuda_renderer *r, bear_cuda_info *info)
{
  if (!r || !info) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  std::memset(info, 0, sizeof(*info));
  info->num_partitions = r->lay.P;
  info->object_splits = r->lay.S;
  info->cmac_splits = r->lay.KS;
  info->history_samples = r->lay.hist;
  info->kernels_per_call = r->kernels_per_call;
  info->device_bytes = r->device_bytes;
  info->shared_bytes = r->bank ? r->bank->bytes : 0;
  int grp[2] = {0, 0};
  if (r->lay.grp_count && cudaMemcpyAsync(grp, r->lay.grp_count, sizeof grp, cudaMemcpyDeviceToHost, r->stream) == cudaSuccess)
    cudaStreamSynchronize(r->stream);
  info->cmac_items_last = (uint32_t)grp[0];
  info->cmac_splits_last = (uint32_t)grp[1];
  info->launches_total = r->launches;
  info->default_direct_delay = r->default_direct_delay;
  info->default_static_delay = r->default_static_delay;
  info->world = (uint32_t)r->lay.W; info->rank = (uint32_t)r->lay.rank;
  info->speakers_begin = (uint32_t)r->lay.l0; info->speakers_end = (uint32_t)r->lay.l1;
  return BEAR_CUDA_OK;
}

int bear_cuda_set_profiling(bear_cuda_renderer *r, int enable)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  r->profiling = enable != 0;
  return BEAR_CUDA_OK;
}

int bear_cuda_get_profile(bear_cuda_renderer *r, bear_cuda_kernel_profile *out, int reset)
{
  if (!r || !out) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    CK(cudaStreamSynchronize(r->stream));
    for (size_t c = 0; c < r->ev_pending.size(); c++) {
      auto &chain = r->ev_pending[c];
      for (size_t i = 0; i + 1 < chain.size(); i++) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, chain[i], chain[i + 1]));
        r->prof_ms[r->ev_kernel[c][i]] += ms;
        r->prof_n[r->ev_kernel[c][i]]++;
      }
      for (cudaEvent_t e : chain) r->ev_free.push_back(e);
    }
    r->ev_pending.clear();
    r->ev_kernel.clear();
    static const char *names[BEAR_CUDA_NUM_KERNELS] = {"mix_kernel", "hoa_mix_kernel", "decorrelate_kernel",
                                                       "bus_fdl_kernel", "cmac_kernel", "ifft_out_kernel",
                                                       "carry_kernel", "control_kernel", "mix_aux_kernels",
                                                       "exchange", "bus_gather_kernel"};
    for (int k = 0; k < BEAR_CUDA_NUM_KERNELS; k++) {
      std::snprintf(out[k].name, sizeof(out[k].name), "%s", names[k]);
      out[k].total_ms = r->prof_ms[k];
      out[k].launches = r->prof_n[k];
      if (reset) { r->prof_ms[k] = 0; r->prof_n[k] = 0; }
    }
  });
}

void bear_cuda_comm_speaker_range(uint32_t rank, uint32_t world, uint32_t L, uint32_t *begin, uint32_t *end)
{
  if (world == 0) world = 1;
  if (begin) *begin = (uint32_t)((uint64_t)rank * L / world);
  if (end) *end = (uint32_t)((uint64_t)(rank + 1) * L / world);
}

int bear_cuda_comm_blob(bear_cuda_renderer *r, uint32_t rank, void *blob)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    if (!blob) throw InvalidArgument{"blob must not be NULL"};
    const Layout &p = r->lay;
    std::memset(blob, 0, BEAR_CUDA_COMM_BLOB_BYTES);
    CommBlob b{};
    b.magic = kCommMagic; b.B = (uint32_t)p.B; b.Tmax = (uint32_t)p.Tmax; b.L = (uint32_t)p.L;
    b.cs = p.cs;
    b.N = (uint32_t)p.N; b.M = (uint32_t)p.M; b.H = (uint32_t)p.H; b.rank = rank;
    if (!r->shared_bus) {
      // peers map ONE allocation: the kBusSets sets of bus rows the calls cycle through
      if (r->calls != 0) throw InvalidArgument{"connect before the first process call (or after bear_cuda_reset)"};
      CK(cudaStreamSynchronize(r->stream));
      r->dfree(r->bus[0]);
      const size_t set = (size_t)3 * r->lay.L * r->lay.cs;
      r->inbox_off = round_up(kBusSets * set, 64);
      r->flags_off = r->inbox_off + round_up((size_t)kMaxRanks * 2 * r->lay.os, 64);
      // (not render state: bear_cuda_reset leaves the counters alone, they only ever grow)
      r->bus[0] = r->dalloc<float>(r->flags_off + 64, false);
      CK(cudaMemsetAsync(r->bus[0], 0, (r->flags_off + 64) * sizeof(float), r->stream));
      CK(cudaStreamSynchronize(r->stream));
      for (int k = 1; k < kBusSets; k++) r->bus[k] = r->bus[0] + k * set;
      r->lay.cdir = r->bus[0];
      r->lay.qdif = r->bus[0] + (size_t)2 * r->lay.L * r->lay.cs;
      r->shared_bus = true;
    }
    CK(cudaIpcGetMemHandle(&b.bus, r->bus[0]));
    if (rank == 0) NCK(nccl().GetUniqueId(&b.id));
    std::memcpy(blob, &b, sizeof b);
  });
}

int bear_cuda_comm_connect(bear_cuda_renderer *r, uint32_t rank, uint32_t world, const void *blobs, int32_t root)
{
  if (!r) return BEAR_CUDA_ERR_INVALID_ARGUMENT;
  return guarded(r, [&] {
    Layout &p = r->lay;
    if (!blobs) throw InvalidArgument{"blobs must not be NULL"};
    if (world < 1 || world > (uint32_t)kMaxRanks) throw InvalidArgument{"world must be in 1..8"};
    if (rank >= world) throw InvalidArgument{"rank must be below world"};
    if (root >= (int32_t)world) throw InvalidArgument{"root must be below world (or negative: every rank)"};
    if (r->calls != 0) throw InvalidArgument{"connect before the first process call (or after bear_cuda_reset)"};
    if (r->comm) throw InvalidArgument{"already connected"};
    if (world == 1) return;
    const unsigned char *bytes = static_cast<const unsigned char *>(blobs);
    CommBlob bl[kMaxRanks];
    bool any_objects = false;
    for (uint32_t w = 0; w < world; w++) {
      std::memcpy(&bl[w], bytes + (size_t)w * BEAR_CUDA_COMM_BLOB_BYTES, sizeof(CommBlob));
      if (bl[w].magic != kCommMagic || bl[w].rank != w) throw InvalidArgument{"blobs are not ordered by rank"};
      if (bl[w].B != (uint32_t)p.B || bl[w].Tmax != (uint32_t)p.Tmax || bl[w].L != (uint32_t)p.L || bl[w].cs != p.cs)
        throw InvalidArgument{"all ranks must use the same period size, max_batch_blocks and data file"};
      any_objects = any_objects || bl[w].N > 0;
      const size_t tbl = sizeof(float) * ((size_t)2 * p.L * bl[w].N + 2 * bl[w].N + (size_t)p.L * bl[w].M + 2 * bl[w].M +
                                          2 * p.L + (size_t)p.A * bl[w].H + 1);
      r->comm_bytes_per_period = std::max(r->comm_bytes_per_period,
                                          (size_t)(bl[w].N + bl[w].M + bl[w].H) * p.B * sizeof(float) + tbl);
    }
    CK(cudaStreamSynchronize(r->stream));
    // the peers' bus rows, mapped into this process (read with ordinary loads by decorrelate_kernel and
    // bus_fdl_kernel; nothing here is ever the source of a bulk or multicast copy)
    for (uint32_t w = 0; w < world; w++) {
      if (w == rank) {
        r->peer_bus[w][0] = r->bus[0];
      } else {
        void *base = nullptr;
        CK(cudaIpcOpenMemHandle(&base, bl[w].bus, cudaIpcMemLazyEnablePeerAccess));
        r->peer_bus[w][0] = static_cast<float *>(base);
      }
      for (int k = 1; k < kBusSets; k++) r->peer_bus[w][k] = r->peer_bus[w][0] + (size_t)k * 3 * p.L * p.cs;
      r->peer_base[w] = r->peer_bus[w][0];
    }
    NCK(nccl().CommInitRank(&r->comm, (int)world, bl[0].id, (int)rank));
    if (const char *env = std::getenv("BEAR_CUDA_SHARED_CHUNKS")) r->shared_chunks = std::max(1, std::atoi(env));
    if (const char *env = std::getenv("BEAR_CUDA_TIMELINE")) r->timeline_path = env;
    {
      int least = 0, greatest = 0;
      CK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
      // the exchange + convolution side is the latency chain of a pipelined call: its (small) kernels go
      // first whenever a mix CTA retires.  BEAR_CUDA_CONV_PRIORITY=0 switches that off (experiments).
      const char *env = std::getenv("BEAR_CUDA_CONV_PRIORITY");
      const int prio = env && !std::atoi(env) ? 0 : greatest;
      CK(cudaStreamCreateWithPriority(&r->conv_stream, cudaStreamNonBlocking, prio));
    }
    CK(cudaEventCreateWithFlags(&r->ev_call, cudaEventDisableTiming));
    if (!r->shared_bus) throw InvalidArgument{"bear_cuda_comm_blob comes first"};
    for (cudaEvent_t &e : r->ev_barrier) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (int b = 0; b < kBusSets; b++) {
      r->sh_view[b] = r->dalloc<int32_t>((size_t)p.Tmax + 1);
      r->sh_dstat[b] = r->dalloc<float>((size_t)(p.Tmax + 2) * 2 * p.L);
    }
    r->mix_events.resize(std::max<size_t>(r->done_events.size(), 4));
    for (cudaEvent_t &e : r->mix_events) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    r->bus_sum = r->dalloc<float>((size_t)3 * p.L * p.cs);
    r->comm_root = root;
    uint32_t l0 = 0, l1 = 0;
    bear_cuda_comm_speaker_range(rank, world, (uint32_t)p.L, &l0, &l1);
    p.W = (int)world; p.rank = (int)rank; p.l0 = (int)l0; p.l1 = (int)l1;
    p.dif = any_objects;
    r->ks_units = std::max(1, (p.l1 - p.l0) + (p.H > 0 ? 1 : 0));
    r->ks_cfg = std::max(1, std::min(r->ks_cfg, r->ks_units));
    p.KS = r->ks_cfg;
    CK(cudaStreamSynchronize(r->stream));
  });
}

void *bear_cuda_host_alloc(size_t bytes)
{
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
  return p;
}
void bear_cuda_host_free(void *p)
{
  if (p) cudaFreeHost(p);
}

}  // extern "C"
