// Device-side layout shared by kernels.cu and renderer.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bearcuda {

constexpr int kMaxSpeakers = 24;  // accumulator tile of the mix kernel (L <= 24)
constexpr int kMixTile = 128;     // samples per CTA in the mix kernels
#ifndef BEAR_MIX_CHUNK
#define BEAR_MIX_CHUNK 12
#endif
constexpr int kMixChunk = BEAR_MIX_CHUNK;   // channels staged in shared memory per pass
constexpr int kMixSlots = 6;      // non-zero loudspeakers a record lists (two panning triplets); channels with
                                  // more take the dense path
constexpr int kMixRecWords = 8 + 4 * kMixSlots;   // words of one (period, channel) record of mix_prep_kernel
constexpr int kMaxObjectSplits = 32;              // object splits of a launch (mix_counts row length)
constexpr int kMaxSplitChannels = 4096;           // channels one object split can sort in shared memory
constexpr int kMaxRanks = 8;      // GPUs that can share one scene (bear_cuda_comm_connect)
// bits of Layout::err_flag (checked by the host after a call has finished)
constexpr int kErrDelayRange = 1; // a table delay is negative or beyond history_samples - 4
constexpr int kErrViewRange = 2;  // a brir_index on the device-resident path is outside [0, V)
constexpr int kErrPeerTimeout = 4;  // a peer GPU did not reach a synchronisation point of a shared scene in ~20 s
constexpr long long kPeerTimeoutCycles = 40LL << 30;   // SM clock cycles (1.9 GHz: 22 s): host work between calls may differ by seconds from rank to rank
#ifndef BEAR_CMAC_GROUP
#define BEAR_CMAC_GROUP 5
#endif
constexpr int kCmacGroup = BEAR_CMAC_GROUP;     // periods of one view that share a filter load in the cMAC kernel

// Everything a kernel needs to find its data.  Passed by value (fits the 4 KB param space).
struct Layout {
  // sizes
  int N, M, H;        // objects, DirectSpeakers, HOA input channels
  int B, logB;        // period (= complex FFT length), log2
  int P;              // BRIR partitions
  int L, A;           // virtual loudspeakers, HOA decoder channels
  int CH;             // FDL channels = 2L + A
  int V;              // views
  int Tmax;           // max blocks per call
  int hist;           // time-domain history (samples) kept in front of every delayed signal
  int S;              // object splits of this launch: mix_kernel CTAs that share a tile
  int KS;             // cMAC splits (partial spectra)
  int R;              // FDL slots per channel (P + Tmax)
  int goff, doff, xoff;  // ramp offsets
  float fs;
  float hoa_delay;    // samples

  // audio in: [N+M+H][xs], block data from column `hist`
  float *xin;  size_t xs;
  // tables, slot 0 = state carried from the previous call
  float *g_dir, *g_dif;   // [Tmax+1][L][N]
  float *g_ds;            // [Tmax+1][L][M]
  float *g_hoa;           // [Tmax+1][A][H]
  float *d_dir;           // [Tmax+1][2N] seconds
  float *d_ds;            // [Tmax+1][2M]
  float *d_stat;          // [Tmax+2][2L]  slots 0,1 = blocks -2,-1
  int32_t *view;          // [Tmax+1]
  // time-domain intermediates (row stride, prefix)
  float *cdir;  size_t cs;   // [2L][cs]      prefix B   direct bus
  float *qdif;               // [L][cs]       prefix B   diffuse bus
  // partial sums of object splits 1..S-1 of one launch, added to cdir/qdif by bus_sum_kernel:
  float *bus_part; size_t bps;   // [S-1][3L][bps], columns relative to the launch; rows 0..2L-1 direct, 2L..3L-1 diffuse
  float *w;     size_t ws;   // [L][ws]       prefix B+hist  (decorrelator outputs)
  float *amb;                // [A][ws]       prefix B+hist  (HOA decoder inputs)
  // spectra (packed: bin 0 = (DC, Nyquist), bins 1..B-1 complex)
  float2 *fdl;               // [CH][R][B], R = P + Tmax: slots 0..P-1 = the P periods before this call, slot P + k = period k
  const float2 *h_brir;      // [V][2L][P][B], pre-scaled by 1/(2B)
  const float2 *h_dec;       // [L][B]
  const float2 *h_hoa;       // [A][2][B]
  float2 *ypart;             // [periods of one launch][2][2][KS][B], KS chosen per launch
  const float2 *tw;          // [2B]: [0,B) exp(-i pi q / B); [B,2B) per-pass radix-8 twiddles
  // out
  float *out;  size_t os;    // [2][os]
  // control tail (optional): data-file delay table, gain-normalisation factors, raw gains of this call
  const float *delays_tab;   // [V][L][2] samples
  const float *gn_factors;   // [gnS][V][2L][2L][2] or null
  int gnS;
  float *raw_obj;            // [Tmax][2L][N]
  float *raw_ds;             // [Tmax][L][M]
  double fs_d, dec_delay_d;
  int front;
  // periods of a launch grouped by view (scratch of view_group_kernel / cmac_kernel)
  int *grp_sorted;           // [Tmax + 1] period indices relative to k0 - 1, sorted by view
  int2 *grp_items;           // [Tmax + 1] (first, count) into grp_sorted
  int *grp_count;            // [2] number of items, cMAC splits in use (<= KS)
  // one scene shared by W GPUs (bear_cuda_comm_connect; W = 1: this renderer alone): every rank mixes its
  // own channels onto its own bus rows (cdir, qdif), and convolves loudspeakers [l0, l1) of the SUM of all
  // ranks' rows, which bus_gather_kernel collects into cin / qin (W = 1: cin = cdir, qin = qdif)
  int W, rank, l0, l1;
  int dif;                   // some rank has objects: the diffuse path runs
  const float *cin, *qin;    // what decorrelate_kernel / bus_fdl_kernel read, same layout as cdir / qdif
  const float *peer_cdir[kMaxRanks], *peer_qdif[kMaxRanks];   // every rank's bus rows (entry `rank` = own)
  float *mix_recs;           // [Tmax][N+M][kMixRecWords] records of mix_prep_kernel
  unsigned *mix_keys;        // [Tmax][N+M] sort keys (loudspeaker sets)
  unsigned *mix_order;       // [Tmax][N+M] channels of every object split sorted by key (mix_order_kernel)
  int *mix_counts;           // [Tmax][kMaxObjectSplits][2] listed / dense channels of every split
  int *err_flag;             // error bits raised by kernels
  int ks_dev;                // this launch's number of splits is grp_count[1] (grouped launch), not KS
  int k0;                    // first period of this launch (a call may be issued in chunks)
};

struct CarryRow {
  float *base;
  int rows;
  long long stride;   // floats between rows
  int count;          // floats to move per row
  long long shift;    // source offset (floats) relative to destination
  long long total;    // > 0: the rows are pieces of ONE run of `total` floats (the last piece is shorter)
};
constexpr int kMaxCarry = 16;
struct CarryPlan {
  CarryRow e[kMaxCarry];
  int row_start[kMaxCarry + 1];
  int n;
  // frequency-domain delay line: slots [T, T+P) -> [0, P) of every channel
  float2 *fdl;
  long long fdl_stride;   // float2 per channel
  int fdl_channels, fdl_B, fdl_P, fdl_T;
};

void launch_filter_fft(const float *taps, int n_filters, int taps_len, int B, int logB, int P, const float2 *tw,
                       float2 *out, float scale, cudaStream_t st);
void launch_mix_prep(const Layout &p, int T, cudaStream_t st);
void launch_mix(const Layout &p, int T, cudaStream_t st);
void launch_bus_sum(const Layout &p, int T, cudaStream_t st);
void launch_bus_gather(const Layout &p, int T, cudaStream_t st, bool thin = false);

// Synchronisation between the GPUs of a shared scene through flags in each other's memory: rank w's
// flag array has one counter per rank, written by that rank with a system-scope release store.
struct PeerSync {
  int *flags[kMaxRanks];     // [w]: rank w's flag array (this rank's own, or the cudaIpc mapping of a peer's)
  int W, rank;
  int value;                 // the counter value of this synchronisation point (same on every rank)
  unsigned signal, wait;     // bit w: store `value` into rank w's array / spin until rank w's counter has reached it
  int *err;
};
void launch_peer_sync(const PeerSync &s, cudaStream_t st);
// out[e][col0 + i] = sum over ranks w of inbox[w][e][col0 + i], i < n (rank order: the result does not depend on timing)
void launch_out_sum(float *out, const float *inbox, size_t os, int W, size_t col0, size_t n, cudaStream_t st);
void launch_validate_views(const Layout &p, int T, cudaStream_t st);
void launch_hoa_mix(const Layout &p, int T, cudaStream_t st);
void launch_decorrelate(const Layout &p, int T, cudaStream_t st);
void launch_bus_fdl(const Layout &p, int T, cudaStream_t st);
int launch_cmac(const Layout &p, int T, cudaStream_t st);   // returns the number of kernels launched
bool cmac_groups_by_view(const Layout &p, int T);            // whether a launch of T periods is grouped by view
void launch_ifft_out(const Layout &p, int T, cudaStream_t st);
void launch_carry(const CarryPlan &plan, cudaStream_t st);

// single-period host path: one packed upload (audio + tables), scattered on the device
struct UnpackItem { const float *src; float *dst; int count; };
constexpr int kMaxUnpack = 10;
struct UnpackPlan { UnpackItem item[kMaxUnpack]; int n; const float *audio; int rows; };
void launch_unpack(const Layout &p, const UnpackPlan &plan, cudaStream_t st);
void launch_control(const Layout &p, int T, cudaStream_t st);
// interleaved file frames (format: 0 f32, 1 s16, 2 packed s24, 3 s32) -> planar input rows
void launch_frames(const Layout &p, const void *frames, int format, int frame_channels, const int *tracks, int rows,
                   int first_frame, int nframes, cudaStream_t st);
cudaError_t configure_kernels(int B);
cudaError_t configure_mix(int hist, int B);
size_t mix_smem_bytes(int hist);

// Fixed-size device scratch that lives in the library image (no allocation), and its device
// address for the current device (resolved on first use per device).
constexpr size_t kStaticScratchBytes = 4u << 20;
cudaError_t static_scratch_pointer(float2 **out);

}  // namespace bearcuda
