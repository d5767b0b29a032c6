/* bear_cuda.h -- C ABI of the B200-native BEAR render core (libbear_cuda.so).
 *
 * The library replaces ONE thing in the reference: the audio DSP graph `bear::DSP`
 * (visr_bear/src/dsp.cpp:7-236) that `bear::Renderer::process()` runs once per period
 * (visr_bear/src/api.cpp:138-154).  Its inputs are exactly the Control->DSP parameter ports of
 * visr_bear/src/top.cpp:34-51 plus the audio; everything above (metadata, libear gain
 * calculation, listener handling) stays on the host in the reference's own code.
 *
 * Conventions: plain pointers and sizes, no C++ types; the caller owns every host buffer for
 * the duration of a call; the library owns all device state; every function returns
 * BEAR_CUDA_OK (0) or a negative error code, and bear_cuda_last_error() gives the message.
 * A renderer is single-threaded like bear::Renderer (visr_bear/src/dynamic_renderer.hpp:14-15).
 * There is no CPU fallback: creation fails if no CUDA device is usable.
 */
#ifndef BEAR_CUDA_H
#define BEAR_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BEAR_CUDA_OK 0
#define BEAR_CUDA_ERR_INVALID_ARGUMENT (-1) /* std::invalid_argument in the reference */
#define BEAR_CUDA_ERR_DATA_FILE (-2)        /* std::runtime_error "BEAR data file error: ..." / tensorfile format_error */
#define BEAR_CUDA_ERR_CUDA (-3)             /* CUDA runtime failure */
#define BEAR_CUDA_ERR_UNSUPPORTED (-4)      /* configuration outside what the kernels support */

typedef struct bear_cuda_renderer bear_cuda_renderer;

/* Replaces bear::ConfigImpl (visr_bear/src/config_impl.hpp:6-14) + DSP construction parameters. */
typedef struct bear_cuda_config {
  uint32_t num_objects_channels;         /* N */
  uint32_t num_direct_speakers_channels; /* M */
  uint32_t num_hoa_channels;             /* H */
  uint32_t period_size;                  /* B: power of two, 128..4096 */
  uint32_t sample_rate;                  /* informational; the data file's fs converts delays */
  uint32_t max_batch_blocks;             /* largest T accepted by the *_batch calls (>=1) */
  int32_t device;                        /* CUDA device ordinal */
  uint32_t max_delay_samples;            /* largest table delay in samples; 0 = the longest delay the data
                                          * file can produce (max loudspeaker delay + decorrelation delay, or
                                          * the HOA delay).  A table delay outside [0, history_samples - 4]
                                          * makes the call fail with BEAR_CUDA_ERR_INVALID_ARGUMENT. */
  uint32_t object_splits;                /* 0 = chosen per launch from N and the launch length */
  uint32_t cmac_splits;                  /* 0 = choose from B and max_batch_blocks */
  /* Ramp end-point conventions of the VISR atoms, ramp[n] = (n + offset)/B, offset 0 or 1
   * (the VISR source is not in the reference tree; see DESIGN.md "Oracle pinning"). */
  uint32_t gain_ramp_offset;
  uint32_t delay_ramp_offset;
  uint32_t xfade_ramp_offset;
  uint32_t chunk_blocks;                 /* periods per launch group / upload chunk (0 = automatic) */
  uint32_t reserved[3];
} bear_cuda_config;

/* The data-file tables `Panner` hands to the DSP constructor (visr_bear/src/panner.cpp:52-137,
 * accessors :250-263,:408-412).  Dense, C order, host memory. */
typedef struct bear_cuda_data {
  uint32_t num_views;                /* V */
  uint32_t num_virtual_loudspeakers; /* L (<= 24) */
  uint32_t brir_length;              /* any length: cut into ceil(brir_length / period_size) partitions */
  uint32_t decorrelator_length;      /* <= period_size (one partition), like hoa_ir_length */
  uint32_t num_hoa_decoder_channels; /* A = (order+1)^2 (<= 64) */
  uint32_t hoa_ir_length;
  const float *brirs;                /* [V][L][2][brir_length] */
  const float *delays;               /* [V][L][2] samples (only used for the default delays) */
  const float *decorrelation_filters;/* [L][decorrelator_length] */
  const float *hoa_irs;              /* [A][2][hoa_ir_length] */
  double fs;
  double decorrelation_delay;        /* samples */
  double hoa_delay;                  /* samples */
  uint32_t front_loudspeaker;
  uint32_t num_gain_norm_delays;     /* S, 0 = the file has no "gain_norm_quick" table */
  const float *gain_norm_factors;    /* [S][V][2L][2L][2] (panner.cpp:113-123) or NULL */
} bear_cuda_data;

/* Per-block tables = the DSP parameter input ports (visr_bear/src/dsp.cpp:30-53,89-96).
 * For the batch calls every pointer addresses T consecutive blocks of the stated per-block
 * shape.  In the single-block call a NULL pointer / negative index means "unchanged since the
 * last block" (ports that the reference only writes on change). */
typedef struct bear_cuda_block_tables {
  const float *direct_gains;           /* [L][N]   direct_gains_in            dsp.cpp:32-34  */
  const float *diffuse_gains;          /* [L][N]   diffuse_gains_in           dsp.cpp:37-39  */
  const float *direct_delays;          /* [2N] s   direct_delays_in           dsp.cpp:30-31  */
  const float *direct_speakers_gains;  /* [L][M]   direct_speakers_gains_in   dsp.cpp:50-53  */
  const float *direct_speakers_delays; /* [2M] s   direct_speakers_delays_in  dsp.cpp:47-49  */
  const float *static_delays;          /* [2L] s   static_delays_in           dsp.cpp:91-92  */
  const int32_t *brir_index;           /* [1]      brir_index_in              dsp.cpp:89     */
  const float *hoa_gains;              /* [A][H]   hoa_gains_in               dsp.cpp:94-96  */
} bear_cuda_block_tables;

/* ---- construction: replaces DSP::DSP + Panner table access (dsp.cpp:7-206) ------------------ */

/* Build a renderer from in-memory tables.  Uploads the tables and transforms all V*2L BRIRs,
 * the decorrelators and the HOA decode filters to partitioned spectra on the device. */
int bear_cuda_create(const bear_cuda_config *config, const bear_cuda_data *data, bear_cuda_renderer **out);

/* Same, reading a BEAR data file (TENF) with the checks and error strings of Panner::Panner
 * (panner.cpp:52-137), check_data_file_version (data_file.cpp:16-20) and tensorfile::read
 * (tensorfile.cpp:187-215).  On failure *out is NULL and bear_cuda_last_error(NULL) has the text. */
int bear_cuda_create_from_file(const bear_cuda_config *config, const char *data_path, bear_cuda_renderer **out);

/* A further renderer that shares the filter spectra and data-file tables of `share_with` (same
 * device, same period size; channel counts, batch size etc. are free).  The tables are reference
 * counted and live until the last renderer using them is destroyed.  Construction is then a few
 * allocations instead of the transform of all V*2L BRIRs -- what DynamicRenderer's background thread
 * exists to hide in the reference (constructor_thread.cpp) -- and N instances for a batch of files
 * hold one copy of the 212 MB table. */
/* (No counterpart in the reference: every bear::Renderer loads and transforms its own tables,
 * api.cpp:238-251.) */
int bear_cuda_create_shared(const bear_cuda_config *config, const bear_cuda_renderer *share_with, bear_cuda_renderer **out);

/* Replaces bear::Renderer::~Renderer (api.cpp:285). */
void bear_cuda_destroy(bear_cuda_renderer *r);

/* Put the renderer back into the state of a freshly constructed one (all histories, the frequency-
 * domain delay line and the carried tables cleared; filter spectra kept), so that one instance can
 * render a batch of independent files one after the other.  The reference builds a new
 * bear::Renderer per file (bear/render_cli.py:503-512), which re-transforms all BRIRs each time. */
int bear_cuda_reset(bear_cuda_renderer *r);

/* Message of the last failure on this renderer (or, with r == NULL, of the last failed create on
 * this thread).  Valid until the next call on the same renderer/thread.  Stands in for the what()
 * of the exceptions the reference throws (std::invalid_argument api.cpp:68-72,158-159;
 * std::runtime_error data_file.cpp:16-20, panner.cpp:43-46): no exception crosses the C ABI. */
const char *bear_cuda_last_error(const bear_cuda_renderer *r);

/* ---- per-period processing: replaces what flow.process() does for the DSP atoms -------------- */

/* One period, host buffers, same buffer contract as bear::Renderer::process (api.cpp:253-259,
 * include/bear/api.hpp:165-177): objects/direct_speakers/hoa are arrays of N/M/H pointers to
 * period_size floats (the array may be NULL when the count is 0), output is 2 pointers. */
int bear_cuda_process(bear_cuda_renderer *r, const float *const *objects_input,
                      const float *const *direct_speakers_input, const float *const *hoa_input,
                      const bear_cuda_block_tables *tables, float *const *output);

/* T consecutive periods in one call (offline rendering: bear-render pushes whole files block by
 * block, bear/render_cli.py:398-465).  objects/direct_speakers/hoa: [count][T*B] channel-major
 * host arrays with the given row strides (in floats); tables: T blocks each; output: [2][T*B]
 * with row stride out_stride.  State carries over between calls exactly as between periods. */
int bear_cuda_process_batch(bear_cuda_renderer *r, uint32_t T, const float *objects, size_t objects_stride,
                            const float *direct_speakers, size_t direct_speakers_stride, const float *hoa,
                            size_t hoa_stride, const bear_cuda_block_tables *tables, float *output,
                            size_t out_stride);

/* The same straight from a file's interleaved sample frames (frame f = one sample of each of the
 * file's frame_channels channels), the form in which bear-render holds a block it has read from
 * the BW64 file before it slices it per track (bear/render_cli.py:205-213; the slicing and the
 * PCM decoding happen in ear.fileio / the Python wrapper, which are not in the reference tree, so
 * this entry point has no single counterpart there).  frames: [T*B][frame_channels] samples of
 * `format`; tracks: [N+M+H] file channel of every input in the order objects, DirectSpeakers,
 * HOA, -1 = silence.  PCM samples are scaled by 2^-(bits-1); S24 is packed 3-byte little-endian.
 * De-interleaving and conversion run on the device, so a 16-bit file costs half the upload of its
 * float form.  Equal, bit for bit, to bear_cuda_process_batch on the converted samples. */
#define BEAR_CUDA_FRAMES_F32 0u
#define BEAR_CUDA_FRAMES_S16 1u
#define BEAR_CUDA_FRAMES_S24 2u
#define BEAR_CUDA_FRAMES_S32 3u
int bear_cuda_process_batch_frames(bear_cuda_renderer *r, uint32_t T, const void *frames, uint32_t format,
                                   uint32_t frame_channels, const int32_t *tracks,
                                   const bear_cuda_block_tables *tables, float *output, size_t out_stride);

/* ---- control tail on the device ----------------------------------------

 * Replaces, for T periods at once, the table producers that sit between the gain calculators and
 * the DSP ports in the reference's Control graph (visr_bear/src/control.cpp:56-85):
 *   DirectDelayCalc (direct_delay_calc.cpp:25-42), GainNorm (gain_norm.cpp:26-44 ->
 *   Panner::compensation_gain, panner.cpp:277-404), DirectDiffuseSplit (direct_diffuse_split.cpp:24-32),
 *   DirectSpeakersDelayCalc, DirectSpeakersGainNorm, StaticDelayCalc (static_delay_calc.cpp:19-30).
 * Inputs are what the gain calculators emit: object gains [2L][N] per period (rows 0..L-1 direct,
 * L..2L-1 diffuse; GainCalcObjects gains_out), DirectSpeakers gains [L][M], the selected view, and
 * the HOA matrix (passed through).  The kernel writes the direct/diffuse gain, delay and static-delay
 * tables of the T periods straight into the renderer's device staging areas; follow with
 * bear_cuda_process_device, or bear_cuda_process_batch / bear_cuda_process with tables == NULL.
 * With a data file that has no gain_norm_quick table the compensation gain is 1 (panner.cpp:368-369). */
int bear_cuda_control_tables(bear_cuda_renderer *r, uint32_t T, const float *object_gains,
                             const float *direct_speakers_gains, const int32_t *brir_index, const float *hoa_gains);

/* ---- device-resident variant (inputs already in HBM) ---------------------------------------- */

/* Device staging areas owned by the renderer.  A producer that already has audio / tables on
 * the device writes them here and calls bear_cuda_process_device; nothing crosses PCIe. */
typedef struct bear_cuda_device_io {
  float *audio_in;         /* [N+M+H][audio_stride]; block data starts at column audio_offset */
  size_t audio_stride;
  size_t audio_offset;
  float *direct_gains;     /* [T][L][N] */
  float *diffuse_gains;    /* [T][L][N] */
  float *direct_delays;    /* [T][2N] */
  float *direct_speakers_gains;  /* [T][L][M] */
  float *direct_speakers_delays; /* [T][2M] */
  float *static_delays;    /* [T][2L] */
  int32_t *brir_index;     /* [T] */
  float *hoa_gains;        /* [T][A][H] */
  float *audio_out;        /* [2][out_stride] */
  size_t out_stride;
  void *stream;            /* cudaStream_t all work of this renderer is ordered on */
} bear_cuda_device_io;

/* No counterpart in the reference (its ports are host memory, dsp.hpp:46-71): the same port tables
 * and audio rows as bear_cuda_process_batch, already on the device. */
int bear_cuda_get_device_io(bear_cuda_renderer *r, bear_cuda_device_io *io);

/* Render T periods from the staging areas; asynchronous on io.stream.  What flow.process() does
 * (api.cpp:151-153) for T periods, with nothing crossing PCIe. */
int bear_cuda_process_device(bear_cuda_renderer *r, uint32_t T);

/* No counterpart in the reference (its DSP graph owns ordinary heap buffers).
 *
 * Same work, inputs, outputs and carried state as bear_cuda_process_device, for hosts that drive
 * the renderer from a place where nothing may be allocated: the audio callback of a plug-in host or
 * sound-card driver thread, where a device allocation can take driver-wide locks and stall the
 * period deadline, or an embedder whose device memory is handed out once at start-up.  This call
 * performs no cudaMalloc/cudaFree and uses no allocating library routine, on the device or on the
 * host heap (argument errors are reported from a buffer reserved at creation; only the report of a
 * CUDA runtime failure may allocate); the one piece of per-call scratch the pipeline needs (the
 * partial output spectra between the multiply-accumulate and the inverse transform) is taken from
 * an array of BEAR_CUDA_STATIC_SCRATCH_BYTES that is part of the library image on every GPU that
 * loads it, instead of from renderer-owned memory.  Per-kernel profiling is off during the call.
 *
 *  - Limit: T * 2 ears * 2 views * cmac_splits * period_size * 8 bytes must fit the array
 *    (bear_cuda_get_info reports cmac_splits); otherwise BEAR_CUDA_ERR_UNSUPPORTED with the numbers
 *    in bear_cuda_last_error and nothing is launched.
 *  - The array's device address is looked up when the renderer is created (with its device
 *    current), so the render call itself does no set-up.
 *  - The array is shared by every renderer and every call on the device and keeps whatever the last
 *    call wrote.  Nothing needs clearing (all that is read was written in the same call), but two
 *    calls whose kernels overlap on the GPU - different renderers on different streams - would
 *    overwrite each other's partial spectra: serialise them. */
#define BEAR_CUDA_STATIC_SCRATCH_BYTES (4u << 20)
int bear_cuda_process_device_static_scratch(bear_cuda_renderer *r, uint32_t T);

/* Block until all work queued on the renderer's stream has finished (no counterpart: the
 * reference's process() is synchronous).  Also reports what the kernels of the asynchronous calls
 * found wrong with device-resident tables (a delay outside [0, history_samples - 4], a brir_index
 * outside [0, V)): BEAR_CUDA_ERR_INVALID_ARGUMENT, once; the offending values were clamped / replaced by
 * view 0 for that call, so no memory outside the renderer's buffers was touched. */
int bear_cuda_synchronize(bear_cuda_renderer *r);

/* ---- one scene on several GPUs ---------------------------------------------
 *
 * No counterpart in the reference (it has no distributed code); the property relied on is that a
 * render is linear in its inputs, which the reference pins in visr_bear/test/test_changes.py:341-361.
 * One process per GPU, each with its own renderer for ITS share of the scene's objects /
 * DirectSpeakers / HOA channels (any split, counts may differ) and the same listener tables.  After
 * bear_cuda_comm_connect the renderers work as one:
 *   - every rank mixes its channels onto its own copy of the 2L + L bus rows (per-ear delays, gain
 *     ramps: the part whose cost grows with the object count);
 *   - rank g then runs decorrelators, static delays, FFTs and the partitioned BRIR multiply-accumulate
 *     for loudspeakers [g L / W, (g+1) L / W) only, on the SUM of all ranks' rows, which it gathers
 *     from the peers' memory over NVLink (cudaIpc mappings, ordinary loads).  The ranks are ordered by
 *     counters they store into each other's memory (release / acquire at system scope): a rank reads
 *     a peer's rows once that peer's counter says they are complete;
 *   - with root >= 0 every rank stores its stereo partial mix straight into an inbox on the root
 *     (ordinary stores over NVLink) and the root adds the W partial mixes up in rank order; with
 *     root < 0 the partial mixes are all-reduced with NCCL and every rank holds the result.  Either
 *     way per chunk of periods, so `output` / audio_out of the root (of every rank) holds the whole
 *     scene's render.  On other ranks the process calls accept output == NULL.
 *   A rank that does not reach a synchronisation point within ~20 s makes its peers' next
 *   bear_cuda_synchronize (or host-buffer call) fail with BEAR_CUDA_ERR_CUDA instead of hanging.
 * Every rank must make the same sequence of process calls (same T, same chunking).  NCCL is bound at
 * run time (dlopen of libnccl.so.2: the copy already in the process, e.g. PyTorch's, else the
 * system's); a renderer that is never connected does not need it.
 *
 * Protocol: each rank fills a blob (bear_cuda_comm_blob; rank 0's carries the NCCL unique id), the
 * host gathers the blobs of all ranks by whatever means it has (torch.distributed, MPI, a file) and
 * hands every rank the concatenation, ordered by rank.  Connect before the first process call. */
#define BEAR_CUDA_COMM_BLOB_BYTES 256
int bear_cuda_comm_blob(bear_cuda_renderer *r, uint32_t rank, void *blob);
int bear_cuda_comm_connect(bear_cuda_renderer *r, uint32_t rank, uint32_t world, const void *blobs, int32_t root);
/* Which loudspeakers rank `rank` of `world` convolves (host only, no GPU needed; no counterpart in the
 * reference). */
void bear_cuda_comm_speaker_range(uint32_t rank, uint32_t world, uint32_t num_virtual_loudspeakers,
                                  uint32_t *begin, uint32_t *end);

/* Pipelined calls of a shared scene (no counterpart in the reference, whose process() returns with the
 * period's output written, renderer.hpp:45-52).  Same work and same results as bear_cuda_process_device,
 * but the call returns with the exchange and the convolution of its T periods still queued on a second
 * stream, so that the mix of the next call's periods runs beside them: with W GPUs the convolution
 * side is short and mostly waits for its peers, and this hides that wait.  Rules:
 *   - the staging areas (audio, tables) may be rewritten on the render stream (bear_cuda_device_io.stream)
 *     as soon as the call has returned: the convolution side works on its own copies of the tables;
 *   - the output of call s is complete on the stream bear_cuda_result_stream returns, in stream order
 *     after the call; read it there before queueing call s+1 (the output rows are reused every call);
 *   - any other entry point first lets the render stream wait for the pending convolution side, and
 *     behaves as documented from then on; bear_cuda_synchronize waits for both streams;
 *   - a renderer that is not connected, has HOA inputs on this rank, a chunk_blocks limit below T, or
 *     per-kernel profiling switched on runs the call exactly as bear_cuda_process_device does. */
int bear_cuda_process_device_pipelined(bear_cuda_renderer *r, uint32_t T);
/* The stream on which the output of a bear_cuda_process_device_pipelined call completes (the renderer's
 * second stream once connected, else the render stream); no counterpart in the reference. */
int bear_cuda_result_stream(bear_cuda_renderer *r, void **stream);

/* ---- introspection --------------------------------------------------------------------------- */

typedef struct bear_cuda_info {
  uint32_t num_partitions;      /* P = ceil(brir_length / B) */
  uint32_t object_splits, cmac_splits;
  uint32_t history_samples;
  uint32_t kernels_per_call;    /* kernel launches one process call issues */
  uint64_t device_bytes;        /* device memory owned by this renderer (state, staging) */
  uint64_t launches_total;      /* kernels launched since creation */
  double default_direct_delay, default_static_delay; /* seconds, panner.cpp:172-186 */
  uint64_t shared_bytes;        /* device memory of the (possibly shared) filter spectra and tables */
  /* the most recent launch whose periods were grouped by BRIR view (>= 8 periods): number of
   * (view, <= 6 periods) work items and of partial spectra per period the device chose; 0 if none yet */
  uint32_t cmac_items_last, cmac_splits_last;
  /* scene sharing (bear_cuda_comm_connect): number of ranks, this rank, and the virtual loudspeakers
   * [speakers_begin, speakers_end) whose convolution this rank runs; 1, 0, 0, L when not connected */
  uint32_t world, rank, speakers_begin, speakers_end;
} bear_cuda_info;

/* No counterpart in the reference (num_partitions is what rbbl::CoreConvolverUniform derives from
 * the filter length, dsp.cpp:73-87; the default delays are panner.cpp:172-186). */
int bear_cuda_get_info(const bear_cuda_renderer *r, bear_cuda_info *info);

/* Per-kernel device timing (CUDA events on the renderer's stream, recorded around every kernel
 * while enabled); no counterpart in the reference (its benchmark times whole process() calls,
 * visr_bear/test/benchmark.cpp:97-201).  bear_cuda_get_profile synchronises the stream and fills
 * BEAR_CUDA_NUM_KERNELS entries, indexed by the BEAR_CUDA_KERNEL_* ids. */
#define BEAR_CUDA_KERNEL_MIX 0
#define BEAR_CUDA_KERNEL_HOA_MIX 1
#define BEAR_CUDA_KERNEL_DECORRELATE 2
#define BEAR_CUDA_KERNEL_BUS_FDL 3
#define BEAR_CUDA_KERNEL_CMAC 4
#define BEAR_CUDA_KERNEL_IFFT_OUT 5
#define BEAR_CUDA_KERNEL_CARRY 6
#define BEAR_CUDA_KERNEL_CONTROL 7
#define BEAR_CUDA_KERNEL_MIX_AUX 8   /* mix_prep_kernel (+ bus_sum_kernel in short launches) */
#define BEAR_CUDA_KERNEL_EXCHANGE 9  /* cross-GPU synchronisation and the sum of the stereo partial mixes of a shared scene */
#define BEAR_CUDA_KERNEL_GATHER 10   /* bus_gather_kernel: the peers' bus rows over NVLink */
#define BEAR_CUDA_NUM_KERNELS 11

typedef struct bear_cuda_kernel_profile {
  char name[32];
  double total_ms;
  uint64_t launches;
} bear_cuda_kernel_profile;

int bear_cuda_set_profiling(bear_cuda_renderer *r, int enable);
int bear_cuda_get_profile(bear_cuda_renderer *r, bear_cuda_kernel_profile *out, int reset);

/* ---- data-file access (host only, no GPU needed) ---------------------------------------------- */

/* Replaces bear::DataFileMetadata::read_from_file (include/bear/api.hpp:261-287,
 * visr_bear/src/api.cpp:330-353) plus the dimensions Panner exposes (panner.hpp:55-57,84-106). */
typedef struct bear_cuda_data_file_info {
  int32_t has_metadata, is_released;
  char label[256];
  char description[1024];
  uint32_t num_views, num_virtual_loudspeakers, brir_length, decorrelator_length;
  uint32_t num_hoa_decoder_channels, hoa_ir_length, front_loudspeaker, reserved;
  double fs, decorrelation_delay, hoa_delay;
} bear_cuda_data_file_info;

int bear_cuda_read_data_file_info(const char *data_path, bear_cuda_data_file_info *info);

/* Read one array of a TENF file as float32 in C order (replaces tensorfile::read +
 * TensorFile::unpack, visr_bear/src/tensorfile.cpp:165-215).  `key` is a '/'-separated path into
 * the JSON metadata, e.g. "hoa/irs".  shape receives up to 8 extents; *ndim their number; at most
 * capacity floats are written; *count receives the element count of the array. */
int bear_cuda_tenf_read_f32(const char *path, const char *key, float *out, size_t capacity, size_t *count,
                            uint32_t *shape, uint32_t *ndim);

/* Pinned host memory helpers for callers that want true asynchronous copies (no counterpart in the
 * reference: its buffers never leave the host). */
void *bear_cuda_host_alloc(size_t bytes);
void bear_cuda_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* BEAR_CUDA_H */
