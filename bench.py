#!/usr/bin/env python
"""Benchmark of the BEAR render hot path (BASELINE.json metric: object-channels rendered in real
time at 48 kHz, N x RTF).

Default workload = BASELINE.json configs[3], the configuration the north_star target is quoted on:
ONE scene of 1024 simultaneous moving Objects (new metadata every period), period 256, the listener
yaw stepping 2 degrees per period so that every period crossfades to a new BRIR view.  A step = one
pass of the hot path over 10 s of that scene (1872 periods).  With --gpus N the scene is sharded by
object group, 1024 / N objects per rank (strong scaling): every rank mixes its objects onto the 48 + 24
bus rows, convolves 24 / N of the loudspeakers on the sum of all ranks' rows (read from the peers over
NVLink) and the stereo partial mixes are summed on rank 0 -- all of it inside the library
(bear_cuda_comm_connect); torch.distributed only carries the 256-byte set-up blobs and the timing
barrier.  `--workload cfg2_ht` is BASELINE configs[1] (64 moving Objects per GPU, period 512, weak
scaling), the headline of round 1; at N = 1 it is also reported as the extra key `configs1_64_objects`.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]

ours:      value = device-resident throughput (audio and tables already in HBM), CUDA events on the
           renderer's stream, max over ranks; e2e = the same metric through the host-buffer C-ABI call
           (bear_cuda_process_batch: pinned host -> device -> kernels -> host).
reference: the CPU implementation of the same path (the oracle port; the reference renderer itself
           needs VISR + Boost, which are not available) on all host cores, one instance per core.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FS = 48000
METRIC = "object-channels rendered real-time at 48 kHz per B200 (N x RTF)"
UNIT = "object-channels x RTF"
GROUP = 64    # objects per generated group (seed = group index): the scene does not depend on the GPU count
WORKLOADS = {
    # name: (period, objects, scaling, audio scale, text)
    "cfg4_ht": (256, 1024, "strong", 0.02,
                "BASELINE configs[3]: 1024 simultaneous moving Objects, per-period metadata, block 256, listener yaw "
                "2 deg/block (view crossfade every block), sharded by object group"),
    "cfg2_ht": (512, 64, "weak", 0.1,
                "BASELINE configs[1]: 64 moving Objects per GPU, per-block metadata, block 512, listener yaw 2 deg/block "
                "(view crossfade every block)"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg4_ht", choices=sorted(WORKLOADS))
    ap.add_argument("--seconds", type=float, default=10.0, help="audio seconds per step")
    ap.add_argument("--cpu-sample-blocks", type=int, default=0, help="blocks of the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="only the headline workload")
    ap.add_argument("--no-frames-e2e", action="store_true", help="skip the 16-bit-frames form of the e2e call")
    ap.add_argument("--no-pipeline", action="store_true",
                    help="N > 1: plain bear_cuda_process_device calls instead of the pipelined ones")
    ap.add_argument("--chunk-blocks", type=int, default=0, help="periods per launch group (0 = library default)")
    ap.add_argument("--cmac-splits", type=int, default=0, help="partial spectra per period (0 = library default)")
    ap.add_argument("--object-splits", type=int, default=0, help="mix CTAs per tile (0 = library default)")
    ap.add_argument("--batch-files", type=int, default=256, help="files of the configs[4] extra (0 = skip)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------- workload
_DATA = {}


def data_and_control(num_views=180, seed=0):
    from bear_b200 import control, data_file
    key = (num_views, seed)
    if key not in _DATA:
        data = data_file.from_tree(data_file.synthetic_tree(seed=seed, num_views=num_views))
        _DATA[key] = (data, control.HostControl(data))
    return _DATA[key]


def scene_tables(ctl, T, B, first_object, count, dense_every=0, num_views=180):
    """Tables of objects [first_object, first_object + count) of the scene, generated group by group
    (seed = group index).  Azimuths ~U(-180,180) change every period (benchmark.cpp:121-147
    `update_every_time`), every 4th object is half diffuse, the listener yaw steps 2 degrees per
    period.  dense_every = k: every k-th object gets all 24 gains non-zero, direct and diffuse (the
    reference benchmark's `extent` objects, visr_bear/test/benchmark.cpp:132-139,274-277)."""
    from bear_b200 import control
    assert first_object % GROUP == 0 and count % GROUP == 0
    yaw = (2.0 * np.arange(T)) % 360.0
    parts = []
    for g in range(first_object // GROUP, (first_object + count) // GROUP):
        rng = np.random.default_rng(1000 + g)
        az = rng.uniform(-180, 180, size=(T, GROUP))
        dif = np.where(np.arange(GROUP) % 4 == 3, 0.5, 0.0)
        tb = ctl.tables(T, obj_az=az, obj_diffuse=dif, yaw_deg=yaw)
        if dense_every:
            L = ctl.L
            for i in range(0, GROUP, dense_every):
                w = rng.uniform(0.05, 1.0, size=(T, L))
                w /= np.sqrt((w ** 2).sum(-1, keepdims=True))
                tb["direct_gains"][:, :, i] = (w * np.sqrt(0.5)).astype(np.float32)
                tb["diffuse_gains"][:, :, i] = (w * np.sqrt(0.5)).astype(np.float32)
        parts.append(tb)
    tb = control.BlockTables(parts[0])
    for key in ("direct_gains", "diffuse_gains"):
        tb[key] = np.ascontiguousarray(np.concatenate([p[key] for p in parts], axis=2))
    tb["direct_delays"] = np.ascontiguousarray(np.concatenate([p["direct_delays"] for p in parts], axis=1))
    if num_views != 180:  # reduced view table for CPU samples: cycle through the views that exist
        tb["brir_index"] = (np.arange(T) % num_views).astype(np.int32)
        tb["static_delays"] = ctl.static_delays(tb["brir_index"]).astype(np.float32)
    return tb


def scene_audio(first_object, count, samples, scale):
    from bear_b200 import control
    return control.synth_audio(count, samples, seed_base=10000 + first_object, scale=scale)


def periods_per_step(args, B):
    return max(8, int(args.seconds * FS / B) // 8 * 8)


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons at 200 ms during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.proc.wait()
        sm, smax, reasons = [], None, set()
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); smax = float(p[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "samples": len(sm),
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------- CPU legs
_W = {}


def _cpu_worker_init(workload, T, num_views):
    sys.path.insert(0, str(ROOT))
    from oracle import bear_ref
    bear_ref.use_native_build()
    B, N, _, scale, _ = WORKLOADS[workload]
    data, ctl = data_and_control(num_views)
    tb = scene_tables(ctl, T, B, 0, N, num_views=num_views)
    audio = scene_audio(0, N, T * B, scale)
    _W.update(r=bear_ref.RefRenderer(data, N, 0, 0, B), blocks=tb.blocks(), audio=audio, N=N, B=B, T=T)


def _cpu_worker_step(_):
    t0 = time.perf_counter()
    _W["r"].render(_W["audio"], None, None, _W["blocks"])
    return time.perf_counter() - t0


def cpu_sample_blocks(args):
    if args.cpu_sample_blocks:
        return args.cpu_sample_blocks
    return 48 if args.workload == "cfg2_ht" else 24


def oracle_build():
    """The timed CPU legs use a build of the oracle for this machine (-O3 -march=native)."""
    from oracle import bear_ref
    return "oracle/bear_ref.cpp f32 (own radix-2 FFT; g++ %s)" % bear_ref.use_native_build()


def cpu_baseline_single(args):
    """Oracle port on one host core, bounded sample of the same workload (~10-20 s of CPU work)."""
    T = cpu_sample_blocks(args)
    _cpu_worker_init(args.workload, T, num_views=min(180, T))
    _cpu_worker_step(0)  # warm
    reps, total = 0, 0.0
    while total < 10.0 and reps < 64:
        total += _cpu_worker_step(0)
        reps += 1
    N, B = _W["N"], _W["B"]
    audio_s = reps * T * B / FS
    return {"value": N * audio_s / total, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "%d x %d periods of %d (%.2f s of audio) of the same workload (%d objects), %s, 1 thread"
                      % (reps, T, B, audio_s, N, oracle_build())}


def libear_convolver_anchor():
    """The reference's own partitioned convolver (libear BlockConvolver compiled in place, oracle/_ref)
    on the convolution load of one channel -- 2976 taps at block 512 -- next to the oracle's convolver
    on the same load: a sanity anchor for how far the port is from reference-grade CPU code (VISR's
    convolver itself is not in the reference tree)."""
    try:
        from oracle import bear_ref
        B, taps, blocks, reps = 512, 2976, 40, 4
        rng = np.random.default_rng(0)
        x = rng.standard_normal(blocks * B).astype(np.float32)
        h = rng.standard_normal(taps).astype(np.float32)
        out = {}
        for name, fn in (("libear_block_convolver", bear_ref.ear_block_convolve), ("oracle_upols", bear_ref.upols)):
            fn(x[:4 * B], h, B)
            t0 = time.perf_counter()
            for _ in range(reps):
                fn(x, h, B)
            dt = (time.perf_counter() - t0) / (reps * blocks)          # seconds per channel-period
            out[name] = {"us_per_channel_period": dt * 1e6, "rtf_48_channels": (B / FS) / (48 * dt)}
        out["note"] = ("one thread; one channel-convolution of 2976 taps per 512-sample period, filter set-up included "
                       "(amortised over %d periods); rtf_48_channels = real-time factor if one thread ran all 48 bus "
                       "channels" % blocks)
        return out
    except Exception as e:  # the _ref library is optional
        return {"unavailable": str(e)[:200]}


def run_reference(args, rank):
    """--impl reference: the CPU path on all host cores (one renderer instance per core, each renders
    the bounded sample; a bear::Renderer is single-threaded, visr_bear/src/dynamic_renderer.hpp:14-15)."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    T = cpu_sample_blocks(args)
    B, N, scaling, _, _ = WORKLOADS[args.workload]
    build = oracle_build()   # compiled once here; the forked workers inherit the choice
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_cpu_worker_init, initargs=(args.workload, T, min(180, T))) as pool:
        for _ in range(max(1, args.warmup)):
            pool.map(_cpu_worker_step, range(cores), chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_cpu_worker_step, range(cores), chunksize=1)
        wall = time.perf_counter() - t0
    audio_s = T * B / FS
    value = cores * N * audio_s * args.steps / wall
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, T, 1, cores),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d instances x %d periods of %d (%d objects each) per step, %s (the reference "
                                   "renderer needs VISR+Boost: unbuildable here)" % (cores, T, B, N, build)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "libear_convolver_anchor": libear_convolver_anchor(),
    }
    emit(line)


def workload_config(args, T, world, instances=None):
    B, N, scaling, _, text = WORKLOADS[args.workload]
    total = N if scaling == "strong" else N * world
    c = {"workload": text, "objects_total": total, "objects_per_gpu": total // world, "period": B, "periods_per_step": T,
         "audio_seconds_per_step": T * B / FS,
         "data_file": "synthetic default shape (180 views x 24 x 2 x 2976 taps, 128-tap decorrelators)",
         "sharding": ("object groups; bus rows of 24/N loudspeakers gathered from the peers over NVLink, stereo partial "
                      "mixes stored into rank 0's inbox and summed there (bear_cuda_comm_connect); the exchange + convolution of a step run "
                      "beside the mix of the next one (pipelined calls)") if world > 1 else "single GPU",
         "l2": "inputs larger than L2 (audio + tables streamed per step >> 126 MB)"}
    if instances:
        c["cpu_instances"] = instances
    return c


# ------------------------------------------------------------------------------------------- GPU leg
class Leg:
    """One renderer with its share of a scene resident on the device and in pinned host memory."""

    def __init__(self, data, ctl, B, T, first_object, count, scale, local_rank, dense_every=0, **kw):
        from bear_b200 import device_io
        from bear_b200.renderer import Renderer, pinned_empty
        self.T, self.B, self.N = T, B, count
        self.tb = scene_tables(ctl, T, B, first_object, count, dense_every)
        self.audio = scene_audio(first_object, count, T * B, scale)
        self.r = Renderer(data, count, 0, 0, B, max_batch_blocks=T, device=local_rank, **kw)
        self.local_rank = local_rank
        self._pinned_empty = pinned_empty
        self._device_io = device_io

    def make_resident(self):
        self.io = self._device_io.DeviceBuffers(self.r, device_index=self.local_rank)
        self.io.load_audio(self.audio)
        self.io.load_tables(self.tb)
        self.r.synchronize()
        self.stream = self.io.stream

    def make_pinned(self, receives=True):
        pe = self._pinned_empty
        self.h_audio = pe(self.audio.shape); self.h_audio[:] = self.audio
        self.h_tb = type(self.tb)({k: pe(v.shape, v.dtype) for k, v in self.tb.items()})
        for k, v in self.tb.items():
            self.h_tb[k][...] = v
        self.h_out = pe((2, self.T * self.B)) if receives else False
        self.h2d = self.audio.nbytes + sum(v.nbytes for v in self.tb.values())
        self.d2h = self.h_out.nbytes if receives else 0

    def step_resident(self):
        self.r.process_device(self.T)

    def step_pipelined(self):
        self.r.process_device_pipelined(self.T)

    def step_e2e(self):
        self.r.process_batch(self.h_audio, None, None, self.h_tb, output=self.h_out)


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from bear_b200 import parallel

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    B, N, scaling, scale, _ = WORKLOADS[args.workload]
    total = N if scaling == "strong" else N * world
    if total % (world * GROUP):
        raise SystemExit("%d objects do not split into %d groups of a multiple of %d" % (total, world, GROUP))
    mine = total // world
    T = periods_per_step(args, B)
    data, ctl = data_and_control()
    leg = Leg(data, ctl, B, T, rank * mine, mine, scale, local_rank, chunk_blocks=args.chunk_blocks,
              cmac_splits=args.cmac_splits, object_splits=args.object_splits)
    r = leg.r
    if world > 1:
        parallel.connect_renderers(r, rank, world, root=0)
    leg.make_resident()
    stream = leg.stream
    # a shared scene is rendered with pipelined calls: the exchange + convolution of step s run on the
    # renderer's second stream beside the mix of step s+1 (the per-kernel pass below runs unpipelined)
    pipelined = world > 1 and not args.no_pipeline
    result_stream = torch.cuda.ExternalStream(r.result_stream(), device=torch.device("cuda", local_rank))
    step = leg.step_pipelined if pipelined else leg.step_resident

    def barrier():
        r.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        stream.wait_stream(result_stream)    # the end of the last step's convolution side
        e1.record(stream)
        e1.synchronize()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # clocks are sampled from the warm-up through both timed regions (the timed regions alone last
    # only tens of milliseconds); a short pre-roll keeps the GPU under load while the sampler starts
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    barrier()                 # the ranks' host-side set-up took different times: start the shared calls together
    for _ in range(100):      # a fixed count: every rank must issue the same number of collectives
        step()
    r.synchronize()
    for _ in range(max(3, args.warmup)):
        step()
    launches0 = r.get_info().launches_total
    ms = timed(step, args.steps)
    launches = r.get_info().launches_total - launches0
    ms_joined = timed(leg.step_resident, args.steps) if pipelined else ms   # the same steps, one after the other
    # per-kernel times: a second pass with an event pair around every kernel (the events themselves
    # cost a little, so the headline above is taken without them)
    r.set_profiling(True)
    r.get_profile(reset=True)
    ms_prof = timed(leg.step_resident, args.steps)
    prof = r.get_profile(reset=True)
    r.set_profiling(False)
    info_resident = r.get_info()     # incl. the work items / splits the device chose for this launch

    # ---- end to end: pinned host buffers through the host-buffer entry point -------------------
    leg.make_pinned(receives=(rank == 0))
    e2e_steps = max(3, min(args.steps, 10))
    barrier()
    for _ in range(2):
        leg.step_e2e()
    ms_e2e = timed(leg.step_e2e, e2e_steps)
    clk = clocks.stop() if rank == 0 else None
    # per-launch kernel times at the chunk size of the host-buffer path (untimed extra call)
    r.set_profiling(True)
    r.get_profile(reset=True)
    leg.step_e2e()
    e2e_prof = {k: v[0] / max(v[1], 1) for k, v in r.get_profile(reset=True).items() if v[1]}
    r.set_profiling(False)

    # the same call fed with the file's interleaved 16-bit PCM frames (bear_cuda_process_batch_frames):
    # half the audio bytes over PCIe, de-interleaving and conversion on the device, same tables
    ms_s16 = None
    if not args.no_frames_e2e:
        from bear_b200.renderer import pinned_empty
        h_frames = pinned_empty((T * B, mine), np.int16)
        for c0 in range(0, mine, 64):   # (in pieces: the float intermediates of the whole scene are 2 GB each)
            h_frames[:, c0:c0 + 64] = np.clip(np.round(leg.audio[c0:c0 + 64].T * 32768.0), -32768, 32767)
        tracks = np.arange(mine, dtype=np.int32)
        step_s16 = lambda: r.process_batch_frames(h_frames, tracks, leg.h_tb, output=leg.h_out)
        barrier()
        step_s16()
        ms_s16 = timed(step_s16, e2e_steps)
        s16_bytes = h_frames.nbytes + sum(v.nbytes for v in leg.tb.values())
        del h_frames

    # the PCIe floor of that call: the same bytes copied host->device from pinned memory, nothing else
    h_flat = torch.from_numpy(leg.h_audio.reshape(-1))
    d_flat = torch.empty(leg.h2d // 4, dtype=torch.float32, device="cuda")

    def step_copy():
        with torch.cuda.stream(stream):
            d_flat[:h_flat.numel()].copy_(h_flat, non_blocking=True)
            d_flat[h_flat.numel():].copy_(h_flat[:d_flat.numel() - h_flat.numel()], non_blocking=True)

    step_copy()
    ms_copy = timed(step_copy, e2e_steps) / e2e_steps
    del d_flat

    audio_s = T * B / FS
    value = total * audio_s * args.steps / (ms * 1e-3)
    e2e_value = total * audio_s * e2e_steps / (ms_e2e * 1e-3)
    h2d_all, d2h_all = leg.h2d * world, 2 * T * B * 4

    extras = {}
    if not args.no_extras:
        extras = run_extras(args, rank, world, local_rank, data, ctl, leg, barrier)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        peak, which = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks else \
            (6650.0, "fallback (B200_PROFILING.md)")
        info = info_resident
        L, P = data.num_virtual_loudspeakers, info.num_partitions
        nl = info.speakers_end - info.speakers_begin
        KS = info.cmac_splits_last or info.cmac_splits
        views = np.asarray(leg.tb["brir_index"]).astype(np.int64)
        n_views = int(np.unique(views).size)
        fB = 4 * T * B                                   # bytes of one float row of the launch
        # Algorithmic (minimum-traffic) bytes per launch of every kernel on THIS rank, DESIGN.md section 4:
        #   mix          audio in + gain/delay tables in, direct (2L rows) and diffuse (L rows) bus out
        #   decorrelate  diffuse bus in, decorrelated bus out            (this rank's nl loudspeakers)
        #   bus_fdl      direct bus + decorrelated bus in, packed spectra (B x 8 bytes a period) out
        #   cmac         filter partitions of every DISTINCT view of the launch once (the kernel groups
        #                the periods by view), every delay-line spectrum once, partial spectra out
        #   ifft_out     partial spectra in (old-view ones too in a view-change period), 2 rows out
        xf = int(np.count_nonzero(np.diff(np.concatenate([[views[0]], views])) != 0))
        alg_by_kernel = {
            "mix_kernel": mine * fB + 4 * T * (2 * 2 * L * mine + 2 * 2 * mine) + 3 * L * fB,
            "decorrelate_kernel": 2 * nl * fB,
            "bus_fdl_kernel": 2 * nl * fB + nl * fB + 2 * nl * 2 * fB,
            "cmac_kernel": 8 * B * (2 * nl * P * n_views + 2 * nl * T + 2 * 2 * KS * T),
            "ifft_out_kernel": 8 * B * 2 * KS * (T + xf) + 2 * fB,
        }
        step_ms_kernels = sum(v[0] for v in prof.values()) / max(args.steps, 1)
        by_kernel = {}
        for name, (t_k, n_k) in prof.items():
            if not n_k or name not in alg_by_kernel:
                continue
            ms_k = t_k / n_k
            by_kernel[name] = {"avg_launch_ms": ms_k, "algorithmic_bytes_per_launch": int(alg_by_kernel[name]),
                               "achieved": alg_by_kernel[name] / (ms_k * 1e-3) / 1e9, "unit": "GB/s",
                               "frac": alg_by_kernel[name] / (ms_k * 1e-3) / 1e9 / peak,
                               "share_of_step": (t_k / max(args.steps, 1)) / step_ms_kernels}
        dominant = max(by_kernel, key=lambda k: by_kernel[k]["share_of_step"])
        traffic_all = {}
        tpath = ROOT / "profiles" / "kernel_traffic.json"
        if tpath.exists() and world == 1:
            try:
                traffic_all = json.loads(tpath.read_text()).get(args.workload, {})
            except Exception:
                traffic_all = {}
        traffic = traffic_all.get(dominant)
        dom = by_kernel[dominant]
        conv = ("decorrelate_kernel", "bus_fdl_kernel", "cmac_kernel", "ifft_out_kernel")
        conv_bytes = sum(alg_by_kernel[k] for k in conv if k in by_kernel)
        conv_ms = sum(by_kernel[k]["avg_launch_ms"] for k in conv if k in by_kernel)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, T, world),
            "rtf": audio_s * args.steps / (ms * 1e-3),
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": dom["achieved"], "peak": peak, "unit": "GB/s",
                         "frac": dom["frac"], "traffic": traffic, "peak_source": which,
                         "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_launch"],
                         "avg_launch_ms": dom["avg_launch_ms"], "share_of_step": dom["share_of_step"],
                         "traffic_source": "profiles/kernel_traffic.json (ncu --set full of the same command, "
                                           "dram read+write per launch)",
                         "note": "dominant kernel by time (rank 0's launches); no tensor-core work on this path, so "
                                 "every kernel is held against HBM.  A fraction well below 1 with traffic ~ algorithmic "
                                 "bytes means the kernel is bound by FP32 issue / latency, not by re-reads "
                                 "(DESIGN.md section 4)"},
            "roofline_by_kernel": {k: dict(v, traffic=traffic_all.get(k)) for k, v in by_kernel.items()},
            "convolution_kernels": {"algorithmic_bytes_per_launch": int(conv_bytes), "ms": conv_ms,
                                    "achieved": conv_bytes / (conv_ms * 1e-3) / 1e9 if conv_ms else None,
                                    "frac": conv_bytes / (conv_ms * 1e-3) / 1e9 / peak if conv_ms else None,
                                    "note": "decorrelate + bus_fdl + cmac + ifft_out together (north_star: roofline "
                                            "fraction in the convolution kernels)"},
            "cmac": {"distinct_views_in_launch": n_views, "work_items": int(info.cmac_items_last), "splits": int(KS)},
            "kernel_ms_per_step": {k: v[0] / max(args.steps, 1) for k, v in prof.items() if v[1]},
            "ms_per_step_with_kernel_events": ms_prof / args.steps,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": ms_e2e / e2e_steps, "steps": e2e_steps,
                    "kernel_ms_per_chunk_launch": e2e_prof,
                    "h2d_copy_only_ms": ms_copy, "h2d_copy_only_gbs_per_gpu": leg.h2d / ms_copy * 1e-6,
                    "api": "bear_cuda_process_batch (pinned host buffers; every rank uploads its objects, rank 0 "
                           "receives the mix)"},
            "e2e_s16_frames": None if ms_s16 is None else {
                "value": total * audio_s * e2e_steps / (ms_s16 * 1e-3), "unit": UNIT, "ms_per_step": ms_s16 / e2e_steps,
                "h2d_bytes_per_step": int(s16_bytes * world), "d2h_bytes_per_step": d2h_all,
                "api": "bear_cuda_process_batch_frames: the same step from interleaved 16-bit PCM frames as a file "
                       "reader holds them (quantised audio, so not the headline e2e)"},
            "gpu_launches": int(launches),
            "clocks": clk,
        }
        if world > 1:
            # bytes each rank pulls over NVLink per step: its nl loudspeakers' 3 bus rows from W - 1 peers
            line["exchange"] = {
                "nvlink_bytes_read_per_rank_per_step": int(3 * nl * fB * (world - 1)),
                "nvlink_bytes_stored_to_root_per_rank_per_step": int(2 * fB),
                "calls": "bear_cuda_process_device_pipelined" if pipelined else "bear_cuda_process_device",
                "ms_per_step_unpipelined": ms_joined / args.steps,
                "ms_per_step": sum(v[0] for k, v in prof.items() if k == "exchange") / max(args.steps, 1),
                "gather_ms_per_step": sum(v[0] for k, v in prof.items() if k == "bus_gather_kernel") / max(args.steps, 1),
                "note": "ms_per_step = the counter exchange that orders the ranks (includes waiting for the slowest "
                        "rank's mix) + the sum of the stereo partial mixes on rank 0; gather_ms_per_step = "
                        "bus_gather_kernel (the peers' rows over NVLink); CUDA events on rank 0's stream in the "
                        "per-kernel pass"}
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_single(args)
        emit(line)
    r.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_extras(args, rank, world, local_rank, data, ctl, leg, barrier):
    """Numbers next to the headline.  At every N: configs[4] (batch of files, sharded by file).  At
    N = 1 also: configs[1] (64 objects, period 512), the dense-gain variant of the headline scene, the
    interleaved-frames entry point, the one-period real-time call and the control tail on the device."""
    import torch
    import torch.distributed as dist
    from bear_b200 import control, parallel
    from bear_b200.renderer import Renderer, pinned_empty
    extras = {}
    B, N, scaling, scale, _ = WORKLOADS[args.workload]
    T = leg.T

    def event_ms(fn, steps, stream):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1) / steps

    # ---- configs[4]: batch of mixed files sharded by file, no collective ------------------------
    if args.batch_files:
        Bf, Tf, distinct = 512, max(8, int(10.0 * FS / 512) // 8 * 8), 2
        files = []
        for f in range(distinct):
            Nf, Mf, Hf, tbf = control.scenario(ctl, "cfg5_ht", Tf, Bf, seed=100 + f)
            audio = [control.synth_audio(c, Tf * Bf, 1000 * f + s, 0.05) for c, s in ((Nf, 0), (Mf, 100), (Hf, 200))]
            pinned = []
            for a in audio:
                p = pinned_empty(a.shape); p[:] = a
                pinned.append(p)
            ptb = type(tbf)({k: pinned_empty(v.shape, v.dtype) for k, v in tbf.items()})
            for k, v in tbf.items():
                ptb[k][...] = v
            files.append((*pinned, ptb))
        outf = pinned_empty((2, Tf * Bf))
        my_files = parallel.file_shard(rank, world, args.batch_files)
        with Renderer(data, Nf, Mf, Hf, Bf, max_batch_blocks=Tf, device=local_rank) as rf:
            for i in range(2):
                rf.reset()
                rf.process_batch(*files[i % distinct][:3], files[i % distinct][3], output=outf)
            barrier()
            t0 = time.perf_counter()
            for idx in my_files:
                o, d, h, tbf = files[idx % distinct]
                rf.reset()
                rf.process_batch(o, d, h, tbf, output=outf)
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        secs = float(dt.item())
        extras["batch_files"] = {
            "workload": "BASELINE configs[4]: %d files x (16 Objects + 12 DirectSpeakers + 16 HOA channels, %.1f s, "
                        "period 512), sharded by file, no collective" % (args.batch_files, Tf * Bf / FS),
            "files": args.batch_files, "files_per_s": args.batch_files / secs, "wall_s": secs,
            "channels_x_rtf": (Nf + Mf + Hf) * args.batch_files * (Tf * Bf / FS) / secs, "scaling": "strong",
            "api": "bear_cuda_reset + bear_cuda_process_batch (pinned host buffers), wall clock, max over ranks"}
        del files, outf
    if world > 1 or rank != 0:
        return extras

    # ---- the other BASELINE object config on one GPU (configs[1]: 64 moving objects, period 512) ----
    other = "cfg2_ht" if args.workload == "cfg4_ht" else "cfg4_ht"
    Bo, No, _, so, text = WORKLOADS[other]
    To = periods_per_step(args, Bo)
    lo = Leg(data, ctl, Bo, To, 0, No, so, local_rank)
    lo.make_resident()
    for _ in range(20):
        lo.step_resident()
    ms_o = event_ms(lo.step_resident, 10, lo.stream)
    lo.r.set_profiling(True)
    lo.r.get_profile(reset=True)
    for _ in range(4):
        lo.step_resident()
    prof_o = {k: v[0] / max(v[1], 1) for k, v in lo.r.get_profile(reset=True).items() if v[1]}
    lo.r.set_profiling(False)
    lo.make_pinned()
    for _ in range(2):
        lo.step_e2e()
    ms_oe = event_ms(lo.step_e2e, 5, lo.stream)
    a_s = To * Bo / FS
    extras["configs1_64_objects" if other == "cfg2_ht" else "configs3_1024_objects"] = {
        "workload": text, "objects": No, "period": Bo, "periods_per_step": To,
        "value": No * a_s / (ms_o * 1e-3), "ms_per_step": ms_o, "kernel_ms_per_launch": prof_o,
        "e2e": {"value": No * a_s / (ms_oe * 1e-3), "ms_per_step": ms_oe, "h2d_bytes_per_step": lo.h2d}}
    if No <= 64:
        # interleaved file frames (float and 16-bit PCM) on that scene: what a file renderer holds after
        # reading a block of a BW64 file; de-interleaving / PCM decoding on the device
        frames = {}
        tracks = np.arange(No, dtype=np.int32)
        for fmt, dtp, sc in (("f32", np.float32, 1.0), ("s16", np.int16, 32768.0)):
            h_frames = pinned_empty((To * Bo, No), dtp)
            h_frames[...] = np.clip(np.round(lo.audio.T * sc), -sc, sc - 1) if fmt != "f32" else lo.audio.T
            fn = lambda: lo.r.process_batch_frames(h_frames, tracks, lo.h_tb, output=lo.h_out)
            fn()
            ms_f = event_ms(fn, 5, lo.stream)
            frames[fmt] = {"ms_per_step": ms_f, "value": No * a_s / (ms_f * 1e-3)}
        extras["e2e_file_frames"] = dict(frames, note="bear_cuda_process_batch_frames on the 64-object scene: "
                                                      "interleaved frames as read from the file (pinned), same tables")
        # real-time use: one period per call through bear_cuda_process (host pointers), wall clock
        with Renderer(data, No, 0, 0, Bo, max_batch_blocks=1, device=local_rank) as rs:
            blk = [lo.tb.block(k) for k in range(8)]
            xin = [np.ascontiguousarray(lo.audio[:, k * Bo:(k + 1) * Bo]) for k in range(8)]
            out1 = np.empty((2, Bo), np.float32)
            for k in range(16):
                rs.process(xin[k % 8], None, None, blk[k % 8], output=out1)
            t0 = time.perf_counter()
            for k in range(200):
                rs.process(xin[k % 8], None, None, blk[k % 8], output=out1)
            extras["stream"] = {"ms_per_period_call": (time.perf_counter() - t0) / 200 * 1e3, "period_ms": 1e3 * Bo / FS,
                                "objects": No,
                                "note": "bear_cuda_process, one period per call incl. host<->device copies and sync"}
    lo.r.close()
    del lo

    # ---- the headline scene with extent-like objects: every 4th object has all 24 gains non-zero ----
    ld = Leg(data, ctl, B, T, 0, N, scale, local_rank, dense_every=4)
    ld.make_resident()
    for _ in range(5):
        ld.step_resident()
    ms_d = event_ms(ld.step_resident, 5, ld.stream)
    ld.r.set_profiling(True)
    ld.r.get_profile(reset=True)
    for _ in range(3):
        ld.step_resident()
    prof_d = {k: v[0] / max(v[1], 1) for k, v in ld.r.get_profile(reset=True).items() if v[1]}
    extras["dense_gains"] = {
        "workload": "the headline scene with every 4th object dense (24 non-zero direct and diffuse gains: the "
                    "reference benchmark's extent objects, visr_bear/test/benchmark.cpp:132-139,274-277)",
        "objects": N, "value": N * (T * B / FS) / (ms_d * 1e-3), "ms_per_step": ms_d, "kernel_ms_per_launch": prof_d}
    ld.r.close()
    del ld

    # ---- control tail on the device (delays + gain normalisation + split), kernel time only ----
    import copy
    d2 = copy.copy(data)
    S = 40   # default.tf: max loudspeaker delay 38 + 2 (finalise.py:214-215)
    d2.gain_norm_factors = np.random.default_rng(0).random((S, data.num_views, 2 * ctl.L, 2 * ctl.L, 2), dtype=np.float32)
    Tc, Nc = 128, 64
    tbc = scene_tables(ctl, Tc, 512, 0, Nc)
    with Renderer(d2, Nc, 0, 0, 512, max_batch_blocks=Tc, device=local_rank) as rc:
        raw = np.concatenate([tbc["direct_gains"], tbc["diffuse_gains"]], axis=1)
        rc.set_profiling(True)
        for _ in range(4):
            rc.control_tables(raw, None, tbc["brir_index"])
        t_ctl, n_ctl = rc.get_profile()["control_kernel"]
        extras["control_tail"] = {"ms_per_launch": t_ctl / max(n_ctl, 1), "periods_per_launch": Tc, "objects": Nc,
                                  "note": "control_kernel (DirectDelayCalc + GainNorm + DirectDiffuseSplit + StaticDelayCalc "
                                          "on the device), gain-norm table of default.tf shape with random values"}
    return extras


_REAL_STDOUT = None


def emit(line):
    """The one JSON line, on the real stdout (fd 1 is pointed at stderr while the run is going so that
    library banners such as NCCL's version line cannot land in front of it)."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, data)
    else:
        sys.stdout.write(data.decode())
        sys.stdout.flush()


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if args.gpus > 1 and world == 1:
            raise SystemExit("launch with torchrun for --gpus > 1 (see the header of this file)")
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
