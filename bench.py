#!/usr/bin/env python
"""Benchmark of the BEAR render hot path (BASELINE.json metric: object-channels rendered in real
time at 48 kHz, N x RTF).

A step = one pass of the hot path over T periods of synthetic input of BASELINE.json configs[1]
("64 moving Objects with per-block metadata changes, block size 512"; the listener yaw steps 2
degrees per block so every block crossfades to a new BRIR view) per GPU.  With --gpus N each rank
renders its own group of 64 objects of one 64*N-object scene and the per-step stereo partial mixes
are summed with an NCCL all-reduce (weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

ours:      value = device-resident throughput (audio and tables already in HBM), CUDA events on the
           renderer's stream; e2e = the same metric through the host-buffer C-ABI call
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
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FS = 48000
METRIC = "object-channels rendered real-time at 48 kHz per B200 (N x RTF)"
UNIT = "object-channels x RTF"
WORKLOADS = {
    # name: (scenario, period, objects per GPU)
    "cfg2_ht": ("cfg2_ht", 512, 64),
    "cfg4_ht": ("cfg4_ht", 256, 1024),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2_ht", choices=sorted(WORKLOADS))
    ap.add_argument("--seconds", type=float, default=10.0, help="audio seconds per step and GPU")
    ap.add_argument("--cpu-sample-blocks", type=int, default=0, help="blocks of the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--chunk-blocks", type=int, default=0, help="periods per launch group (0 = library default)")
    ap.add_argument("--cmac-splits", type=int, default=0, help="partial spectra per period (0 = library default)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------- workload
def make_workload(workload, T, group=0, num_views=180, seed=0):
    """Synthetic default-shaped data file + tables + audio for T periods of one object group."""
    from bear_b200 import control, data_file
    scen, B, _ = WORKLOADS[workload]
    data = data_file.from_tree(data_file.synthetic_tree(seed=seed, num_views=num_views))
    ctl = control.HostControl(data)
    N, M, H, tb = control.scenario(ctl, scen, T, B, seed=1000 + group)
    if num_views != 180:  # reduced view table for CPU samples: cycle through the views that exist
        tb["brir_index"] = (np.arange(T) % num_views).astype(np.int32)
        tb["static_delays"] = ctl.static_delays(tb["brir_index"]).astype(np.float32)
    scale = 0.1 if N <= 64 else 0.02
    audio = control.synth_audio(N, T * B, seed_base=10000 * group, scale=scale)
    return data, ctl, N, B, tb, audio


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
    data, ctl, N, B, tb, audio = make_workload(workload, T, num_views=num_views)
    _W.update(r=bear_ref.RefRenderer(data, N, 0, 0, B), blocks=tb.blocks(), audio=audio, N=N, B=B, T=T)


def _cpu_worker_step(_):
    t0 = time.perf_counter()
    _W["r"].render(_W["audio"], None, None, _W["blocks"])
    return time.perf_counter() - t0


def cpu_sample_blocks(args):
    if args.cpu_sample_blocks:
        return args.cpu_sample_blocks
    return 48 if args.workload == "cfg2_ht" else 24


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
            "sample": "%d x %d periods of %d (%.2f s of audio) of the same workload, oracle/bear_ref f32, 1 thread"
                      % (reps, T, B, audio_s)}


def run_reference(args, rank):
    """--impl reference: the CPU path on all host cores (one renderer instance per core, each renders
    the bounded sample; a bear::Renderer is single-threaded, visr_bear/src/dynamic_renderer.hpp:14-15)."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    T = cpu_sample_blocks(args)
    scen, B, N = WORKLOADS[args.workload]
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
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, T, cores),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d instances x %d periods of %d per step, oracle/bear_ref f32 (the reference "
                                   "renderer needs VISR+Boost: unbuildable here)" % (cores, T, B)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(args, T, instances=None):
    scen, B, N = WORKLOADS[args.workload]
    c = {"workload": "BASELINE configs[1]: 64 moving Objects, per-block metadata, block 512, listener yaw 2 deg/block "
                     "(view crossfade every block)" if args.workload == "cfg2_ht" else
                     "BASELINE configs[3] shape: 1024 moving Objects, block 256, view crossfade every block",
         "objects_per_gpu": N, "period": B, "periods_per_step": T, "audio_seconds_per_step": T * B / FS,
         "data_file": "synthetic default shape (180 views x 24 x 2 x 2976 taps, 128-tap decorrelators)",
         "sharding": "object groups, NCCL all-reduce of stereo partial mixes" if args.gpus > 1 else "single GPU",
         "l2": "inputs larger than L2 (audio + spectra streamed per step >> 126 MB)"}
    if instances:
        c["cpu_instances"] = instances
    return c


# ------------------------------------------------------------------------------------------- GPU leg
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from bear_b200 import device_io
    from bear_b200.renderer import Renderer, pinned_empty

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    scen, B, N = WORKLOADS[args.workload]
    T = max(8, int(args.seconds * FS / B) // 8 * 8)
    data, ctl, N, B, tb, audio = make_workload(args.workload, T, group=rank)
    r = Renderer(data, N, 0, 0, B, max_batch_blocks=T, device=local_rank, chunk_blocks=args.chunk_blocks,
                 cmac_splits=args.cmac_splits)
    io = device_io.DeviceBuffers(r, device_index=local_rank)
    io.load_audio(audio)
    io.load_tables(tb)
    r.synchronize()
    stream = io.stream
    out_dev = io.output(T)

    def barrier():
        r.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step_resident():
        r.process_device(T)
        if world > 1:
            with torch.cuda.stream(stream):
                dist.all_reduce(out_dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
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
    for _ in range(400):      # a fixed count: every rank must issue the same number of collectives
        step_resident()
    r.synchronize()
    for _ in range(max(3, args.warmup)):
        step_resident()
    r.set_profiling(True)
    r.get_profile(reset=True)
    launches0 = r.get_info().launches_total
    ms = timed(step_resident, args.steps)
    launches = r.get_info().launches_total - launches0
    prof = r.get_profile(reset=True)
    r.set_profiling(False)
    info_resident = r.get_info()     # incl. the work items / splits the device chose for this launch

    # ---- end to end: pinned host buffers through the host-buffer entry point -------------------
    h_audio = pinned_empty(audio.shape); h_audio[:] = audio
    h_tb = type(tb)({k: pinned_empty(v.shape, v.dtype) for k, v in tb.items()})
    for k, v in tb.items():
        h_tb[k][...] = v
    h_out = pinned_empty((2, T * B))
    h2d = audio.nbytes + sum(v.nbytes for v in tb.values())
    d2h = h_out.nbytes

    def step_e2e():
        # host-buffer entry point (chunked uploads overlap the kernels); with several GPUs the partial
        # mix it leaves in the device output area is then summed over the ranks and fetched again
        r.process_batch(h_audio, None, None, h_tb, output=h_out)
        if world > 1:
            with torch.cuda.stream(stream):
                dist.all_reduce(out_dev)
                torch.from_numpy(h_out).copy_(out_dev, non_blocking=True)
            stream.synchronize()

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, e2e_steps)
    clk = clocks.stop() if rank == 0 else None
    # per-launch kernel times at the chunk size of the host-buffer path (untimed extra call)
    r.set_profiling(True)
    r.get_profile(reset=True)
    step_e2e()
    e2e_prof = {k: v[0] / max(v[1], 1) for k, v in r.get_profile(reset=True).items() if v[1]}
    r.set_profiling(False)

    # the PCIe floor of that call: the same bytes copied host->device from pinned memory, nothing else
    h_flat = torch.from_numpy(h_audio.reshape(-1))
    d_flat = torch.empty(h2d // 4, dtype=torch.float32, device="cuda")

    def step_copy():
        with torch.cuda.stream(stream):
            d_flat[:h_flat.numel()].copy_(h_flat, non_blocking=True)
            d_flat[h_flat.numel():].copy_(h_flat[:d_flat.numel() - h_flat.numel()], non_blocking=True)

    step_copy()
    ms_copy = timed(step_copy, e2e_steps) / e2e_steps

    # ---- the same call fed with the file's interleaved frames (float and 16-bit PCM): what a file
    # renderer holds after reading a block of a BW64 file; de-interleaving / PCM decoding on the device
    frames_res = {}
    if world == 1:
        tracks = np.arange(N, dtype=np.int32)
        for fmt, dt, scale in (("f32", np.float32, 1.0), ("s16", np.int16, 32768.0)):
            h_frames = pinned_empty((T * B, N), dt)
            h_frames[...] = np.clip(np.round(audio.T * scale), -scale, scale - 1) if fmt != "f32" else audio.T

            def step_frames():
                r.process_batch_frames(h_frames, tracks, h_tb, output=h_out)

            step_frames()
            ms_f = timed(step_frames, e2e_steps) / e2e_steps
            frames_res[fmt] = {"ms_per_step": ms_f, "h2d_bytes_per_step": int(h_frames.nbytes + h2d - audio.nbytes),
                               "value": N * audio_s_of(T, B) / (ms_f * 1e-3)}

    # ---- real-time use: one period per call through bear_cuda_process (host pointers), wall clock ----
    stream_ms = None
    stream_kernels = None
    if rank == 0:
        with Renderer(data, N, 0, 0, B, max_batch_blocks=1, device=local_rank) as rs:
            blk = [tb.block(k) for k in range(8)]
            xin = [np.ascontiguousarray(audio[:, k * B:(k + 1) * B]) for k in range(8)]
            out1 = np.empty((2, B), np.float32)
            for k in range(16):
                rs.process(xin[k % 8], None, None, blk[k % 8], output=out1)
            t0 = time.perf_counter()
            for k in range(200):
                rs.process(xin[k % 8], None, None, blk[k % 8], output=out1)
            stream_ms = (time.perf_counter() - t0) / 200 * 1e3
            rs.set_profiling(True)
            for k in range(50):
                rs.process(xin[k % 8], None, None, blk[k % 8], output=out1)
            stream_kernels = {kk: v[0] / max(v[1], 1) for kk, v in rs.get_profile().items() if v[1]}

    # ---- the "1000 objects in real time" target: 1024 moving objects, period 256 (BASELINE configs[3]
    # shape on ONE GPU), 2 s of audio per launch, device-resident, same timing method ----
    many = None
    if rank == 0 and world == 1 and args.workload == "cfg2_ht":
        from bear_b200 import control as _control
        B4, T4 = 256, 376
        N4, _, _, tb4 = _control.scenario(ctl, "cfg4_ht", T4, B4, seed=77)
        audio4 = _control.synth_audio(N4, T4 * B4, seed_base=500000, scale=0.02)
        with Renderer(data, N4, 0, 0, B4, max_batch_blocks=T4, device=local_rank) as r4:
            io4 = device_io.DeviceBuffers(r4, device_index=local_rank)
            io4.load_audio(audio4)
            io4.load_tables(tb4)
            r4.synchronize()
            for _ in range(5):
                r4.process_device(T4)
            r4.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(io4.stream)
            for _ in range(10):
                r4.process_device(T4)
            e1.record(io4.stream)
            e1.synchronize()
            ms4 = e0.elapsed_time(e1) / 10
        rtf4 = (T4 * B4 / FS) / (ms4 * 1e-3)
        many = {"objects": N4, "period": B4, "periods_per_launch": T4, "ms_per_launch": ms4, "rtf": rtf4,
                "object_channels_x_rtf": N4 * rtf4,
                "note": "1024 moving objects with per-period metadata and a view change every period on one GPU"}
        del audio4, tb4

    # ---- control tail on the device (delays + gain normalisation + split), kernel time only ----
    control_ms = None
    if rank == 0:
        import copy
        d2 = copy.copy(data)
        S = 40   # default.tf: max loudspeaker delay 38 + 2 (finalise.py:214-215)
        d2.gain_norm_factors = np.random.default_rng(0).random((S, data.num_views, 2 * ctl.L, 2 * ctl.L, 2),
                                                               dtype=np.float32)
        Tc = min(T, 128)
        with Renderer(d2, N, 0, 0, B, max_batch_blocks=Tc, device=local_rank) as rc:
            raw = np.concatenate([tb["direct_gains"][:Tc], tb["diffuse_gains"][:Tc]], axis=1)
            rc.set_profiling(True)
            for _ in range(4):
                rc.control_tables(raw, None, tb["brir_index"][:Tc])
            t_ctl, n_ctl = rc.get_profile()["control_kernel"]
            control_ms = {"ms_per_launch": t_ctl / max(n_ctl, 1), "periods_per_launch": Tc, "objects": N,
                          "ms_per_10s_audio": t_ctl / max(n_ctl, 1) * T / Tc,
                          "note": "control_kernel (DirectDelayCalc + GainNorm + DirectDiffuseSplit + StaticDelayCalc on the "
                                  "device), gain-norm table of default.tf shape with random values"}

    audio_s = T * B / FS
    value = world * N * audio_s * args.steps / (ms * 1e-3)
    e2e_value = world * N * audio_s * e2e_steps / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        peak, which = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks else \
            (6650.0, "fallback (B200_PROFILING.md)")
        info = info_resident
        L, A, P = data.num_virtual_loudspeakers, data.n_hoa_channels, info.num_partitions
        KS = info.cmac_splits_last or info.cmac_splits
        views = np.asarray(tb["brir_index"]).astype(np.int64)
        n_views = int(np.unique(views).size)
        S = info.object_splits
        fB = 4 * T * B                                   # bytes of one float row of the launch
        # Algorithmic (minimum-traffic) bytes per launch of every kernel, DESIGN.md section 4:
        #   mix          audio in + gain/delay tables in, direct (2L rows) and diffuse (L rows) bus out
        #   decorrelate  diffuse bus in, decorrelated bus out
        #   bus_fdl      direct bus + decorrelated bus in, 2L packed spectra (B x 8 bytes a period) out
        #   cmac         filter partitions of every DISTINCT view of the launch once (the kernel groups
        #                the periods by view), every delay-line spectrum once, partial spectra out
        #   ifft_out     partial spectra in (old-view ones too in a view-change period), 2 rows out
        xf = int(np.count_nonzero(np.diff(np.concatenate([[views[0]], views])) != 0))
        alg_by_kernel = {
            "mix_kernel": N * fB + 4 * T * (2 * 2 * L * N + 2 * 2 * N) + S * 3 * L * fB,
            "decorrelate_kernel": S * L * fB + L * fB,
            "bus_fdl_kernel": S * 2 * L * fB + L * fB + 2 * L * 2 * fB,
            "cmac_kernel": 8 * B * (2 * L * P * n_views + 2 * L * T + 2 * 2 * KS * T),
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
        # cmac_kernel under the two older accountings, for comparison across rounds: one filter set per
        # PERIOD (what a kernel without view grouping must read), and SURVEY 8(d)'s single-period
        # streaming figure (delay-line partitions re-read by each of the P periods that use them,
        # filters of old and new view both read in a view-change period)
        alg_per_period = T * 8 * B * (2 * L * P + 2 * L + 2 * 2 * KS)
        alg_survey = T * 2 * 8 * B * (L * P + L * P * 2.0 + KS * 2.0)
        t_cmac, n_cmac = prof["cmac_kernel"]
        cmac_ms = t_cmac / max(n_cmac, 1)
        traffic_all = {}
        tpath = ROOT / "profiles" / "kernel_traffic.json"
        if tpath.exists():
            try:
                traffic_all = json.loads(tpath.read_text()).get(args.workload, {})
            except Exception:
                traffic_all = {}
        traffic = traffic_all.get(dominant)
        dom = by_kernel[dominant]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, T),
            "rtf_per_gpu": audio_s * args.steps / (ms * 1e-3),
            "roofline": {"bound": "hbm", "kernel": dominant, "achieved": dom["achieved"], "peak": peak, "unit": "GB/s",
                         "frac": dom["frac"], "traffic": traffic, "peak_source": which,
                         "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_launch"],
                         "avg_launch_ms": dom["avg_launch_ms"], "share_of_step": dom["share_of_step"],
                         "traffic_source": "profiles/kernel_traffic.json (ncu --set full of the same command, "
                                           "dram read+write per launch)",
                         "note": "dominant kernel by time; no tensor-core work on this path, so every kernel is "
                                 "held against HBM.  A fraction well below 1 with traffic ~ algorithmic bytes "
                                 "means the kernel is bound by FP32 issue / latency, not by re-reads "
                                 "(DESIGN.md section 4)"},
            "roofline_by_kernel": {k: dict(v, traffic=traffic_all.get(k)) for k, v in by_kernel.items()},
            "cmac_accounting": {
                "distinct_views_in_launch": n_views, "work_items": int(info.cmac_items_last),
                "splits": int(KS), "avg_launch_ms": cmac_ms,
                "per_period_filter_bytes_per_launch": alg_per_period,
                "achieved_per_period_accounting": alg_per_period / (cmac_ms * 1e-3) / 1e9 if cmac_ms > 0 else 0.0,
                "survey_8d_streaming_bytes_per_launch": alg_survey,
                "achieved_survey_8d": alg_survey / (cmac_ms * 1e-3) / 1e9 if cmac_ms > 0 else 0.0,
                "note": "GB/s-equivalents above the HBM peak are possible because the kernel reads each view's "
                        "filters once per launch (grouping) and the delay line mostly from L2"},
            "kernel_ms_per_step": {k: v[0] / max(args.steps, 1) for k, v in prof.items() if v[1]},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / e2e_steps, "steps": e2e_steps,
                    "kernel_ms_per_chunk_launch": e2e_prof,
                    "h2d_copy_only_ms": ms_copy, "h2d_copy_only_gbs": h2d / ms_copy * 1e-6,
                    "api": "bear_cuda_process_batch (pinned host buffers)" if world == 1 else
                           "bear_cuda_process_batch (pinned host buffers) -> NCCL all-reduce of the partial mixes -> pinned host"},
            "e2e_file_frames": dict(frames_res, note="bear_cuda_process_batch_frames: interleaved frames as read from "
                                                     "the file (pinned), same tables; not the headline e2e (that is the "
                                                     "reference-facing float call above)") if frames_res else None,
            "gpu_launches": int(launches),
            "many_objects": many,
            "control_tail": control_ms,
            "stream": {"ms_per_period_call": stream_ms, "period_ms": 1e3 * B / FS,
                       "kernel_ms_per_period": stream_kernels if rank == 0 else None,
                       "note": "bear_cuda_process, one period per call incl. host<->device copies and sync (real-time use)"},
            "clocks": clk,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_single(args)
        emit(line)
    r.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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


def audio_s_of(T, B):
    return T * B / FS


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
