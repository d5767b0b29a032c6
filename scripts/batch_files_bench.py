#!/usr/bin/env python
"""BASELINE configs[4]: offline batch render of F synthetic mixed ADM files (16 Objects + 12
DirectSpeakers + 16 HOA channels, 10 s each), sharded BY FILE across the GPUs of one box.

One process per GPU (torchrun for N > 1); each rank renders files rank, rank+N, ... on one renderer
instance with bear_cuda_reset between files, through the host-buffer entry point (pinned host ->
device -> kernels -> host).  No collective is needed: files are independent.  Prints one JSON line.

    python scripts/batch_files_bench.py [--files 256] [--seconds 10]
    python -m torch.distributed.run --nproc-per-node 8 ... scripts/batch_files_bench.py --gpus 8
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
FS, B = 48000, 512


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--files", type=int, default=256)
    ap.add_argument("--seconds", type=float, default=10.0)
    ap.add_argument("--distinct", type=int, default=4, help="distinct synthetic files cycled through")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    from bear_b200 import control, data_file, parallel
    from bear_b200.renderer import Renderer, pinned_empty

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    T = max(8, int(args.seconds * FS / B) // 8 * 8)
    data = data_file.from_tree(data_file.synthetic_tree(seed=0))
    ctl = control.HostControl(data)
    files = []
    for f in range(args.distinct):
        N, M, H, tb = control.scenario(ctl, "cfg5_ht", T, B, seed=100 + f)
        audio = [control.synth_audio(c, T * B, 1000 * f + s, 0.05) for c, s in ((N, 0), (M, 100), (H, 200))]
        pinned = []
        for a in audio:
            p = pinned_empty(a.shape)
            p[:] = a
            pinned.append(p)
        ptb = type(tb)({k: pinned_empty(v.shape, v.dtype) for k, v in tb.items()})
        for k, v in tb.items():
            ptb[k][...] = v
        files.append((*pinned, ptb))
    out = pinned_empty((2, T * B))
    mine = parallel.file_shard(rank, world, args.files)
    with Renderer(data, N, M, H, B, max_batch_blocks=T, device=local_rank) as r:
        for i in range(2):                                  # warm-up
            r.reset()
            r.process_batch(*files[i % len(files)][:3], files[i % len(files)][3], output=out)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for idx in mine:
            obj, ds, hoa, tb = files[idx % len(files)]
            r.reset()
            r.process_batch(obj, ds, hoa, tb, output=out)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device="cuda")
        launches = r.get_info().launches_total
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    if rank == 0:
        secs = float(dt.item())
        audio_s = T * B / FS
        line = {"workload": "BASELINE configs[4]: batch of mixed files (16 Objects + 12 DirectSpeakers + 16 HOA), "
                            "sharded by file, no collective",
                "files": args.files, "n_gpus": world, "seconds_per_file": audio_s, "wall_s": secs,
                "files_per_s": args.files / secs, "rtf_per_file_stream": args.files * audio_s / secs,
                "channels_x_rtf": (N + M + H) * args.files * audio_s / secs,
                "h2d_bytes_per_file": int(sum(a.nbytes for a in files[0][:3]) + sum(v.nbytes for v in files[0][3].values())),
                "scaling": "weak" if False else "strong", "api": "bear_cuda_reset + bear_cuda_process_batch (pinned host buffers)",
                "gpu_launches_rank0": int(launches)}
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
