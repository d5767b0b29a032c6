"""One scene on several GPUs, run under torch.distributed.run by tests/test_multi_gpu.py:

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mp_shared_scene.py

Every rank renders its object group (DirectSpeakers on the last rank, HOA on rank 0) through
bear_cuda_comm_connect; rank 0 compares the result with the same scene rendered by one unconnected
renderer and with the CPU oracle (the reference property: rendering items together equals the sum of
rendering them apart, visr_bear/test/test_changes.py:341-361, atol 1e-5)."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    import torch
    import torch.distributed as dist
    from bear_b200 import control, data_file, device_io, parallel
    from bear_b200.renderer import Renderer
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    data = data_file.from_tree(data_file.synthetic_tree(seed=1, num_views=12, brir_length=700))
    ctl = control.HostControl(data)
    B, T, N, M, H = 256, 20, 70, 2, 4
    rng = np.random.default_rng(3)
    tb = ctl.tables(2 * T, obj_az=rng.uniform(-180, 180, size=(2 * T, N)), obj_diffuse=np.where(np.arange(N) % 3 == 0, .4, 0.),
                    obj_gain=0.3, ds_speakers=[2, 5], hoa_gains=np.eye(16, 4, dtype=np.float32), H=H,
                    yaw_deg=30.0 * np.arange(2 * T))
    obj, ds, hoa = (control.synth_audio(c, 2 * T * B, s) for c, s in ((N, 0), (M, 100), (H, 200)))

    sl = parallel.object_group(rank, world, N)
    n, m, h = sl.stop - sl.start, (M if rank == world - 1 else 0), (H if rank == 0 else 0)
    mine = control.BlockTables(tb)
    mine["direct_gains"] = np.ascontiguousarray(tb["direct_gains"][:, :, sl])
    mine["diffuse_gains"] = np.ascontiguousarray(tb["diffuse_gains"][:, :, sl])
    mine["direct_delays"] = np.ascontiguousarray(tb["direct_delays"][:, 2 * sl.start:2 * sl.stop])
    if not m:
        mine["direct_speakers_gains"] = np.zeros((2 * T, ctl.L, 0), np.float32)
        mine["direct_speakers_delays"] = np.zeros((2 * T, 0), np.float32)
    if not h:
        mine["hoa_gains"] = np.zeros((2 * T, ctl.A, 0), np.float32)

    failures = []

    def check(label, got, want_args, receives):
        if receives:
            Nw, Mw, Hw, hoa_w, tb_w = want_args
            with Renderer(data, Nw, Mw, Hw, B, max_batch_blocks=T, device=local) as whole:
                want = whole.render(obj, ds, hoa_w, tb_w)
            err = float(np.abs(got - want).max())
            print("rank %d %s: max-abs vs one GPU %.3e, peak %.3f" % (rank, label, err, np.abs(want).max()), flush=True)
            if not (err <= 1e-5 and np.abs(want).max() > 0.05):
                failures.append((label, err))
            if rank == 0:
                from oracle import bear_ref
                ref = bear_ref.RefRenderer(data, Nw, Mw, Hw, B).render(obj, ds, hoa_w, tb_w.blocks())
                err = float(np.abs(got - ref).max())
                print("rank 0 %s: max-abs vs oracle %.3e" % (label, err), flush=True)
                if err > 1e-5:
                    failures.append((label, "oracle", err))

    for root, chunk in ((0, 0), (-1, 7)):
        with Renderer(data, n, m, h, B, max_batch_blocks=T, device=local, chunk_blocks=chunk) as r:
            parallel.connect_renderers(r, rank, world, root=root)
            info = r.get_info()
            assert info.world == world and info.rank == rank
            receives = root < 0 or root == rank
            # two calls: host-buffer path, then the device-resident path (the bus rows alternate)
            first = r.process_batch(obj[sl, :T * B], ds[:, :T * B] if m else None, hoa[:, :T * B] if h else None,
                                    mine.slice(0, T), output=None if receives else False)
            io = device_io.DeviceBuffers(r, device_index=local)
            rows = [obj[sl, T * B:]] + ([ds[:, T * B:]] if m else []) + ([hoa[:, T * B:]] if h else [])
            io.load_audio(np.concatenate(rows))
            io.load_tables(mine.slice(T, 2 * T))
            r.process_device(T)
            r.synchronize()
            second = io.read_output(T) if receives else None
        dist.barrier()
        check("root %d chunk %d" % (root, chunk), np.concatenate([first, second], axis=1) if receives else None,
              (N, M, H, hoa, tb), receives)
        dist.barrier()

    # Pipelined calls (bear_cuda_process_device_pipelined): the convolution side of call s is still queued
    # when the inputs of call s+1 are written and its mix starts.  Five calls of 8 periods; the scene has no
    # HOA (a rank with HOA inputs runs these calls unpipelined, which the second pass covers: HOA on rank 0).
    for with_hoa in (False, True):
        hh = h if with_hoa else 0
        tb_p, mine_p = control.BlockTables(tb), control.BlockTables(mine)
        if not with_hoa:
            tb_p["hoa_gains"] = np.zeros((2 * T, ctl.A, 0), np.float32)
        if not hh:
            mine_p["hoa_gains"] = np.zeros((2 * T, ctl.A, 0), np.float32)
        Tc = 8
        with Renderer(data, n, m, hh, B, max_batch_blocks=Tc, device=local) as r:
            parallel.connect_renderers(r, rank, world, root=0)
            io = device_io.DeviceBuffers(r, device_index=local)

            def load(c):
                a, b = c * Tc * B, (c + 1) * Tc * B
                rows = [obj[sl, a:b]] + ([ds[:, a:b]] if m else []) + ([hoa[:, a:b]] if hh else [])
                io.load_audio(np.concatenate(rows))
                io.load_tables(mine_p.slice(c * Tc, (c + 1) * Tc))

            outs = []
            ncalls = 2 * T // Tc
            load(0)
            r.process_device_pipelined(Tc)
            for c in range(1, ncalls):
                load(c)   # queued on the render stream while call c-1's convolution side is still running
                outs.append(io.read_output(Tc, pipelined=True))
                r.process_device_pipelined(Tc)
            outs.append(io.read_output(Tc, pipelined=True))
            r.synchronize()
        dist.barrier()
        check("pipelined hoa=%d" % with_hoa, np.concatenate(outs, axis=1) if rank == 0 else None,
              (N, M, H if with_hoa else 0, hoa if with_hoa else None, tb_p), rank == 0)
        dist.barrier()
    flag = torch.tensor([len(failures)])
    dist.all_reduce(flag)
    dist.destroy_process_group()
    if int(flag.item()):
        raise SystemExit("shared-scene render differs: %s" % failures)
    if rank == 0:
        print("SHARED_SCENE_OK world=%d" % world, flush=True)


if __name__ == "__main__":
    main()
