"""Multi-process host logic (world_size 2, gloo, CPU): object-group sharding + stereo partial-mix
all-reduce reproduces the single-process render.  The local renderer is the CPU oracle here (tests
may use it); on the GPU box the same ShardedScene wraps bear_b200.renderer.Renderer."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from bear_b200 import control, data_file, parallel


def test_object_groups_partition():
    for N in (0, 1, 7, 64, 1024):
        for world in (1, 2, 3, 8):
            sls = [parallel.object_group(r, world, N) for r in range(world)]
            assert sls[0].start == 0 and sls[-1].stop == N
            assert all(a.stop == b.start for a, b in zip(sls, sls[1:]))
            sizes = [s.stop - s.start for s in sls]
            assert max(sizes) - min(sizes) <= 1
    assert sorted(sum((parallel.file_shard(r, 4, 10) for r in range(4)), [])) == list(range(10))


def test_loudspeaker_ranges_of_a_shared_scene_partition():
    """bear_cuda_comm_speaker_range (host only): the ranks' convolution shares tile [0, L)."""
    import ctypes as C
    from bear_b200 import renderer
    lib = renderer.lib()
    for L in (1, 3, 24):
        for world in (1, 2, 4, 5, 8):
            r = []
            for rank in range(world):
                b, e = C.c_uint32(), C.c_uint32()
                lib.bear_cuda_comm_speaker_range(rank, world, L, C.byref(b), C.byref(e))
                r.append((b.value, e.value))
            assert r[0][0] == 0 and r[-1][1] == L
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1


class _OracleRenderer:
    def __init__(self, data, n, m, h, B):
        from oracle import bear_ref
        self.r = bear_ref.RefRenderer(data, n, m, h, B)

    def render(self, obj, ds, hoa, tb):
        return self.r.render(obj, ds, hoa, tb.blocks()).astype(np.float32)


def _scene():
    data = data_file.from_tree(data_file.synthetic_tree(seed=1, num_views=6, brir_length=600))
    ctl = control.HostControl(data)
    B, T, N = 256, 5, 6
    rng = np.random.default_rng(0)
    tb = ctl.tables(T, obj_az=rng.uniform(-180, 180, size=(T, N)), obj_diffuse=[0, .5, 0, 1, 0, .3],
                    ds_speakers=[2, 5], hoa_gains=np.eye(16, 4, dtype=np.float32), H=4, yaw_deg=60.0 * np.arange(T))
    audio = [control.synth_audio(c, T * B, s) for c, s in ((N, 0), (2, 100), (4, 200))]
    return data, B, N, tb, audio


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    data, B, N, tb, (obj, ds, hoa) = _scene()
    scene = parallel.ShardedScene(lambda n, m, h: _OracleRenderer(data, n, m, h, B), N, 2, 4, rank, world)
    out = scene.render(obj, ds, hoa, tb)
    q.put((rank, scene.counts, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_scene_matches_single_process_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=240) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert results[0][1] == (3, 2, 4) and results[1][1] == (3, 0, 0)
    data, B, N, tb, (obj, ds, hoa) = _scene()
    whole = _OracleRenderer(data, N, 2, 4, B).render(obj, ds, hoa, tb)
    for _, _, out in results:   # all-reduce: every rank holds the full mix
        np.testing.assert_allclose(out, whole, atol=1e-5)
    np.testing.assert_array_equal(results[0][2], results[1][2])
